/* sanm_b200/csrc/shard.h -- see shard.cpp. */
#pragma once

#include "graph.h"
#include "remap.h"

namespace sanm_b200 {

//! copy of g whose [N, ...] constants are restricted to items [lo, hi)
std::shared_ptr<Graph> shard_graph(const Graph& g, uint64_t N, uint64_t lo,
                                   uint64_t hi);
//! rows of items [lo, hi) of a map onto an [N, ...] tensor (remap_in)
std::shared_ptr<Remap> shard_remap_rows(const Remap& r, uint64_t N, uint64_t lo,
                                        uint64_t hi);
//! entries reading items [lo, hi) of a map from an [N, ...] tensor (remap_out)
std::shared_ptr<Remap> shard_remap_cols(const Remap& r, uint64_t N, uint64_t lo,
                                        uint64_t hi);

}  // namespace sanm_b200
