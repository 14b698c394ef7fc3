/* sanm_b200/csrc/capi_util.h -- shared pieces of the C ABI translation units. */
#pragma once

#include "../../include/sanm_b200.h"
#include "graph.h"
#include "remap.h"
#include "taylor.h"

#include <memory>

namespace sanm_b200 {
class SparseLU;
class AnmSolver;
namespace capi {
std::string& last_error();

template <typename Fn>
int guard(Fn&& fn) {
    try {
        fn();
        return SANM_OK;
    } catch (Error& e) {
        last_error() = e.what();
        return e.code;
    } catch (std::exception& e) {
        last_error() = e.what();
        return SANM_ERR_GENERIC;
    }
}
}  // namespace capi
}  // namespace sanm_b200

struct sanm_graph {
    std::shared_ptr<sanm_b200::Graph> g;
};
struct sanm_prop {
    std::unique_ptr<sanm_b200::TaylorProp> p;
};
struct sanm_remap {
    std::shared_ptr<sanm_b200::Remap> r;
};
