/* sanm_b200/csrc/comm.h -- the one exchange step of the multi-GPU path.
 *
 * The reference parallelises over contiguous batch (tet) ranges inside one
 * process and gathers by memcpy (libsanm/symbolic.cpp:418-439, 525-529).  Here
 * each range lives on its own GPU (one process per GPU) and every rank keeps
 * the full nodal vectors; the only data-path exchange is the sum over ranks of
 * nodal quantities assembled from the local tets: f(x0), the right-hand side
 * b_k of every order, and once per ANM iteration the CSR values of K (and
 * grad_t).  All of them go through allreduce_sum(): ncclAllReduce on the
 * library stream over NVLink / NVSwitch, in place, no host synchronisation.
 * After the reduction every rank holds bitwise identical data and runs the
 * (replicated) factorisation, back-solves and scalar decisions itself, so no
 * broadcast of x_k is needed.
 *
 * A host callback transport exists for the CPU-only tests (gloo, world 2).
 */
#pragma once

#include "rt.h"

namespace sanm_b200 {
namespace comm {

int world();
int rank();
//! in-place sum over ranks of a device vector; no-op for a single rank
void allreduce_sum(double* dev, size_t n);
//! contiguous shard [lo, hi) of `total` items owned by `rank` (the split of
//! libsanm/symbolic.cpp:525-529)
inline void shard_range(uint64_t total, int rank, int world, uint64_t* lo,
                        uint64_t* hi) {
    *lo = total * (uint64_t)rank / (uint64_t)world;
    *hi = total * (uint64_t)(rank + 1) / (uint64_t)world;
}

}  // namespace comm
}  // namespace sanm_b200
