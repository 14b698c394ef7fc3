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

#include <vector>

namespace sanm_b200 {
namespace comm {

int world();
int rank();
//! in-place sum over ranks of a device vector; no-op for a single rank
void allreduce_sum(double* dev, size_t n);
/*! Halo exchange for nodal vectors that stay DISTRIBUTED: every rank owns the
 * contiguous index range [lo, hi) that its local tets touch (x-slab sharding
 * of a lexicographically numbered mesh); ranges of neighbouring ranks overlap
 * on the interface nodes only.  halo_sum() adds, inside the overlaps, the
 * contributions of the ranks that share them (ncclSend/ncclRecv between the
 * sharing pairs, summed in rank order so both sides get the same bits);
 * entries outside [lo, hi) are not touched.  Message size = the interface,
 * not n: this is what keeps the element sweep weak-scalable, whereas the
 * all-reduce of full vectors is what the replicated solver needs.  Without
 * NCCL (callback transport of the CPU tests) it falls back to allreduce_sum,
 * which gives the same values inside [lo, hi). */
struct Halo {
    uint64_t lo = 0, hi = 0, n = 0;
    struct Peer {
        int rank;
        uint64_t lo, cnt;  //!< overlap [lo, lo + cnt)
        size_t scratch_off;
    };
    std::vector<Peer> peers;
    rt::DBuf<double> scratch;
    bool ready = false;
};
//! collective: exchange the ranges of all ranks and plan the pairwise overlaps
void halo_setup(Halo& h, uint64_t lo, uint64_t hi, uint64_t n);
//! collective, in place on the library stream
void halo_sum(double* dev, Halo& h);

//! contiguous shard [lo, hi) of `total` items owned by `rank` (the split of
//! libsanm/symbolic.cpp:525-529)
inline void shard_range(uint64_t total, int rank, int world, uint64_t* lo,
                        uint64_t* hi) {
    *lo = total * (uint64_t)rank / (uint64_t)world;
    *hi = total * (uint64_t)(rank + 1) / (uint64_t)world;
}

}  // namespace comm
}  // namespace sanm_b200
