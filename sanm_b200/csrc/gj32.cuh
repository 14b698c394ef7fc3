/* sanm_b200/csrc/gj32.cuh -- Gauss-Jordan inversion with partial pivoting of
 * the 32 x 32 pivot block of a front by one CTA of 4 warps (CUDA only; the
 * host-emulation build uses the generic gj_steps of sparse_lu.cu).
 *
 * This is the sequential heart of the blocked factorisation (one call per 32
 * pivots of every large front), so it is written for latency:
 *  - rows never move: every lane tracks the position `pos` of row `lane` in
 *    the permuted matrix (identically in all warps);
 *  - warp w owns columns 8w..8w+7 of the update; column k+1 is additionally
 *    published through a ping-pong buffer, so one __syncthreads per pivot is
 *    the only block-wide synchronisation;
 *  - the arg-max uses two 32-bit warp reductions on the bit pattern of |a|
 *    (non-negative doubles order like unsigned integers), the reciprocal of
 *    every candidate is computed while the arg-max is in flight;
 *  - the 16 shared-memory loads of a pivot are issued before the arithmetic
 *    (the compiler otherwise serialises load -> DFMA -> store per column).
 */
#pragma once
#include <cstdint>

namespace sanm_b200 {

//! barrier of the 128 threads that run block_gj32 (a named barrier, so that
//! the routine can also be called by the first four warps of a larger CTA)
__device__ __forceinline__ void gj32_sync() {
    asm volatile("bar.sync 1, 128;" ::: "memory");
}

/*! D: shared, column-major, ld 32, rows/cols >= nb zero.  On return D holds
 * inv(P A) and lperm[i] = original row at position i.  acol: shared [64].
 * Must be called by threads 0..127 of the CTA, all of them. */
__device__ inline void block_gj32(double* __restrict__ D, uint32_t nb,
                                  uint32_t* __restrict__ lperm,
                                  double* __restrict__ acol, double thresh,
                                  int* nr_perturbed) {
    constexpr int NBK = 32, CW = 8;  // block size, columns per warp
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t j0 = w * CW;
    uint32_t pos = lane;
    if (w == 0) acol[lane] = D[lane];  // column 0
    gj32_sync();
    for (uint32_t k = 0; k < nb; ++k) {
        const bool eligible = pos >= k && pos < nb;
        const double a = acol[(k & 1) * NBK + lane];
        // speculative, off the critical path (1 for rows that cannot be chosen:
        // a zero would send the division down its slow path)
        const double inv_a = 1.0 / ((eligible && a != 0.0) ? a : 1.0);
        const unsigned long long bits =
                eligible ? (unsigned long long)__double_as_longlong(fabs(a)) : 0ull;
        const uint32_t hi = (uint32_t)(bits >> 32), lo = (uint32_t)bits;
        const uint32_t mhi = __reduce_max_sync(0xffffffffu, hi);
        const uint32_t mlo =
                __reduce_max_sync(0xffffffffu, (eligible && hi == mhi) ? lo : 0u);
        const uint32_t cand =
                __ballot_sync(0xffffffffu, eligible && hi == mhi && lo == mlo);
        const uint32_t who = __ffs(cand) - 1;  // ties -> lowest lane
        const double best = __longlong_as_double(
                (long long)(((unsigned long long)mhi << 32) | mlo));
        const uint32_t ppos = __shfl_sync(0xffffffffu, pos, who);
        if (pos == k) pos = ppos;
        if (lane == who) pos = k;
        double ip = __shfl_sync(0xffffffffu, inv_a, who);
        if (!(best > thresh)) {  // uniform: static pivot perturbation
            const double piv = __shfl_sync(0xffffffffu, a, who);
            ip = 1.0 / (piv < 0 ? -thresh : thresh);
            if (threadIdx.x == 0) atomicAdd(nr_perturbed, 1);
        }
        double v[CW], pr[CW];
#pragma unroll
        for (int j = 0; j < CW; ++j) v[j] = D[(j0 + j) * NBK + lane];
#pragma unroll
        for (int j = 0; j < CW; ++j) pr[j] = D[(j0 + j) * NBK + who];
        __syncwarp();
        asm volatile("" ::: "memory");  // all 16 loads in flight before the math
        const bool is_piv = lane == who;
        const double na = is_piv ? 0.0 : -a;
#pragma unroll
        for (int j = 0; j < CW; ++j) {
            pr[j] *= ip;
            double nv = fma(na, pr[j], v[j]);
            if (is_piv) nv = pr[j];
            if (j0 + j == k) nv = is_piv ? ip : na * ip;
            D[(j0 + j) * NBK + lane] = nv;
            if (j0 + j == k + 1) acol[((k + 1) & 1) * NBK + lane] = nv;
        }
        gj32_sync();
    }
    // move the rows to their positions
    double r[CW];
#pragma unroll
    for (int j = 0; j < CW; ++j) r[j] = D[(j0 + j) * NBK + lane];
    gj32_sync();
    if (lane < nb) {
#pragma unroll
        for (int j = 0; j < CW; ++j) D[(j0 + j) * NBK + pos] = r[j];
        if (w == 0) lperm[pos] = lane;
    }
    gj32_sync();
}

}  // namespace sanm_b200
