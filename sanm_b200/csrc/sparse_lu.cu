/* sanm_b200/csrc/sparse_lu.cu -- numeric multifrontal factorisation and solves
 * (see sparse_lu.h for the algorithm). */
#include "sparse_lu.h"

#include <cstdlib>
#include "cache.h"
#include "sparse_symbolic.h"
#include "prof.h"
#ifndef SANM_HOSTEMU
#include "gj32.cuh"
#endif

#include <algorithm>
#include <chrono>
#include <cmath>

namespace sanm_b200 {

namespace {

//! per-kernel-type device times of the factor / solve phases (SANM_LU_PROF=1;
//! the event pairs cost ~2 us per launch, so off by default)
const bool g_lu_prof = getenv("SANM_LU_PROF") != nullptr;
struct LuScope {
    std::unique_ptr<prof::Scope> s;
    explicit LuScope(const char* tag) {
        if (g_lu_prof) s.reset(new prof::Scope{tag});
    }
};

constexpr int NB = 32;          //!< pivot block of the large-front path
//! fronts up to this size are factored inside shared memory by one CTA
//! (104^2 doubles = 86 KB: two CTAs per SM; at most 168 fits the 227 KB limit)
uint32_t small_nr() {
    static const uint32_t v = [] {
        const char* e = getenv("SANM_LU_SMALL_NR");
        return e ? (uint32_t)std::min(168, std::max(32, atoi(e))) : 104u;
    }();
    return v;
}
constexpr int ASM_TILE = 4096;  //!< entries per assembly task
constexpr int ABSMAX_BLOCKS = 256;
constexpr int LPERM_PER_FRONT = 128;  //!< >= 2 * NB (narrow path) and >= WB (wide path)
//! fronts of at least this many rows take the superblock (DMMA) path
//! (read at every analyze(): tests force small fronts through it)
uint32_t wide_min_nr() {
    const char* e = getenv("SANM_LU_WIDE_MIN");
    return e ? (uint32_t)std::max(129, atoi(e)) : 1024u;
}
//! levels whose largest front has more rows than this keep the solve's work
//! vector in global memory (L2) instead of shared memory
uint32_t solve_smem_nr() {
    const char* e = getenv("SANM_LU_SOLVE_SMEM_NR");
    return e ? (uint32_t)std::max(64, atoi(e)) : 8192u;
}

struct FrontD {
    double* F;        //!< nr x nr, column-major
    double* vec;      //!< solve workspace [nr]: [w1' ; update]
    const uint32_t* st;  //!< row structure (permuted indices)
    const int32_t* inv;  //!< map parent row -> own update row (size nr_parent)
    uint32_t* rowperm;   //!< [nc] row permutation of the pivot block
    uint32_t* lperm;     //!< [2][NB] row permutation of the pivot step, by step parity
    const uint32_t* gptr;  //!< [nr + 1] child contributions of row i: gsrc[gptr[i] .. gptr[i+1])
    uint32_t nr, nc, c0;
    uint32_t child_begin, child_end;
    int32_t parent;
};

struct Task {
    uint32_t front, chunk;
};

/* ---------------------------------------------------------------- kernels */

struct ScatterA {
    double* fronts;
    const uint64_t* dest;
    const double* vals;
    SANM_D void operator()(size_t p) const { fronts[dest[p]] = vals[p]; }
};

struct AbsMaxKernel {
    const double* v;
    uint64_t n, nblk;
    double* out;  // out[bid] = max |v| over the block's share
    SANM_D void operator()(const rt::BlockCtx& c) const {
        double* red = reinterpret_cast<double*>(c.smem);
        double m = 0;
        const uint64_t per = (n + nblk - 1) / nblk;
        const uint64_t i0 = c.bid * per, i1 = i0 + per < n ? i0 + per : n;
        for (uint64_t i = i0 + c.tid; i < i1; i += c.nthr) {
            double a = fabs(v[i]);
            m = a > m ? a : m;
        }
        red[c.tid] = m;
        c.sync();
        for (int s = c.nthr / 2; s > 0; s >>= 1) {
            if (c.tid < s) red[c.tid] = red[c.tid] > red[c.tid + s] ? red[c.tid] : red[c.tid + s];
            c.sync();
        }
        if (c.tid == 0) out[c.bid] = red[0];
    }
};

//! F[i,j] += sum over children of their Schur complement entry (gather form)
struct AssembleChildren {
    const FrontD* fronts;
    const uint32_t* child_list;
    const Task* tasks;
    struct Pre {
        Task tk;
        FrontD f;
    };
    //! descriptors only (rt.h HasPre)
    SANM_D Pre pre(const rt::BlockCtx& c) const {
        Pre p;
        p.tk = tasks[c.bid];
        p.f = fronts[p.tk.front];
        return p;
    }
    //! what an entry needs of a child: kept in shared memory for the block
    //! instead of re-reading the 100-byte front descriptor per entry and child
    struct ChildD {
        const double* U;  //!< update block: U[jj * nr + ii]
        const int32_t* inv;
        uint32_t nr;
    };
    static constexpr uint32_t MAX_CH = 16;
    SANM_D void operator()(const rt::BlockCtx& c, const Pre& p) const {
        const Task tk = p.tk;
        const FrontD& f = p.f;
        const uint64_t total = (uint64_t)f.nr * f.nr;
        const uint64_t e0 = (uint64_t)tk.chunk * ASM_TILE;
        const uint64_t e1 = e0 + ASM_TILE < total ? e0 + ASM_TILE : total;
        const uint32_t nch = f.child_end - f.child_begin;
        if (nch <= MAX_CH) {
            ChildD* ch = reinterpret_cast<ChildD*>(c.smem);
            for (uint32_t q = c.tid; q < nch; q += c.nthr) {
                const FrontD d = fronts[child_list[f.child_begin + q]];
                ch[q] = ChildD{d.F + (uint64_t)d.nc * d.nr + d.nc, d.inv, d.nr};
            }
            c.sync();
            for (uint64_t e = e0 + c.tid; e < e1; e += c.nthr) {
                const uint32_t e32 = (uint32_t)e;  // nr^2 < 2^32 (checked at analysis)
                const uint32_t i = e32 % f.nr, j = e32 / f.nr;
                double acc = 0;
                bool any = false;
                for (uint32_t q = 0; q < nch; ++q) {
                    // child order: the sum is the one of the plain loop below
                    const int32_t ii = ch[q].inv[i], jj = ch[q].inv[j];
                    if (ii >= 0 && jj >= 0) {
                        acc += ch[q].U[(uint64_t)jj * ch[q].nr + ii];
                        any = true;
                    }
                }
                // entries no child reaches are not touched (no read-modify-write)
                if (any) f.F[e] += acc;
            }
            return;
        }
        for (uint64_t e = e0 + c.tid; e < e1; e += c.nthr) {
            const uint32_t e32 = (uint32_t)e;
            const uint32_t i = e32 % f.nr, j = e32 / f.nr;
            double acc = 0;
            for (uint32_t q = f.child_begin; q < f.child_end; ++q) {
                const FrontD ch = fronts[child_list[q]];
                const int32_t ii = ch.inv[i], jj = ch.inv[j];
                if (ii >= 0 && jj >= 0) {
                    acc += ch.F[(uint64_t)(ch.nc + jj) * ch.nr + ch.nc + ii];
                }
            }
            f.F[e] += acc;
        }
    }
};

/*!
 * In-place Gauss-Jordan on the first nb pivots starting at k0 of an m x m
 * column-major matrix M (leading dim ld), pivoting among rows [k, plim).
 * Runs inside one CTA; `sh` is one shared int used to broadcast the pivot row.
 */
SANM_D inline void gj_steps(const rt::BlockCtx& c, double* M, uint32_t ld,
                            uint32_t m, uint32_t k0, uint32_t nb, uint32_t plim,
                            uint32_t* rowperm, int* sh, double thresh,
                            int* nr_perturbed) {
    for (uint32_t k = k0; k < k0 + nb; ++k) {
#ifndef SANM_HOSTEMU
        if (c.tid < 32) {
            // warp 0: arg max of |M[i,k]| over i in [k, plim), ties -> lowest i
            double best = -1.0;
            uint32_t r = k;
            for (uint32_t i = k + c.tid; i < plim; i += 32) {
                double a = fabs(M[(uint64_t)k * ld + i]);
                if (a > best) {
                    best = a;
                    r = i;
                }
            }
            for (int off = 16; off > 0; off >>= 1) {
                double ob = __shfl_xor_sync(0xffffffffu, best, off);
                uint32_t orr = __shfl_xor_sync(0xffffffffu, r, off);
                if (ob > best || (ob == best && orr < r)) {
                    best = ob;
                    r = orr;
                }
            }
            if (c.tid == 0) {
                if (!(best > thresh)) {
                    double p = M[(uint64_t)k * ld + r];
                    M[(uint64_t)k * ld + r] = p < 0 ? -thresh : thresh;
                    rt_atomic_inc(nr_perturbed);
                }
                *sh = (int)r;
            }
        }
#else
        if (c.tid == 0) {
            uint32_t r = k;
            double best = fabs(M[(uint64_t)k * ld + k]);
            for (uint32_t i = k + 1; i < plim; ++i) {
                double a = fabs(M[(uint64_t)k * ld + i]);
                if (a > best) {
                    best = a;
                    r = i;
                }
            }
            if (!(best > thresh)) {
                // static pivot perturbation (as PARDISO does for tiny pivots)
                double p = M[(uint64_t)k * ld + r];
                M[(uint64_t)k * ld + r] = p < 0 ? -thresh : thresh;
                rt_atomic_inc(nr_perturbed);
            }
            *sh = (int)r;
        }
#endif
        c.sync();
        const uint32_t r = (uint32_t)*sh;
        if (r != k) {
            for (uint32_t j = c.tid; j < m; j += c.nthr) {
                double a = M[(uint64_t)j * ld + k];
                M[(uint64_t)j * ld + k] = M[(uint64_t)j * ld + r];
                M[(uint64_t)j * ld + r] = a;
            }
            if (c.tid == 0 && rowperm) {
                uint32_t a = rowperm[k];
                rowperm[k] = rowperm[r];
                rowperm[r] = a;
            }
        }
        c.sync();
        const double ip = 1.0 / M[(uint64_t)k * ld + k];
        for (uint32_t j = c.tid; j < m; j += c.nthr) {
            if (j != k) M[(uint64_t)j * ld + k] *= ip;
        }
        c.sync();
        {
            // warp w takes columns w, w + nw, ...; lanes run down the column
            const uint32_t wl = c.nthr >= 32 ? 32u : (uint32_t)c.nthr;
            const uint32_t lane = c.tid % wl, w = c.tid / wl, nw = c.nthr / wl;
            const double* __restrict__ colk = M + (uint64_t)k * ld;
            for (uint32_t j = w; j < m; j += nw) {
                if (j == k) continue;
                double* __restrict__ col = M + (uint64_t)j * ld;
                const double rj = col[k];
                for (uint32_t i = lane; i < m; i += wl) {
                    if (i != k) col[i] -= colk[i] * rj;
                }
            }
        }
        c.sync();
        for (uint32_t i = c.tid; i < m; i += c.nthr) {
            double* p = M + (uint64_t)k * ld + i;
            *p = i == k ? ip : -(*p) * ip;
        }
        c.sync();
    }
}

//! whole small front in shared memory, one CTA per front
struct SmallFrontKernel {
    const FrontD* fronts;
    const uint32_t* list;
    const double* amax;
    int* nr_perturbed;
    SANM_D FrontD pre(const rt::BlockCtx& c) const { return fronts[list[c.bid]]; }
    SANM_D void operator()(const rt::BlockCtx& c, const FrontD& f) const {
        double* M = reinterpret_cast<double*>(c.smem);
        int* sh = reinterpret_cast<int*>(M + (uint64_t)f.nr * f.nr);
        const uint64_t tot = (uint64_t)f.nr * f.nr;
        for (uint64_t e = c.tid; e < tot; e += c.nthr) M[e] = f.F[e];
        for (uint32_t i = c.tid; i < f.nc; i += c.nthr) f.rowperm[i] = i;
        c.sync();
        gj_steps(c, M, f.nr, f.nr, 0, f.nc, f.nc, f.rowperm, sh,
                 amax[0] * 1e-13, nr_perturbed);
        for (uint64_t e = c.tid; e < tot; e += c.nthr) f.F[e] = M[e];
    }
};

/* ---- large fronts: blocked Gauss-Jordan in global memory, NB pivots a step */

//! invert the diagonal block (with pivoting inside it) and apply its row
//! swaps to the rest of the block row
struct BigDiagKernel {
    const FrontD* fronts;
    const uint32_t* list;
    uint32_t kb;
    const double* amax;
    int* nr_perturbed;
    //! descriptor only (rt.h HasPre)
    SANM_D FrontD pre(const rt::BlockCtx& c) const { return fronts[list[c.bid]]; }
    SANM_D void operator()(const rt::BlockCtx& c) const { (*this)(c, pre(c)); }
    SANM_D void operator()(const rt::BlockCtx& c, const FrontD& f) const {
        if (kb >= f.nc) return;
        const uint32_t nb = f.nc - kb < (uint32_t)NB ? f.nc - kb : (uint32_t)NB;
        double* D = reinterpret_cast<double*>(c.smem);  // nb x nb, ld NB
        uint32_t* lperm = reinterpret_cast<uint32_t*>(D + NB * NB);
        int* sh = reinterpret_cast<int*>(lperm + NB);
#ifndef SANM_HOSTEMU
        {
            // 4 warps, warp w owns columns 8w..8w+7 (gj32.cuh)
            const uint32_t lane = c.tid & 31, j0 = (c.tid >> 5) * 8;
            double* __restrict__ G = f.F + (uint64_t)kb * f.nr + kb;
            double v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                v[j] = (lane < nb && j0 + j < nb) ? G[(uint64_t)(j0 + j) * f.nr + lane] : 0.0;
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) D[(j0 + j) * NB + lane] = v[j];
            double* acol = reinterpret_cast<double*>(sh + 2);
            gj32_sync();
            block_gj32(D, nb, lperm, acol, amax[0] * 1e-13, nr_perturbed);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (lane < nb && j0 + j < nb) G[(uint64_t)(j0 + j) * f.nr + lane] = D[(j0 + j) * NB + lane];
            }
        }
#else
        for (uint32_t e = c.tid; e < nb * nb; e += c.nthr) {
            uint32_t i = e % nb, j = e / nb;
            D[j * NB + i] = f.F[(uint64_t)(kb + j) * f.nr + kb + i];
        }
        for (uint32_t i = c.tid; i < nb; i += c.nthr) lperm[i] = i;
        c.sync();
        gj_steps(c, D, NB, nb, 0, nb, nb, lperm, sh, amax[0] * 1e-13,
                 nr_perturbed);
        for (uint32_t e = c.tid; e < nb * nb; e += c.nthr) {
            uint32_t i = e % nb, j = e / nb;
            f.F[(uint64_t)(kb + j) * f.nr + kb + i] = D[j * NB + i];
        }
#endif
        // row i of the block now holds original row lperm[i]; the rest of
        // the block row is permuted by BigRowPanelKernel while it loads it
        // by step parity: the row panel of step kb - NB may still be reading its own
        for (uint32_t i = c.tid; i < nb; i += c.nthr) f.lperm[((kb / NB) & 1) * NB + i] = lperm[i];
        c.sync();
        // pivoting never leaves the block: position kb + i of the front holds
        // the original row kb + lperm[i]
        for (uint32_t i = c.tid; i < nb; i += c.nthr) f.rowperm[kb + i] = kb + lperm[i];
    }
};

/* The tile kernels of the narrow path run one wave of 64 x 64 tiles per launch
 * on the top fronts, two warps per scheduler: what a launch costs is the
 * number of instructions a warp executes, not bandwidth (ncu on a 121-tile
 * launch: 12 cycles per issued instruction, two thirds of the samples in the
 * index arithmetic and predicates of the tile fills and of the epilogue).  So
 * the tiles have compile-time shapes: fills map element -> (row, column) with
 * shifts, every load of a thread is issued before its first shared-memory
 * store, entries outside the front or past the last pivot are ZERO so that the
 * product loops run a fixed NB iterations, and the pivot rows / columns are
 * masked once per thread. */
#ifdef SANM_HOSTEMU
#define SANM_TILE_THREADS(c) ((uint32_t)(c).nthr)
#else
#define SANM_TILE_THREADS(c) 256u
#endif

//! dst[e] = src(e) (address or null -> 0) for e < TOTAL; CUDA: 256 threads,
//! TOTAL / 256 loads in flight per thread (read-only path)
template <int TOTAL, class Src>
SANM_D inline void stage_fixed(const rt::BlockCtx& c, double* dst, Src src) {
#ifdef SANM_HOSTEMU
    for (uint32_t e = c.tid; e < (uint32_t)TOTAL; e += c.nthr) {
        const double* p = src(e);
        dst[e] = p ? *p : 0.0;
    }
#else
    constexpr int PER = TOTAL / 256;
    double r[PER];
#pragma unroll
    for (int it = 0; it < PER; ++it) {
        const double* p = src((uint32_t)c.tid + it * 256);
        r[it] = p ? SANM_LDG(p) : 0.0;
    }
#pragma unroll
    for (int it = 0; it < PER; ++it) dst[c.tid + it * 256] = r[it];
#endif
}

//! R' = Dinv * R for a tile of 64 columns outside the pivot block
struct BigRowPanelKernel {
    const FrontD* fronts;
    const uint32_t* list;
    uint32_t kb, tiles;
    SANM_D FrontD pre(const rt::BlockCtx& c) const { return fronts[list[c.bid / tiles]]; }
    SANM_D void operator()(const rt::BlockCtx& c) const { (*this)(c, pre(c)); }
    SANM_D void operator()(const rt::BlockCtx& c, const FrontD& f) const {
        const uint32_t tile = (uint32_t)(c.bid % tiles);
        if (kb >= f.nc || tile * 64 >= f.nr) return;
        const uint32_t nb = f.nc - kb < (uint32_t)NB ? f.nc - kb : (uint32_t)NB;
        double* Di = reinterpret_cast<double*>(c.smem);  // [NB][NB] : Di[m*NB+i]
        double* R = Di + NB * NB;                        // [64][NB] : R[j*NB+m]
        const uint32_t j0 = tile * 64;
        const uint32_t nj = f.nr - j0 < 64 ? f.nr - j0 : 64;
        // rows of the pivot block were swapped by BigDiagKernel (lperm)
        const uint32_t* lp = f.lperm + ((kb / NB) & 1) * NB;
        const double* Fk = f.F + kb;
        stage_fixed<NB * NB>(c, Di, [&](uint32_t e) -> const double* {
            const uint32_t i = e % NB, m = e / NB;
            return i < nb && m < nb ? Fk + (uint64_t)(kb + m) * f.nr + i : nullptr;
        });
        stage_fixed<NB * 64>(c, R, [&](uint32_t e) -> const double* {
            const uint32_t m = e % NB, j = e / NB;
            return m < nb && j < nj ? Fk + (uint64_t)(j0 + j) * f.nr + lp[m] : nullptr;
        });
        c.sync();
        // 4 rows x 2 columns per thread (8 x 32 thread grid in the CUDA build):
        // 8 FMAs per 6 shared loads instead of 1 per 2 -- the plain dot-product
        // form is bound by shared-memory bandwidth
        for (uint32_t q = c.tid; q < 256; q += SANM_TILE_THREADS(c)) {
            const uint32_t i4 = (q % 8) * 4, j2 = (q / 8) * 2;
            double acc[4][2] = {};
#pragma unroll 8
            for (uint32_t m = 0; m < (uint32_t)NB; ++m) {
                double a[4], b[2];
                for (int u = 0; u < 4; ++u) a[u] = Di[m * NB + i4 + u];
                for (int v = 0; v < 2; ++v) b[v] = R[(j2 + v) * NB + m];
                for (int u = 0; u < 4; ++u)
                    for (int v = 0; v < 2; ++v) acc[u][v] += a[u] * b[v];
            }
            for (int v = 0; v < 2; ++v) {
                const uint32_t gj = j0 + j2 + v;
                if (j2 + v >= nj || (gj >= kb && gj < kb + nb)) continue;
                for (int u = 0; u < 4; ++u) {
                    if (i4 + u < nb) f.F[(uint64_t)gj * f.nr + kb + i4 + u] = acc[u][v];
                }
            }
        }
    }
};

/*! T -= C * R' on a 64x64 tile (rows and columns of the pivot block excluded).
 * R' = Dinv * R is recomputed per tile from the untouched block row (a third
 * of the tile's flops, in a latency-bound kernel) instead of being read back
 * from a row-panel launch before it: BigRowPanelKernel stores R' in the launch
 * AFTER this one, beside the column panel and the next pivot block. */
struct BigGemmKernel {
    const FrontD* fronts;
    const uint32_t* list;
    uint32_t kb, tiles;  // tiles per dimension
    SANM_D FrontD pre(const rt::BlockCtx& c) const {
        return fronts[list[c.bid / (tiles * tiles)]];
    }
    SANM_D void operator()(const rt::BlockCtx& c) const { (*this)(c, pre(c)); }
    SANM_D void operator()(const rt::BlockCtx& c, const FrontD& f) const {
        const uint32_t per = tiles * tiles;
        const uint32_t t = (uint32_t)(c.bid % per), ti = t % tiles, tj = t / tiles;
        if (kb >= f.nc || ti * 64 >= f.nr || tj * 64 >= f.nr) return;
        const uint32_t nb = f.nc - kb < (uint32_t)NB ? f.nc - kb : (uint32_t)NB;
        double* Cs = reinterpret_cast<double*>(c.smem);  // [NB][64] : Cs[m*64+i]
        double* Rs = Cs + NB * 64;                       // [64][NB] : Rs[j*NB+m]
        double* Rr = Rs + NB * 64;                       // [64][NB] : R before Dinv
        double* Di = Rr + NB * 64;                       // [NB][NB] : Di[m*NB+i]
        const uint32_t* lp = f.lperm + ((kb / NB) & 1) * NB;
        const uint32_t i0 = ti * 64, j0 = tj * 64;
        const uint32_t ni = f.nr - i0 < 64 ? f.nr - i0 : 64;
        const uint32_t nj = f.nr - j0 < 64 ? f.nr - j0 : 64;
        // local rows / columns that belong to the pivot block of this step
        const uint32_t pr0 = kb > i0 ? kb - i0 : 0, pr1 = kb + nb > i0 ? kb + nb - i0 : 0;
        const uint32_t pc0 = kb > j0 ? kb - j0 : 0, pc1 = kb + nb > j0 ? kb + nb - j0 : 0;
        auto row_ok = [&](uint32_t li) { return li < ni && !(li >= pr0 && li < pr1); };
        auto col_ok = [&](uint32_t lj) { return lj < nj && !(lj >= pc0 && lj < pc1); };
        // The accumulators start from the tile itself (T - C R' instead of a
        // read-modify-write epilogue of 16 dependent round trips per thread)
        double* T0 = f.F + (uint64_t)j0 * f.nr + i0;
        auto load_tile = [&](double (*acc)[4], uint32_t q) {
            const uint32_t qi = (q % 16) * 4, qj = (q / 16) * 4;
            for (int v = 0; v < 4; ++v)
                for (int u = 0; u < 4; ++u)
                    acc[u][v] = row_ok(qi + u) && col_ok(qj + v)
                                        ? T0[(uint64_t)(qj + v) * f.nr + qi + u]
                                        : 0.0;
        };
        double tacc[4][4];
        const bool early = SANM_TILE_THREADS(c) >= 256;  // one 4 x 4 block per thread
        if (early) load_tile(tacc, c.tid);
        const double* Fk = f.F + kb;
        stage_fixed<NB * 64>(c, Cs, [&](uint32_t e) -> const double* {
            const uint32_t i = e % 64, m = e / 64;
            return i < ni && m < nb ? f.F + (uint64_t)(kb + m) * f.nr + i0 + i : nullptr;
        });
        stage_fixed<NB * 64>(c, Rr, [&](uint32_t e) -> const double* {
            const uint32_t m = e % NB, j = e / NB;
            return m < nb && j < nj ? Fk + (uint64_t)(j0 + j) * f.nr + lp[m] : nullptr;
        });
        stage_fixed<NB * NB>(c, Di, [&](uint32_t e) -> const double* {
            const uint32_t i = e % NB, m = e / NB;
            return i < nb && m < nb ? Fk + (uint64_t)(kb + m) * f.nr + i : nullptr;
        });
        c.sync();
        // same operation order as BigRowPanelKernel
        for (uint32_t q = c.tid; q < 256; q += SANM_TILE_THREADS(c)) {
            const uint32_t i4 = (q % 8) * 4, j2 = (q / 8) * 2;
            double acc[4][2] = {};
#pragma unroll 8
            for (uint32_t m = 0; m < (uint32_t)NB; ++m) {
                double a[4], b[2];
                for (int u = 0; u < 4; ++u) a[u] = Di[m * NB + i4 + u];
                for (int v = 0; v < 2; ++v) b[v] = Rr[(j2 + v) * NB + m];
                for (int u = 0; u < 4; ++u)
                    for (int v = 0; v < 2; ++v) acc[u][v] += a[u] * b[v];
            }
            for (int v = 0; v < 2; ++v)
                for (int u = 0; u < 4; ++u) Rs[(j2 + v) * NB + i4 + u] = acc[u][v];
        }
        c.sync();
        // 16 x 16 thread grid in the CUDA build (256 threads), 4x4 per thread;
        // generic strided form so that the host build (1 thread) is the same code
        for (uint32_t q = c.tid; q < 256; q += SANM_TILE_THREADS(c)) {
            const uint32_t qi = (q % 16) * 4, qj = (q / 16) * 4;
            if (!early) load_tile(tacc, q);
#pragma unroll 8
            for (uint32_t m = 0; m < (uint32_t)NB; ++m) {
                double a[4], b[4];
                for (int u = 0; u < 4; ++u) a[u] = Cs[m * 64 + qi + u];
                for (int u = 0; u < 4; ++u) b[u] = Rs[(qj + u) * NB + m];
                for (int u = 0; u < 4; ++u)
                    for (int v = 0; v < 4; ++v) tacc[u][v] -= a[u] * b[v];
            }
            bool ro[4], co[4];
            for (int u = 0; u < 4; ++u) ro[u] = row_ok(qi + u);
            for (int v = 0; v < 4; ++v) co[v] = col_ok(qj + v);
            for (int v = 0; v < 4; ++v) {
                if (!co[v]) continue;
                double* col = T0 + (uint64_t)(qj + v) * f.nr + qi;
                for (int u = 0; u < 4; ++u) {
                    if (ro[u]) col[u] = tacc[u][v];
                }
            }
        }
    }
};

//! C' = -C * Dinv for a tile of 64 rows outside the pivot block
struct BigColPanelKernel {
    const FrontD* fronts;
    const uint32_t* list;
    uint32_t kb, tiles;
    SANM_D FrontD pre(const rt::BlockCtx& c) const { return fronts[list[c.bid / tiles]]; }
    SANM_D void operator()(const rt::BlockCtx& c) const { (*this)(c, pre(c)); }
    SANM_D void operator()(const rt::BlockCtx& c, const FrontD& f) const {
        const uint32_t tile = (uint32_t)(c.bid % tiles);
        if (kb >= f.nc || tile * 64 >= f.nr) return;
        const uint32_t nb = f.nc - kb < (uint32_t)NB ? f.nc - kb : (uint32_t)NB;
        double* Di = reinterpret_cast<double*>(c.smem);  // [NB][NB] : Di[q*NB+m]
        double* Cs = Di + NB * NB;                       // [NB][64] : Cs[m*64+i]
        const uint32_t i0 = tile * 64;
        const uint32_t ni = f.nr - i0 < 64 ? f.nr - i0 : 64;
        stage_fixed<NB * NB>(c, Di, [&](uint32_t e) -> const double* {
            const uint32_t i = e % NB, j = e / NB;
            return i < nb && j < nb ? f.F + (uint64_t)(kb + j) * f.nr + kb + i : nullptr;
        });
        stage_fixed<NB * 64>(c, Cs, [&](uint32_t e) -> const double* {
            const uint32_t i = e % 64, m = e / 64;
            return i < ni && m < nb ? f.F + (uint64_t)(kb + m) * f.nr + i0 + i : nullptr;
        });
        c.sync();
        // 4 rows x 2 columns per thread (16 x 16 thread grid), see the row panel
        for (uint32_t t = c.tid; t < 256; t += SANM_TILE_THREADS(c)) {
            const uint32_t i4 = (t % 16) * 4, q2 = (t / 16) * 2;
            double acc[4][2] = {};
#pragma unroll 8
            for (uint32_t m = 0; m < (uint32_t)NB; ++m) {
                double a[4], b[2];
                for (int u = 0; u < 4; ++u) a[u] = Cs[m * 64 + i4 + u];
                for (int v = 0; v < 2; ++v) b[v] = Di[(q2 + v) * NB + m];
                for (int u = 0; u < 4; ++u)
                    for (int v = 0; v < 2; ++v) acc[u][v] += a[u] * b[v];
            }
            for (int v = 0; v < 2; ++v) {
                if (q2 + v >= nb) continue;
                for (int u = 0; u < 4; ++u) {
                    const uint32_t gi = i0 + i4 + u;
                    if (i4 + u >= ni || (gi >= kb && gi < kb + nb)) continue;
                    f.F[(uint64_t)(kb + q2 + v) * f.nr + gi] = -acc[u][v];
                }
            }
        }
    }
};

/*! Second launch of a pivot step: the two panels of step kb and the pivot
 * block of step kb + NB.  The three are independent once the GEMM of step kb
 * is done: the row panel rewrites rows kb .. kb+NB, the column panel columns
 * kb .. kb+NB (both outside the inverted block, which they only read), and
 * the next pivot block lies in rows and columns kb+NB ..  The pivot block is
 * a single 128-thread block per front (~10 us of dependent shared-memory
 * round trips); beside the panels it costs no launch of its own.  Blocks
 * [0, ndiag) are the pivot blocks (scheduled first; ndiag = 0 on the last
 * step), then npanel column tiles, then npanel row tiles. */
struct BigPanelsDiagKernel {
    BigColPanelKernel col;
    BigRowPanelKernel row;
    BigDiagKernel diag;
    uint32_t ndiag, npanel;
    SANM_D FrontD pre(const rt::BlockCtx& c) const {
        rt::BlockCtx c2 = c;
        if (c.bid < ndiag) return diag.pre(c2);
        if (c.bid < ndiag + npanel) {
            c2.bid = c.bid - ndiag;
            return col.pre(c2);
        }
        c2.bid = c.bid - ndiag - npanel;
        return row.pre(c2);
    }
    SANM_D void operator()(const rt::BlockCtx& c, const FrontD& f) const {
        rt::BlockCtx c2 = c;
        if (c.bid < ndiag) {
            // gj32.cuh wants exactly 4 warps; warps 4..7 leave as a whole
            if (c.tid >= 128) return;
            if (c2.nthr > 128) c2.nthr = 128;
            diag(c2, f);
        } else if (c.bid < ndiag + npanel) {
            c2.bid = c.bid - ndiag;
            col(c2, f);
        } else {
            c2.bid = c.bid - ndiag - npanel;
            row(c2, f);
        }
    }
};

#include "sparse_lu_wide.h"

/* ---- solves */

/* Solve kernels.  A task is one chunk of ROWS rows of one front; a block of
 * ROWS x GROUPS threads.  Thread (g, r) owns row r of the chunk and
 * accumulates columns g, g + GROUPS, ...; the partial sums of the groups are
 * added in a fixed order through shared memory, so results are bitwise
 * reproducible.  What bounds a level is memory latency, not bandwidth (the
 * factors are L2-resident and a level has few fronts): chunks are therefore
 * only 8 rows (64-byte column segments, still whole sectors) so that a big
 * front spreads over nr/8 blocks, and every thread keeps 8 independent loads
 * in flight. */
//! backward sweep (rows of nr columns): 8-row chunks, 32 column groups for
//! levels of small fronts, 128 where a front has many columns
constexpr int BWD_ROWS = 8, BWD_GROUPS_NARROW = 32, BWD_GROUPS_WIDE = 128;
//! forward sweep (rows of only nc columns, and every chunk first rebuilds the
//! front's w1 from the children): fewer, larger chunks
constexpr int FWD_ROWS = 32, FWD_GROUPS_NARROW = 8, FWD_GROUPS_WIDE = 32;
constexpr uint32_t SOLVE_WIDE_COLS = 192;
constexpr int SOLVE_STAGE = 8;  //!< partials summed per thread in stage 1

//! first batch of solve_partial's factor loads (columns g, g + GROUPS, ...):
//! issued by the kernels' pre-wait prologues
template <int GROUPS>
SANM_HD inline void solve_prefetch(double* a, const double* __restrict__ col0,
                                   uint64_t ld, uint32_t ncol, int g) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
        const uint32_t jj = g + u * GROUPS;
        a[u] = jj < ncol ? SANM_LDG(col0 + (uint64_t)jj * ld) : 0.0;
    }
}

//! partial dot product of one row of a column-major block with v; a0 (or
//! null) = the first batch of factor entries, already loaded by solve_prefetch
template <int GROUPS>
SANM_HD inline double solve_partial(const double* __restrict__ col0, uint64_t ld,
                                    uint32_t ncol, const double* __restrict__ v,
                                    int g, const double* a0 = nullptr) {
    constexpr int U = 8;
    double acc[U] = {0, 0, 0, 0, 0, 0, 0, 0};
    uint32_t j = g;
    if (a0 && j < ncol) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t jj = j + u * GROUPS;
            acc[u] += a0[u] * (jj < ncol ? v[jj] : 0.0);
        }
        j += U * GROUPS;
    }
    for (; j < ncol; j += U * GROUPS) {
        double a[U], x[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t jj = j + u * GROUPS;
            const bool ok = jj < ncol;
            // the factors are read-only during a solve: ld.global.nc
            a[u] = ok ? SANM_LDG(col0 + (uint64_t)jj * ld) : 0.0;
            x[u] = ok ? v[jj] : 0.0;
        }
        // all U loads in flight before the first use (needs the bounded
        // launch, see rt.h k_for_each_block_b)
        SANM_TIE8(a);
#pragma unroll
        for (int u = 0; u < U; ++u) acc[u] += a[u] * x[u];
    }
    return ((acc[0] + acc[1]) + (acc[2] + acc[3])) +
           ((acc[4] + acc[5]) + (acc[6] + acc[7]));
}

/*! fixed-order sum of the GROUPS partials of every row: part[g * ROWS + r] ->
 * returns (for threads with c.tid < ROWS, after two syncs) the row sums in
 * part[0 .. ROWS) */
template <int SOLVE_ROWS, int GROUPS>
SANM_D inline void solve_reduce(const rt::BlockCtx& c, double* part) {
    constexpr int S1 = GROUPS / SOLVE_STAGE;  // stage-1 threads per row
    double* st1 = part + SOLVE_ROWS * GROUPS;
    c.sync();
    for (int t = c.tid; t < SOLVE_ROWS * S1; t += c.nthr) {
        const int r = t % SOLVE_ROWS, q = t / SOLVE_ROWS;
        double s = 0;
#pragma unroll
        for (int u = 0; u < SOLVE_STAGE; ++u) s += part[(q * SOLVE_STAGE + u) * SOLVE_ROWS + r];
        st1[q * SOLVE_ROWS + r] = s;
    }
    c.sync();
    for (int r = c.tid; r < SOLVE_ROWS; r += c.nthr) {
        double s = 0;
#pragma unroll
        for (int q = 0; q < S1; ++q) s += st1[q * SOLVE_ROWS + r];
        part[r] = s;
    }
    // the threads that continue with part[r] are the ones that wrote it
}
template <int SOLVE_ROWS, int GROUPS>
constexpr size_t solve_scratch_doubles() {
    return (size_t)SOLVE_ROWS * GROUPS + (size_t)SOLVE_ROWS * (GROUPS / SOLVE_STAGE);
}

/*! A gathered sum  base + sum_p vecs[gsrc[p]]  whose index loads were issued
 * before the kernel's dependency wait: the first two sources (a front rarely
 * has more than two children touching a row), the rest as a range. */
struct GatherIdx {
    uint32_t s0, s1, rest0, rest1;
};
constexpr uint32_t NO_SRC = 0xFFFFFFFFu;
SANM_HD inline GatherIdx gather_idx(const uint32_t* __restrict__ gptr,
                                    const uint32_t* __restrict__ gsrc, uint32_t row) {
    GatherIdx g;
    const uint32_t p0 = gptr[row], p1 = gptr[row + 1];
    g.s0 = p0 < p1 ? gsrc[p0] : NO_SRC;
    g.s1 = p0 + 1 < p1 ? gsrc[p0 + 1] : NO_SRC;
    g.rest0 = p0 + 2 < p1 ? p0 + 2 : p1;
    g.rest1 = p1;
    return g;
}
//! same summation order as the plain loop (child order)
SANM_HD inline double gather_sum(double base, const GatherIdx& g,
                                 const uint32_t* __restrict__ gsrc,
                                 const double* vecs) {
    // both loads in flight together
    const double c0 = vecs[g.s0 != NO_SRC ? g.s0 : 0];
    const double c1 = vecs[g.s1 != NO_SRC ? g.s1 : 0];
    double v = base;
    if (g.s0 != NO_SRC) v += c0;
    if (g.s1 != NO_SRC) v += c1;
    for (uint32_t p = g.rest0; p < g.rest1; ++p) v += vecs[gsrc[p]];
    return v;
}

//! forward: vec = [w1' ; w2 + F21' w1'] with w = b + gathered child updates;
//! b is read through the fill-reducing permutation
//! VG: the front's w1' was written to f.vec[0 .. nc) by SolvePrepKernel and is
//! read from global memory (fronts too large for a shared-memory copy)
//!
//! hoist: everything that does not depend on the level below -- index lists,
//! the row permutation, b, and the first batch of the thread's factor row --
//! is loaded in pre(), i.e. before griddepcontrol.wait, while the previous
//! level is still computing; what remains after the wait is one round trip for
//! the child updates, the shared-memory permutation and the reduction.  Only
//! data written at least two kernels earlier may be hoisted (a kernel's
//! prologue can overlap its predecessor, whose own wait guarantees everything
//! before it): the first launch of a sweep runs with hoist = 0.
template <int SOLVE_ROWS, int GROUPS, bool VG = false>
struct ForwardKernel {
    static constexpr int SOLVE_THREADS = SOLVE_ROWS * GROUPS;
    const FrontD* fronts;
    const uint32_t* gsrc;  // flattened child-update sources (offsets into vecs)
    const double* vecs;
    const Task* tasks;
    const double* b;       // right-hand side, original numbering
    const uint32_t* perm;  // new index -> original index
    int hoist;
    struct Pre {
        Task tk;
        FrontD f;
        GatherIdx gw, gt;  //!< w1 row tid; update row r0 + tid (tid < SOLVE_ROWS)
        double bv;
        uint32_t rp;
        double a[8];
    };
    SANM_D Pre pre(const rt::BlockCtx& c) const {
        Pre p;
        p.tk = tasks[c.bid];
        p.f = fronts[p.tk.front];
        if (!VG && hoist) {
            const FrontD& f = p.f;
            const uint32_t i = c.tid;
            if (i < f.nc) {
                p.gw = gather_idx(f.gptr, gsrc, i);
                p.bv = b[perm[f.c0 + i]];
                p.rp = f.rowperm[i];
            }
            const uint32_t nu = f.nr - f.nc, r0 = p.tk.chunk * SOLVE_ROWS;
            if (c.tid < SOLVE_ROWS && r0 + c.tid < nu) {
                p.gt = gather_idx(f.gptr, gsrc, f.nc + r0 + c.tid);
            }
            if (c.tid < SOLVE_THREADS) {
                const int lane = c.tid % SOLVE_ROWS, g = c.tid / SOLVE_ROWS;
                if (r0 + lane < nu) {
                    solve_prefetch<GROUPS>(p.a, f.F + f.nc + r0 + lane, f.nr, f.nc, g);
                }
            }
        }
        return p;
    }
    SANM_D void operator()(const rt::BlockCtx& c) const { (*this)(c, pre(c)); }
    SANM_D void operator()(const rt::BlockCtx& c, const Pre& p) const {
        const Task tk = p.tk;
        const FrontD& f = p.f;
        const bool h = !VG && hoist;
        double* w1 = reinterpret_cast<double*>(c.smem);
        const double* w1p = VG ? f.vec : w1 + f.nc;
        double* part = VG ? w1 : w1 + 2 * (size_t)f.nc;
        const uint32_t nu = f.nr - f.nc;
        const uint32_t r0 = tk.chunk * SOLVE_ROWS;
        // the update rows' child contributions do not depend on this front's
        // product: gathered up front, beside the w1 gather
        double tail = 0;
        if (h && c.tid < SOLVE_ROWS && r0 + c.tid < nu) {
            tail = gather_sum(0.0, p.gt, gsrc, vecs);
        }
        if (!VG) {
            double* w1p_s = w1 + f.nc;
            for (uint32_t i = c.tid; i < f.nc; i += c.nthr) {
                // child updates in child order (flattened at analysis time:
                // offsets into the work vectors), so the sum is the one of the
                // tree walk
                if (h && i == (uint32_t)c.tid) {
                    w1[i] = gather_sum(p.bv, p.gw, gsrc, vecs);
                } else {
                    double v = b[perm[f.c0 + i]];
                    for (uint32_t q = f.gptr[i], qe = f.gptr[i + 1]; q < qe; ++q) v += vecs[gsrc[q]];
                    w1[i] = v;
                }
            }
            c.sync();
            for (uint32_t i = c.tid; i < f.nc; i += c.nthr) {
                const uint32_t rp = (h && i == (uint32_t)c.tid) ? p.rp : f.rowperm[i];
                w1p_s[i] = w1[rp];
                if (tk.chunk == 0) f.vec[i] = w1p_s[i];
            }
            c.sync();
        }
        for (int th = c.tid; th < SOLVE_THREADS; th += c.nthr) {
            const int lane = th % SOLVE_ROWS, g = th / SOLVE_ROWS;
            const uint32_t r = r0 + lane;
            part[th] = r < nu ? solve_partial<GROUPS>(f.F + f.nc + r, f.nr, f.nc, w1p, g,
                                                      h && th == c.tid ? p.a : nullptr)
                              : 0.0;
        }
        solve_reduce<SOLVE_ROWS, GROUPS>(c, part);
        for (int lane = c.tid; lane < SOLVE_ROWS; lane += c.nthr) {
            const uint32_t r = r0 + lane;
            if (r >= nu) continue;
            double acc = tail;
            if (!(h && lane == c.tid)) {
                acc = 0;
                for (uint32_t q = f.gptr[f.nc + r], qe = f.gptr[f.nc + r + 1]; q < qe; ++q) {
                    acc += vecs[gsrc[q]];
                }
            }
            f.vec[f.nc + r] = acc + part[lane];
        }
    }
};

//! backward: x1 = G w1' - F12' x2, written to the permuted work vector (read
//! by descendants) and to the caller's vector in original numbering
//! VG: [w1' ; -x2] is f.vec (the tail written by SolvePrepKernel), read from
//! global memory
//! hoist: see ForwardKernel; after the wait only x2 (the ancestors' solution)
//! is fetched
template <int SOLVE_ROWS, int GROUPS, bool VG = false>
struct BackwardKernel {
    static constexpr int SOLVE_THREADS = SOLVE_ROWS * GROUPS;
    const FrontD* fronts;
    const Task* tasks;
    double* xp;  // permuted solution
    double* x;   // solution, original numbering
    const uint32_t* perm;
    int hoist;
    struct Pre {
        Task tk;
        FrontD f;
        double w;          //!< f.vec[tid] (tid < nc)
        uint32_t sti;      //!< f.st[nc + tid]
        uint32_t pm;       //!< perm[c0 + r0 + tid] (tid < SOLVE_ROWS)
        double a[8];
    };
    SANM_D Pre pre(const rt::BlockCtx& c) const {
        Pre p;
        p.tk = tasks[c.bid];
        p.f = fronts[p.tk.front];
        if (!VG && hoist) {
            const FrontD& f = p.f;
            const uint32_t i = c.tid, r0 = p.tk.chunk * SOLVE_ROWS;
            if (i < f.nc) p.w = f.vec[i];
            if (f.nc + i < f.nr) p.sti = f.st[f.nc + i];
            if (c.tid < SOLVE_ROWS && r0 + i < f.nc) p.pm = perm[f.c0 + r0 + i];
            if (c.tid < SOLVE_THREADS) {
                const int lane = c.tid % SOLVE_ROWS, g = c.tid / SOLVE_ROWS;
                if (r0 + lane < f.nc) {
                    solve_prefetch<GROUPS>(p.a, f.F + r0 + lane, f.nr, f.nr, g);
                }
            }
        }
        return p;
    }
    SANM_D void operator()(const rt::BlockCtx& c) const { (*this)(c, pre(c)); }
    SANM_D void operator()(const rt::BlockCtx& c, const Pre& p) const {
        const Task tk = p.tk;
        const FrontD& f = p.f;
        const bool h = !VG && hoist;
        double* vs = reinterpret_cast<double*>(c.smem);  // [w1' ; -x2]
        const double* v = VG ? f.vec : vs;
        double* part = VG ? vs : vs + f.nr;
        if (!VG) {
            for (uint32_t i = c.tid; i < f.nc; i += c.nthr) {
                vs[i] = (h && i == (uint32_t)c.tid) ? p.w : f.vec[i];
            }
            for (uint32_t i = f.nc + c.tid; i < f.nr; i += c.nthr) {
                const uint32_t sti = (h && i == f.nc + c.tid) ? p.sti : f.st[i];
                vs[i] = -xp[sti];
            }
            c.sync();
        }
        const uint32_t r0 = tk.chunk * SOLVE_ROWS;
        for (int th = c.tid; th < SOLVE_THREADS; th += c.nthr) {
            const int lane = th % SOLVE_ROWS, g = th / SOLVE_ROWS;
            const uint32_t r = r0 + lane;
            part[th] = r < f.nc ? solve_partial<GROUPS>(f.F + r, f.nr, f.nr, v, g,
                                                        h && th == c.tid ? p.a : nullptr)
                                : 0.0;
        }
        solve_reduce<SOLVE_ROWS, GROUPS>(c, part);
        for (int lane = c.tid; lane < SOLVE_ROWS; lane += c.nthr) {
            const uint32_t r = r0 + lane;
            if (r >= f.nc) continue;
            const double acc = part[lane];
            xp[f.c0 + r] = acc;
            x[(h && lane == c.tid) ? p.pm : perm[f.c0 + r]] = acc;
        }
    }
};

/*! Levels with fronts beyond the shared-memory budget (LuTopo::Level::
 * solve_global): one extra launch per sweep builds each front's input vector
 * once in global memory -- forward: w1' = P_r (b + child updates) into
 * vec[0 .. nc); backward: -x2 into vec[nc .. nr) (the forward sweep's update
 * part is dead by then) -- and the row-chunk kernels read it through L2. */
template <bool FORWARD>
struct SolvePrepKernel {
    const FrontD* fronts;
    const uint32_t* gsrc;
    const double* vecs;
    const Task* tasks;
    const double* b;
    const uint32_t* perm;
    const double* xp;
    struct Pre {
        Task tk;
        FrontD f;
    };
    SANM_D Pre pre(const rt::BlockCtx& c) const {
        Pre p;
        p.tk = tasks[c.bid];
        p.f = fronts[p.tk.front];
        return p;
    }
    SANM_D void operator()(const rt::BlockCtx& c, const Pre& p) const {
        const FrontD& f = p.f;
        const uint32_t i0 = p.tk.chunk * 256;
        const uint32_t i1 = i0 + 256 < f.nr ? i0 + 256 : f.nr;
        for (uint32_t i = i0 + c.tid; i < i1; i += c.nthr) {
            if (FORWARD) {
                if (i >= f.nc) break;
                const uint32_t j = f.rowperm[i];
                double v = b[perm[f.c0 + j]];
                for (uint32_t q = f.gptr[j], qe = f.gptr[j + 1]; q < qe; ++q) v += vecs[gsrc[q]];
                f.vec[i] = v;
            } else if (i >= f.nc) {
                f.vec[i] = -xp[f.st[i]];
            }
        }
    }
};

/* ---- the whole solve as ONE persistent launch: phase l < nlev is forward
 * level l, phase nlev + m is backward level nlev-1-m; a grid-wide barrier
 * (~1 us) between phases replaces 2 nlev dependent kernel launches.  Pays off
 * on small systems, where a level is only a few blocks of work; larger ones
 * run better as one launch per level, whose blocks have that level's shape. */
struct LevelD {
    uint32_t fwd_off, fwd_n, bwd_off, bwd_n;
    uint32_t fwd_wide, bwd_wide;
};

struct SolvePhased {
    const FrontD* fronts;
    const uint32_t* gsrc;
    const double* vecs;
    const Task* fwd_tasks;
    const Task* bwd_tasks;
    const LevelD* levels;
    uint32_t nlev;
    const double* b;
    const uint32_t* perm;
    double* xp;
    double* x;
    SANM_D void operator()(const rt::BlockCtx& c, uint32_t phase) const {
        rt::BlockCtx c2 = c;
        if (phase < nlev) {
            const LevelD L = levels[phase];
            for (size_t t = c.bid; t < L.fwd_n; t += c.nblk) {
                c2.bid = t;
                if (L.fwd_wide) {
                    ForwardKernel<FWD_ROWS, FWD_GROUPS_WIDE>{
                            fronts, gsrc, vecs, fwd_tasks + L.fwd_off, b, perm, 1}(c2);
                } else {
                    ForwardKernel<FWD_ROWS, FWD_GROUPS_NARROW>{
                            fronts, gsrc, vecs, fwd_tasks + L.fwd_off, b, perm, 1}(c2);
                }
                c.sync();  // shared memory is reused by the next task
            }
        } else {
            const LevelD L = levels[2 * nlev - 1 - phase];
            for (size_t t = c.bid; t < L.bwd_n; t += c.nblk) {
                c2.bid = t;
                if (L.bwd_wide) {
                    BackwardKernel<BWD_ROWS, BWD_GROUPS_WIDE>{
                            fronts, bwd_tasks + L.bwd_off, xp, x, perm, 1}(c2);
                } else {
                    BackwardKernel<BWD_ROWS, BWD_GROUPS_NARROW>{
                            fronts, bwd_tasks + L.bwd_off, xp, x, perm, 1}(c2);
                }
                c.sync();
            }
        }
    }
};

struct SpmvKernel {
    const uint32_t* rowptr;
    const uint32_t* colidx;
    const double* vals;
    const double* x;
    double* y;
    SANM_D void operator()(size_t i) const {
        double s = 0;
        for (uint32_t p = rowptr[i]; p < rowptr[i + 1]; ++p) s += vals[p] * x[colidx[p]];
        y[i] = s;
    }
};

}  // namespace

//! everything that depends on the sparsity pattern only: built once per
//! topology, shared (read-only, host and device) by all solver instances --
//! the ANM drivers create a new solver every iteration (libsanm/anm.cpp:205)
struct LuTopo {
    std::shared_ptr<const Symbolic> sym;
    std::vector<uint32_t> h_rowptr, h_colidx;
    rt::DBuf<uint32_t> rowptr, colidx, perm, structs, child_list;
    rt::DBuf<uint32_t> gptr, gsrc;
    rt::DBuf<int32_t> inv;
    rt::DBuf<uint64_t> a_dest;
    // per-level launch lists
    struct Level {
        uint32_t asm_off = 0, asm_n = 0;      //!< assembly tasks
        uint32_t small_off = 0, small_n = 0;  //!< list of small fronts
        uint32_t big_off = 0, big_n = 0;      //!< list of large fronts
        uint32_t big_max_nc = 0, big_max_nr = 0, small_max_nr = 0;
        uint32_t wide_off = 0, wide_n = 0;    //!< list of wide fronts (superblock path)
        uint32_t wide_max_nc = 0, wide_max_nr = 0;
        //! solve: work vectors of this level's fronts stay in global memory
        bool solve_global = false;
        uint32_t prep_off = 0, prep_n = 0;    //!< (front, 256-row chunk) tasks
        uint32_t fwd_off = 0, fwd_n = 0, bwd_off = 0, bwd_n = 0;
        uint32_t max_nc = 0, max_nr = 0;
        //! the level's tasks are ordered narrow fronts first (forward: nc, backward:
        //! nr below SOLVE_WIDE_COLS): [off, off + split) can run in narrow blocks
        uint32_t fwd_split = 0, bwd_split = 0;
        uint32_t fwd_narrow_nc = 0, bwd_narrow_nr = 0;  //!< maxima over the narrow part
    };
    std::vector<Level> levels;
    rt::DBuf<Task> asm_tasks, fwd_tasks, bwd_tasks;
    rt::DBuf<LevelD> level_d;
    size_t solve_smem = 0;
    rt::DBuf<uint32_t> small_list, big_list, wide_list;
    rt::DBuf<Task> prep_tasks;
    uint32_t max_front = 0;
    uint32_t wide_min = 0, smem_nr = 0;  //!< thresholds the lists were built with
};

struct SparseLU::Impl {
    std::shared_ptr<LuTopo> topo;
    // numeric state of this instance
    rt::DBuf<uint32_t> rowperm, lperm;
    rt::DBuf<double> fronts, vecs, vals, bp, xp, scal;
    rt::DBuf<int> counters;
    rt::DBuf<FrontD> fd;
    bool factored = false;
    bool counters_read = false;
    bool bench_done = false;
    rt::DBuf<double> ref_b, ref_r, ref_dx;  //!< iterative refinement (on demand)
};

SparseLU::SparseLU() : m_impl{new Impl} {}
SparseLU::~SparseLU() = default;

const uint32_t* SparseLU::d_rowptr() const { return m_impl->topo->rowptr.ptr(); }
const uint32_t* SparseLU::d_colidx() const { return m_impl->topo->colidx.ptr(); }

void SparseLU::analyze(uint64_t n, const uint32_t* rowptr,
                       const uint32_t* colidx, const double* node_keys) {
    rt::check_device();
    sanm_check(n > 0 && n < (1ull << 31));
    // SANM_LU_DEBUG: host wall time of the stages of this call
    const bool dbg = getenv("SANM_LU_DEBUG") != nullptr;
    auto lap_t0 = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (!dbg) return;
        rt::sync();
        auto t1 = std::chrono::steady_clock::now();
        printf("[lu-setup] %-34s %8.2f ms\n", what,
               std::chrono::duration<double, std::milli>(t1 - lap_t0).count());
        lap_t0 = t1;
    };
    Impl& I = *m_impl;
    m_n = n;
    m_nnz = rowptr[n];
    // Everything that depends on the pattern only (ordering, symbolic
    // factorisation, launch lists, their device copies) is shared between
    // solver instances: the ANM drivers build a new solver every iteration.
    static LruCache<LuTopo> topo_cache{6};
    const uint64_t key = hash_bytes(
            colidx, m_nnz * sizeof(uint32_t),
            hash_bytes(rowptr, (n + 1) * sizeof(uint32_t), n * 0x51ED27ull + m_nnz));
    const uint32_t wide_min = wide_min_nr(), smem_nr = solve_smem_nr();
    I.topo = topo_cache.get(key);
    bool fresh = false;
    if (!I.topo || I.topo->wide_min != wide_min || I.topo->smem_nr != smem_nr ||
        I.topo->h_rowptr.size() != n + 1 ||
        I.topo->h_colidx.size() != m_nnz ||
        memcmp(I.topo->h_rowptr.data(), rowptr, (n + 1) * sizeof(uint32_t)) ||
        memcmp(I.topo->h_colidx.data(), colidx, m_nnz * sizeof(uint32_t))) {
        fresh = true;
        I.topo = std::make_shared<LuTopo>();
        I.topo->wide_min = wide_min;
        I.topo->smem_nr = smem_nr;
        I.topo->h_rowptr.assign(rowptr, rowptr + n + 1);
        I.topo->h_colidx.assign(colidx, colidx + m_nnz);
        I.topo->sym = std::make_shared<const Symbolic>(
                symbolic_analyze((uint32_t)n, rowptr, colidx, node_keys));
        topo_cache.put(key, I.topo);
    }
    LuTopo& T = *I.topo;
    const Symbolic& S = *T.sym;
    lap(fresh ? "pattern hash + symbolic analysis" : "pattern hash + cache hit");

    if (fresh) {
        T.rowptr.upload(T.h_rowptr);
        T.colidx.upload(T.h_colidx);
        T.perm.upload(S.perm);
        T.structs.upload(S.structs);
        T.child_list.upload(S.child_list);
        T.inv.upload(S.inv);
        T.a_dest.upload(S.a_dest);
    }
    I.fronts.alloc(S.front_total);
    I.vecs.alloc(S.soff[S.nsn]);
    I.rowperm.alloc(n);
    I.lperm.alloc((size_t)S.nsn * LPERM_PER_FRONT);
    I.vals.alloc(m_nnz);
    I.bp.alloc(n);
    I.xp.alloc(n);
    I.scal.alloc(4 + ABSMAX_BLOCKS);
    I.counters.alloc(4);
    I.counters.zero();
    lap("uploads + numeric buffers");

    // flattened gather lists of the forward sweep: for row i of front s the
    // positions (in the work-vector array) of the children's update entries
    // that land on it, in child order
    if (fresh) {
        std::vector<uint32_t> gptr(S.soff[S.nsn] + S.nsn + 1), gsrc;
        size_t at = 0;
        for (uint32_t s = 0; s < S.nsn; ++s) {
            const uint32_t nr = S.sn_nr[s];
            for (uint32_t i = 0; i < nr; ++i) {
                gptr[at++] = (uint32_t)gsrc.size();
                for (uint32_t q = S.child_ptr[s]; q < S.child_ptr[s + 1]; ++q) {
                    const uint32_t ch = S.child_list[q];
                    const int32_t ii = S.inv[S.inv_off[ch] + i];
                    if (ii >= 0) gsrc.push_back((uint32_t)(S.soff[ch] + S.sn_nc[ch] + ii));
                }
            }
            gptr[at++] = (uint32_t)gsrc.size();
        }
        sanm_check(gsrc.size() < (1ull << 32) && S.soff[S.nsn] < (1ull << 32));
        if (gsrc.empty()) gsrc.push_back(0);
        T.gptr.upload(gptr);
        T.gsrc.upload(gsrc);
    }

    std::vector<FrontD> fd(S.nsn);
    uint32_t max_front = 0;
    for (uint32_t s = 0; s < S.nsn; ++s) {
        FrontD& f = fd[s];
        f.F = I.fronts.ptr() + S.foff[s];
        f.vec = I.vecs.ptr() + S.soff[s];
        f.st = T.structs.ptr() + S.soff[s];
        f.inv = S.sn_parent[s] >= 0 ? T.inv.ptr() + S.inv_off[s] : nullptr;
        f.rowperm = I.rowperm.ptr() + S.sn_c0[s];
        f.lperm = I.lperm.ptr() + (size_t)s * LPERM_PER_FRONT;
        f.gptr = T.gptr.ptr() + S.soff[s] + s;
        f.nr = S.sn_nr[s];
        f.nc = S.sn_nc[s];
        f.c0 = S.sn_c0[s];
        f.child_begin = S.child_ptr[s];
        f.child_end = S.child_ptr[s + 1];
        f.parent = S.sn_parent[s];
        max_front = std::max(max_front, f.nr);
    }
    sanm_check(max_front < 65536,
               "front of %u rows: the assembly kernel indexes a front with 32 bits",
               max_front);
    I.fd.upload(fd);
    lap("gather lists + front descriptors");

    const uint32_t nlev = (uint32_t)S.level_ptr.size() - 1;
    if (fresh) {
    T.levels.assign(nlev, {});
    std::vector<Task> asm_t, fwd_t, bwd_t, prep_t;
    std::vector<uint32_t> small_l, big_l, wide_l;
    for (uint32_t l = 0; l < nlev; ++l) {
        auto& L = T.levels[l];
        L.asm_off = (uint32_t)asm_t.size();
        L.small_off = (uint32_t)small_l.size();
        L.big_off = (uint32_t)big_l.size();
        L.wide_off = (uint32_t)wide_l.size();
        L.prep_off = (uint32_t)prep_t.size();
        L.fwd_off = (uint32_t)fwd_t.size();
        L.bwd_off = (uint32_t)bwd_t.size();
        for (uint32_t q = S.level_ptr[l]; q < S.level_ptr[l + 1]; ++q) {
            const uint32_t s = S.level_list[q];
            const uint32_t nr = S.sn_nr[s], nc = S.sn_nc[s], nu = nr - nc;
            if (S.child_ptr[s + 1] > S.child_ptr[s]) {
                uint64_t tot = (uint64_t)nr * nr;
                for (uint64_t t = 0; t * ASM_TILE < tot; ++t) {
                    asm_t.push_back({s, (uint32_t)t});
                }
            }
            if (nr <= small_nr()) {
                small_l.push_back(s);
                L.small_max_nr = std::max(L.small_max_nr, nr);
            } else if (nr >= wide_min) {
                wide_l.push_back(s);
                L.wide_max_nc = std::max(L.wide_max_nc, nc);
                L.wide_max_nr = std::max(L.wide_max_nr, nr);
            } else {
                big_l.push_back(s);
                L.big_max_nc = std::max(L.big_max_nc, nc);
                L.big_max_nr = std::max(L.big_max_nr, nr);
            }
            L.max_nc = std::max(L.max_nc, nc);
            L.max_nr = std::max(L.max_nr, nr);
        }
        for (int pass = 0; pass < 2; ++pass) {
            for (uint32_t q = S.level_ptr[l]; q < S.level_ptr[l + 1]; ++q) {
                const uint32_t s = S.level_list[q];
                const uint32_t nr = S.sn_nr[s], nc = S.sn_nc[s], nu = nr - nc;
                if ((nc < SOLVE_WIDE_COLS) == (pass == 0)) {
                    uint32_t nf = std::max<uint32_t>(1, (nu + FWD_ROWS - 1) / FWD_ROWS);
                    for (uint32_t t = 0; t < nf; ++t) fwd_t.push_back({s, t});
                    if (pass == 0) L.fwd_narrow_nc = std::max(L.fwd_narrow_nc, nc);
                }
                if ((nr < SOLVE_WIDE_COLS) == (pass == 0)) {
                    uint32_t nbk = (nc + BWD_ROWS - 1) / BWD_ROWS;
                    for (uint32_t t = 0; t < nbk; ++t) bwd_t.push_back({s, t});
                    if (pass == 0) L.bwd_narrow_nr = std::max(L.bwd_narrow_nr, nr);
                }
            }
            if (pass == 0) {
                L.fwd_split = (uint32_t)fwd_t.size() - L.fwd_off;
                L.bwd_split = (uint32_t)bwd_t.size() - L.bwd_off;
            }
        }
        L.asm_n = (uint32_t)asm_t.size() - L.asm_off;
        L.small_n = (uint32_t)small_l.size() - L.small_off;
        L.big_n = (uint32_t)big_l.size() - L.big_off;
        L.wide_n = (uint32_t)wide_l.size() - L.wide_off;
        L.solve_global = L.max_nr > smem_nr;
        if (L.solve_global) {
            for (uint32_t q = S.level_ptr[l]; q < S.level_ptr[l + 1]; ++q) {
                const uint32_t s = S.level_list[q];
                for (uint32_t t = 0; t * 256 < S.sn_nr[s]; ++t) prep_t.push_back({s, t});
            }
        }
        L.prep_n = (uint32_t)prep_t.size() - L.prep_off;
        L.fwd_n = (uint32_t)fwd_t.size() - L.fwd_off;
        L.bwd_n = (uint32_t)bwd_t.size() - L.bwd_off;
    }
    {
        std::vector<LevelD> ld(nlev);
        size_t sm = 0;
        for (uint32_t l = 0; l < nlev; ++l) {
            const auto& L = T.levels[l];
            const bool fw = L.max_nc >= SOLVE_WIDE_COLS, bw = L.max_nr >= SOLVE_WIDE_COLS;
            ld[l] = {L.fwd_off, L.fwd_n, L.bwd_off, L.bwd_n, fw, bw};
            if (L.solve_global) continue;  // (the persistent kernel is not used then)
            sm = std::max(sm, 2 * (size_t)L.max_nc +
                                      (fw ? solve_scratch_doubles<FWD_ROWS, FWD_GROUPS_WIDE>()
                                          : solve_scratch_doubles<FWD_ROWS, FWD_GROUPS_NARROW>()));
            sm = std::max(sm, (size_t)L.max_nr +
                                      (bw ? solve_scratch_doubles<BWD_ROWS, BWD_GROUPS_WIDE>()
                                          : solve_scratch_doubles<BWD_ROWS, BWD_GROUPS_NARROW>()));
        }
        T.level_d.upload(ld);
        T.solve_smem = sm * sizeof(double);
    }
    T.asm_tasks.upload(asm_t);
    T.fwd_tasks.upload(fwd_t);
    T.bwd_tasks.upload(bwd_t);
    T.small_list.upload(small_l);
    T.big_list.upload(big_l);
    if (wide_l.empty()) wide_l.push_back(0);
    T.wide_list.upload(wide_l);
    if (prep_t.empty()) prep_t.push_back({0, 0});
    T.prep_tasks.upload(prep_t);
    }  // fresh
    lap("level task lists");

    if (getenv("SANM_LU_DEBUG")) {
        printf("[lu] n=%zu nnz=%zu nsn=%u levels=%u factor_entries=%zu "
               "front_total=%zu flops=%.3g max_front=%u\n",
               (size_t)n, (size_t)m_nnz, S.nsn, nlev, (size_t)S.factor_entries,
               (size_t)S.front_total, S.flops, max_front);
        for (uint32_t l = 0; l < nlev; ++l) {
            uint32_t cnt = S.level_ptr[l + 1] - S.level_ptr[l], mnc = 0, mnr = 0;
            uint64_t ent = 0;
            for (uint32_t q = S.level_ptr[l]; q < S.level_ptr[l + 1]; ++q) {
                uint32_t s = S.level_list[q];
                mnc = std::max(mnc, S.sn_nc[s]);
                mnr = std::max(mnr, S.sn_nr[s]);
                ent += (uint64_t)S.sn_nc[s] * (2 * S.sn_nr[s] - S.sn_nc[s]);
            }
            printf("[lu]  level %2u: fronts=%5u max_nc=%4u max_nr=%4u "
                   "entries=%9zu fwd_tasks=%u bwd_tasks=%u\n",
                   l, cnt, mnc, mnr, (size_t)ent, T.levels[l].fwd_n,
                   T.levels[l].bwd_n);
        }
        fflush(stdout);
    }
    m_stats = {};
    m_stats.n = n;
    m_stats.nnz_a = m_nnz;
    m_stats.factor_entries = S.factor_entries;
    m_stats.front_entries = S.front_total;
    m_stats.flops = S.flops;
    m_stats.nr_supernodes = S.nsn;
    m_stats.nr_levels = nlev;
    m_stats.max_front = max_front;
    m_analyzed = true;
    I.factored = false;
}

void SparseLU::factor(const double* vals_dev) {
    sanm_check(m_analyzed, "factor() before analyze()");
    Impl& I = *m_impl;
    LuTopo& T = *I.topo;
    if (vals_dev != I.vals.ptr()) {
        rt::d2d(I.vals.ptr(), vals_dev, m_nnz * sizeof(double));
    }
    I.fronts.zero();
    I.counters.zero();
    // max |a_ij| (scale of the static pivot threshold): two stages
    rt::for_each_block<256>(ABSMAX_BLOCKS, 256, 256 * sizeof(double),
                       AbsMaxKernel{I.vals.ptr(), m_nnz, ABSMAX_BLOCKS,
                                    I.scal.ptr() + 4});
    rt::for_each_block<256>(1, 256, 256 * sizeof(double),
                       AbsMaxKernel{I.scal.ptr() + 4, ABSMAX_BLOCKS, 1,
                                    I.scal.ptr()});
    rt::for_each(m_nnz, ScatterA{I.fronts.ptr(), T.a_dest.ptr(), I.vals.ptr()});
    const double* amax = I.scal.ptr();
    int* npert = I.counters.ptr();
    for (auto& L : T.levels) {
        if (L.asm_n) {
            LuScope ps{"lu_asm"};
            rt::for_each_block<256>(L.asm_n, 256,
                               AssembleChildren::MAX_CH * sizeof(AssembleChildren::ChildD),
                               AssembleChildren{I.fd.ptr(), T.child_list.ptr(),
                                                T.asm_tasks.ptr() + L.asm_off});
        }
        if (L.small_n) {
            size_t smem = (size_t)L.small_max_nr * L.small_max_nr * sizeof(double) + 16;
            LuScope ps{"lu_small"};
            rt::for_each_block<256>(L.small_n, 256, smem,
                               SmallFrontKernel{I.fd.ptr(),
                                                T.small_list.ptr() + L.small_off,
                                                amax, npert});
        }
        if (L.wide_n) {
            // superblocks of WB pivots: inverse of the pivot block, row panel,
            // rank-WB trailing update (DMMA), column panel (sparse_lu_wide.h)
            const uint32_t* list = T.wide_list.ptr() + L.wide_off;
            const uint32_t tiles = (L.wide_max_nr + WT - 1) / WT;
            // Look-ahead (SANM_LU_LOOKAHEAD=1, persistent trailing update only): the
            // next pivot block is updated first and inverted on a side stream
            // beside the trailing update, which leaves that front's CTA an SM.
            // Opt-in: measured 1.322 -> 1.301 s per factorisation of the 1.08 M-tet
            // cuboid (the single-CTA inverse is hidden, but the trailing update
            // runs on fewer SMs and the pivot tile costs a launch of its own) --
            // not enough for a second stream in the default path.  Default: the
            // plain order diag -> row -> gemm -> col.
#ifndef SANM_HOSTEMU
            const bool lookahead =
                    wide_gemm_cfg() == 4 && getenv("SANM_LU_LOOKAHEAD") != nullptr;
#else
            const bool lookahead = false;
#endif
            void* diag_done = nullptr;  // fence: inverse of the pivot block of step kb
            for (uint32_t kb = 0; kb < L.wide_max_nc; kb += WB) {
                if (diag_done) {
                    rt::main_wait(diag_done);
                    diag_done = nullptr;
                } else {
                    LuScope ps{"lu_wide_diag"};
                    rt::for_each_block<512>(L.wide_n, 512, W_DIAG_SMEM,
                                            WideDiagKernel{I.fd.ptr(), list, kb, amax, npert});
                }
                const bool ahead = lookahead && kb + WB < L.wide_max_nc;
                {
                    LuScope ps{"lu_wide_row"};
                    rt::for_each_block<WidePanelCfg::THREADS, WidePanelCfg::MINB>(
                            (size_t)L.wide_n * tiles, WidePanelCfg::THREADS,
                            WidePanelCfg::SMEM,
                            WidePanelKernel<true>{I.fd.ptr(), list, kb, tiles});
                }
                {
                    if (ahead) {
                        {
                            LuScope ps{"lu_wide_pivot_tile"};
                            rt::for_each_block<WideGemmBig::THREADS, 1>(
                                    L.wide_n, WideGemmBig::THREADS, WideGemmBig::SMEM,
                                    WidePivotTileKernel{I.fd.ptr(), list, kb});
                        }
                        rt::pipe_wait(rt::PIPE_AUX, rt::fence_main());
                        {
                            rt::ScopedPipe side{rt::PIPE_AUX};
                            rt::for_each_block<512>(
                                    L.wide_n, 512, W_DIAG_SMEM,
                                    WideDiagKernel{I.fd.ptr(), list, kb + WB, amax, npert});
                        }
                        diag_done = rt::fence_pipe(rt::PIPE_AUX);
                    }
                    LuScope ps{"lu_wide_gemm"};
                    const int fast = getenv("SANM_LU_WIDE_SLOW") == nullptr;
                    auto gemm = [&](auto cfg) {
                        using G = decltype(cfg);
                        const uint32_t tiles_m = (L.wide_max_nr + G::TM - 1) / G::TM;
                        const uint32_t tiles_n = (L.wide_max_nr + G::TN - 1) / G::TN;
                        rt::for_each_block<G::THREADS, G::MINB>(
                                (size_t)L.wide_n * tiles_m * tiles_n, G::THREADS, G::SMEM,
                                WideGemmKernelT<G>{I.fd.ptr(), list, kb, tiles_m, tiles_n,
                                                   fast});
                    };
                    switch (wide_gemm_cfg()) {
                        case 1:
                            gemm(WideGemmSplit{});
                            break;
                        case 2:
                            gemm(WideGemmHalfN{});
                            break;
                        case 3:
                            gemm(WideGemmBig3{});
                            break;
#ifndef SANM_HOSTEMU
                        case 4: {
                            const uint64_t ntile = (uint64_t)L.wide_n * tiles * tiles;
                            // with the look-ahead, SMs are left to the pivot-block
                            // CTAs that run beside this kernel
                            uint64_t nsm = (uint64_t)rt::sm_count();
                            if (ahead) nsm -= std::min<uint64_t>(L.wide_n, 8);
                            rt::for_each_block<WidePersist::THREADS, 1>(
                                    (size_t)(ntile < nsm ? ntile : nsm), WidePersist::THREADS,
                                    WidePersist::SMEM,
                                    WideGemmPersistKernel{I.fd.ptr(), list, kb, tiles,
                                                          L.wide_n, ahead ? 1 : 0});
                            break;
                        }
#endif
                        default:
                            gemm(WideGemmBig{});
                            break;
                    }
                }
                {
                    LuScope ps{"lu_wide_col"};
                    rt::for_each_block<WidePanelCfg::THREADS, WidePanelCfg::MINB>(
                            (size_t)L.wide_n * tiles, WidePanelCfg::THREADS,
                            WidePanelCfg::SMEM,
                            WidePanelKernel<false>{I.fd.ptr(), list, kb, tiles});
                }
            }
        }
        if (L.big_n) {
            const uint32_t* list = T.big_list.ptr() + L.big_off;
            const uint32_t tiles = (L.big_max_nr + 63) / 64;
            const size_t diag_smem =
                    (NB * NB + 2 * NB) * sizeof(double) + NB * sizeof(uint32_t) + 16;
            const size_t panel_smem = (NB * NB + NB * 64) * sizeof(double);
            {
                LuScope ps{"lu_diag"};
                // 4 warps per front (gj32.cuh)
                rt::for_each_block<128>(L.big_n, 128, diag_smem,
                                        BigDiagKernel{I.fd.ptr(), list, 0, amax, npert});
            }
            for (uint32_t kb = 0; kb < L.big_max_nc; kb += NB) {
                {
                    LuScope ps{"lu_gemm"};
                    // two CTAs per SM (128 registers): a front of 700 rows is
                    // 121 tiles, a level several of them
                    rt::for_each_block<256, 2>((size_t)L.big_n * tiles * tiles, 256,
                                       (3 * NB * 64 + NB * NB) * sizeof(double),
                                       BigGemmKernel{I.fd.ptr(), list, kb, tiles});
                }
                {
                    // both panels of this step beside the pivot block of the next
                    LuScope ps{"lu_panels_diag"};
                    const uint32_t ndiag = kb + NB < L.big_max_nc ? L.big_n : 0;
                    const uint32_t npanel = L.big_n * tiles;
                    rt::for_each_block<256>(
                            (size_t)ndiag + 2 * (size_t)npanel, 256,
                            std::max(diag_smem, panel_smem),
                            BigPanelsDiagKernel{
                                    BigColPanelKernel{I.fd.ptr(), list, kb, tiles},
                                    BigRowPanelKernel{I.fd.ptr(), list, kb, tiles},
                                    BigDiagKernel{I.fd.ptr(), list, kb + NB, amax, npert},
                                    ndiag, npanel});
                }
            }
        }
    }
    I.factored = true;
    I.counters_read = false;
    m_stats.nr_perturbed = 0;
    m_stats.nr_refine_steps = 0;
}

namespace {
//! r = b - y
struct ResidualKernel {
    const double* b;
    const double* y;
    double* r;
    SANM_D void operator()(size_t i) const { r[i] = b[i] - y[i]; }
};
struct AxpyKernel {
    double* x;
    const double* dx;
    SANM_D void operator()(size_t i) const { x[i] += dx[i]; }
};
}  // namespace

void SparseLU::solve(double* x_dev, const double* b_dev) {
    Impl& I = *m_impl;
    sanm_check(I.factored, "solve() before factor()");
#ifndef SANM_HOSTEMU
    static const bool solve_bench = getenv("SANM_LU_SOLVE_BENCH") != nullptr;  // diagnostic
    if (solve_bench && !I.bench_done) {
        // diagnostic: is a solve bound by the host's launch rate or by the
        // device?  20 solves enqueued back to back (host time to enqueue vs
        // device time), then the same chain replayed from a CUDA graph
        I.bench_done = true;
        cudaStream_t st = rt::stream();
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        constexpr int REP = 20;
        auto run = [&](const char* what, auto&& body) {
            rt::sync();
            auto t0 = std::chrono::steady_clock::now();
            cudaEventRecord(e0, st);
            for (int i = 0; i < REP; ++i) body();
            cudaEventRecord(e1, st);
            auto t1 = std::chrono::steady_clock::now();
            rt::sync();
            float ms = 0;
            cudaEventElapsedTime(&ms, e0, e1);
            fprintf(stderr, "[lu-bench] %-8s n=%llu levels=%zu: host enqueue %.1f us, device %.1f us per solve\n",
                    what, (unsigned long long)m_n, I.topo->levels.size(),
                    std::chrono::duration<double, std::micro>(t1 - t0).count() / REP,
                    ms * 1e3 / REP);
        };
        run("launches", [&] { solve_once(x_dev, b_dev); });
        run("launches", [&] { solve_once(x_dev, b_dev); });
        cudaGraph_t g = nullptr;
        cudaGraphExec_t ge = nullptr;
        SANM_CUDA_CHECK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        solve_once(x_dev, b_dev);
        SANM_CUDA_CHECK(cudaStreamEndCapture(st, &g));
        SANM_CUDA_CHECK(cudaGraphInstantiate(&ge, g, 0));
        run("graph", [&] { cudaGraphLaunch(ge, st); });
        run("graph", [&] { cudaGraphLaunch(ge, st); });
        cudaGraphExecDestroy(ge);
        cudaGraphDestroy(g);
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
    }
#endif
    if (!I.counters_read) {
        // one 4-byte download per factorisation (the stream is drained here
        // once; the ANM driver synchronises after its first solve anyway)
        int cnt = 0;
        rt::d2h(&cnt, I.counters.ptr(), sizeof(int));
        m_stats.nr_perturbed = (uint64_t)cnt;
        I.counters_read = true;
    }
    if (!m_stats.nr_perturbed) {
        solve_once(x_dev, b_dev);
        return;
    }
    // perturbed pivots: the factors are those of a nearby matrix; refine
    // against the true one (same rule as PARDISO's default iparm[7])
    I.ref_b.alloc(m_n);
    I.ref_r.alloc(m_n);
    I.ref_dx.alloc(m_n);
    rt::d2d(I.ref_b.ptr(), b_dev, m_n * sizeof(double));  // b may alias x
    solve_once(x_dev, I.ref_b.ptr());
    for (int it = 0; it < 2; ++it) {
        apply(I.ref_r.ptr(), x_dev);
        rt::for_each(m_n, ResidualKernel{I.ref_b.ptr(), I.ref_r.ptr(), I.ref_r.ptr()});
        solve_once(I.ref_dx.ptr(), I.ref_r.ptr());
        rt::for_each(m_n, AxpyKernel{x_dev, I.ref_dx.ptr()});
        ++m_stats.nr_refine_steps;
    }
}

void SparseLU::solve_once(double* x_dev, const double* b_dev) {
    Impl& I = *m_impl;
    LuTopo& T = *I.topo;
    // b may alias x: the forward sweep has consumed b before the backward
    // sweep writes x
    // One launch per level is the default: with programmatic dependent launch
    // (rt.h) the chain costs ~3.5 us per level and beats the persistent
    // grid-synchronised kernel at every size measured (cuboid n = 9 k: 0.110 vs
    // 0.144 ms).  SANM_LU_PHASED=1 selects the persistent kernel.
    bool phased = getenv("SANM_LU_PHASED") != nullptr;  // per call: tests toggle it
    for (auto& L : T.levels) phased = phased && !L.solve_global;
    if (phased && !g_lu_prof) {
        const uint32_t nlev = (uint32_t)T.levels.size();
        rt::for_each_block_phased<512>(
                512, T.solve_smem, 2 * nlev,
                SolvePhased{I.fd.ptr(), T.gsrc.ptr(), I.vecs.ptr(),
                            T.fwd_tasks.ptr(), T.bwd_tasks.ptr(), T.level_d.ptr(),
                            nlev, b_dev, T.perm.ptr(), I.xp.ptr(), x_dev});
        return;
    }
    char tag[32];
    static const bool no_hoist = getenv("SANM_LU_NO_HOIST") != nullptr;
    // narrow tasks of a mixed level get their own launch from this many on
    // (SANM_LU_SOLVE_SPLIT; 0 = never)
    const uint32_t split_min = [] {  // per call: tests switch it
        const char* e = getenv("SANM_LU_SOLVE_SPLIT");
        const int v = e ? atoi(e) : 4096;
        return v > 0 ? (uint32_t)v : 0xFFFFFFFFu;
    }();
    for (auto& L : T.levels) {
        snprintf(tag, sizeof(tag), "lu_fwd_%02d", (int)(&L - T.levels.data()));
        LuScope ps{tag};
        constexpr int R = FWD_ROWS;
        // pre-wait prologues (ForwardKernel): not in the first launch of the
        // sweep, whose predecessor may have produced b
        const int hoist = no_hoist ? 0 : (&L != T.levels.data());
        if (L.solve_global) {
            constexpr int G = FWD_GROUPS_WIDE;
            rt::for_each_block<256>(
                    L.prep_n, 256, 0,
                    SolvePrepKernel<true>{I.fd.ptr(), T.gsrc.ptr(), I.vecs.ptr(),
                                          T.prep_tasks.ptr() + L.prep_off, b_dev,
                                          T.perm.ptr(), I.xp.ptr()});
            rt::for_each_block<R * G>(
                    L.fwd_n, R * G, solve_scratch_doubles<R, G>() * sizeof(double),
                    ForwardKernel<R, G, true>{I.fd.ptr(), T.gsrc.ptr(), I.vecs.ptr(),
                                              T.fwd_tasks.ptr() + L.fwd_off, b_dev,
                                              T.perm.ptr(), 0});
        } else if (L.max_nc >= SOLVE_WIDE_COLS) {
            // a large level that mixes wide and narrow fronts (the bottom of a
            // big problem): the narrow ones in narrow blocks, as their own launch
            uint32_t skip = 0;
            if (L.fwd_split >= split_min && L.fwd_split < L.fwd_n) {
                constexpr int G = FWD_GROUPS_NARROW;
                skip = L.fwd_split;
                rt::for_each_block<R * G>(
                        skip, R * G,
                        (2 * (size_t)L.fwd_narrow_nc + solve_scratch_doubles<R, G>()) *
                                sizeof(double),
                        ForwardKernel<R, G>{I.fd.ptr(), T.gsrc.ptr(), I.vecs.ptr(),
                                            T.fwd_tasks.ptr() + L.fwd_off, b_dev,
                                            T.perm.ptr(), hoist});
            }
            constexpr int G = FWD_GROUPS_WIDE;
            rt::for_each_block<R * G>(
                    L.fwd_n - skip, R * G,
                    (2 * (size_t)L.max_nc + solve_scratch_doubles<R, G>()) * sizeof(double),
                    ForwardKernel<R, G>{I.fd.ptr(), T.gsrc.ptr(), I.vecs.ptr(),
                                     T.fwd_tasks.ptr() + L.fwd_off + skip, b_dev,
                                     T.perm.ptr(), hoist});
        } else {
            constexpr int G = FWD_GROUPS_NARROW;
            rt::for_each_block<R * G>(
                    L.fwd_n, R * G,
                    (2 * (size_t)L.max_nc + solve_scratch_doubles<R, G>()) * sizeof(double),
                    ForwardKernel<R, G>{I.fd.ptr(), T.gsrc.ptr(), I.vecs.ptr(),
                                     T.fwd_tasks.ptr() + L.fwd_off, b_dev,
                                     T.perm.ptr(), hoist});
        }
    }
    for (size_t l = T.levels.size(); l-- > 0;) {
        auto& L = T.levels[l];
        snprintf(tag, sizeof(tag), "lu_bwd_%02d", (int)l);
        LuScope ps{tag};
        constexpr int R = BWD_ROWS;
        // not in the first launch: the root's w1' comes from the launch before it
        const int hoist = no_hoist ? 0 : (l + 1 != T.levels.size());
        if (L.solve_global) {
            constexpr int G = BWD_GROUPS_WIDE;
            rt::for_each_block<256>(
                    L.prep_n, 256, 0,
                    SolvePrepKernel<false>{I.fd.ptr(), T.gsrc.ptr(), I.vecs.ptr(),
                                           T.prep_tasks.ptr() + L.prep_off, b_dev,
                                           T.perm.ptr(), I.xp.ptr()});
            rt::for_each_block<R * G>(
                    L.bwd_n, R * G, solve_scratch_doubles<R, G>() * sizeof(double),
                    BackwardKernel<R, G, true>{I.fd.ptr(), T.bwd_tasks.ptr() + L.bwd_off,
                                               I.xp.ptr(), x_dev, T.perm.ptr(), 0});
        } else if (L.max_nr >= SOLVE_WIDE_COLS) {
            uint32_t skip = 0;
            if (L.bwd_split >= split_min && L.bwd_split < L.bwd_n) {
                constexpr int G = BWD_GROUPS_NARROW;
                skip = L.bwd_split;
                rt::for_each_block<R * G>(
                        skip, R * G,
                        ((size_t)L.bwd_narrow_nr + solve_scratch_doubles<R, G>()) *
                                sizeof(double),
                        BackwardKernel<R, G>{I.fd.ptr(), T.bwd_tasks.ptr() + L.bwd_off,
                                             I.xp.ptr(), x_dev, T.perm.ptr(), hoist});
            }
            constexpr int G = BWD_GROUPS_WIDE;
            rt::for_each_block<R * G>(
                    L.bwd_n - skip, R * G,
                    ((size_t)L.max_nr + solve_scratch_doubles<R, G>()) * sizeof(double),
                    BackwardKernel<R, G>{I.fd.ptr(), T.bwd_tasks.ptr() + L.bwd_off + skip,
                                      I.xp.ptr(), x_dev, T.perm.ptr(), hoist});
        } else {
            constexpr int G = BWD_GROUPS_NARROW;
            rt::for_each_block<R * G>(
                    L.bwd_n, R * G,
                    ((size_t)L.max_nr + solve_scratch_doubles<R, G>()) * sizeof(double),
                    BackwardKernel<R, G>{I.fd.ptr(), T.bwd_tasks.ptr() + L.bwd_off,
                                      I.xp.ptr(), x_dev, T.perm.ptr(), hoist});
        }
    }
}

void SparseLU::apply(double* y_dev, const double* x_dev) const {
    const Impl& I = *m_impl;
    const LuTopo& T = *I.topo;
    rt::for_each(m_n, SpmvKernel{T.rowptr.ptr(), T.colidx.ptr(), I.vals.ptr(),
                                 x_dev, y_dev});
}

void SparseLU::apply_transposed(double*, const double*) const {
    sanm_throw(SANM_ERR_GENERIC, "apply_transposed is not implemented");
}

double SparseLU::coeff_l2() const {
    std::vector<double> v = m_impl->vals.to_host();
    double s = 0;
    for (double x : v) s += x * x;
    return std::sqrt(s);
}

}  // namespace sanm_b200
