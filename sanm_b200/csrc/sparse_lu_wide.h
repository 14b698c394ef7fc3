/* sanm_b200/csrc/sparse_lu_wide.h -- the wide-front path of the numeric
 * factorisation (included by sparse_lu.cu inside its anonymous namespace).
 *
 * Fronts of a few thousand rows and more (the top separators of 3-D meshes:
 * 10-20 k rows for a 1 M-tet cube) are reduced WB = 128 pivots at a time.  The
 * block Gauss-Jordan sweep composes: sweeping a set P of pivots at once is
 *
 *     D   = (P_r F[P,P])^-1                 WideDiagKernel, one CTA, shared memory
 *     R'  = D  (P_r F[P,Q])                 WidePanelKernel<ROW>   (128 x nr GEMM)
 *     F[Q,Q] -= F[Q,P] R'                   WideGemmKernel         (rank-128 update)
 *     C'  = -F[Q,P] D                       WidePanelKernel<COL>
 *
 * with Q = all other rows/columns of the front (already swept pivots
 * included, as in any Gauss-Jordan step) and P_r the row permutation of the
 * partial pivoting inside the 128 x 128 block.  The trailing update is a real
 * GEMM (K = 128: 2 nr^2 128 flops for 16 nr^2 bytes of front traffic, i.e.
 * compute-bound) and runs on the FP64 tensor path: mma.sync.m8n8k4.f64 (DMMA),
 * operands staged in shared memory by cp.async, 128 x 128 tile per CTA, 64 x 32
 * per warp.  tcgen05 has no FP64 kind on sm_100a; DMMA and the FMA pipe have
 * the same peak here (37.1 vs 36.7 TFLOP/s, tools/micro/fp64_probe.cu), but a
 * DMMA needs 1 A and 1 B double per thread for 256 multiply-adds, so the
 * kernel is not bound by shared-memory or register bandwidth as the rank-32
 * FMA kernels of the narrow path are.
 *
 * Replaces PARDISO's numeric phase for these fronts
 * (libsanm/sparse_solver.cpp:71-127, phase 12).
 */
#pragma once

#ifndef SANM_WIDE_GEMM_DEFAULT
#define SANM_WIDE_GEMM_DEFAULT 4
#endif
constexpr int WB = 128;  //!< pivots per superblock
constexpr int WT = 128;  //!< tile edge of the panel / trailing GEMMs

/*! Shape of the tile kernel: WM x WN warps, each MT x NT DMMA tiles of 8 x 8
 * (tile = TM x TN with TM = WM MT 8 rows and TN = WN NT 8 columns), K staged WK
 * deep in STAGES buffers.  Shared-memory strides TM + 4 / WK + 4 (= 4 mod 16
 * doubles) make the fragment loads of a half-warp hit 16 distinct 8-byte
 * banks. */
template <int WM_, int WN_, int MT_, int NT_, int WK_, int STAGES_, int MINB_>
struct WideCfgT {
    static constexpr int WM = WM_, WN = WN_, MT = MT_, NT = NT_, WK = WK_,
                         STAGES = STAGES_, MINB = MINB_;
    static constexpr int TM = WM * MT * 8, TN = WN * NT * 8;
    static constexpr int THREADS = 32 * WM * WN;
    static constexpr int LDA = TM + 4, LDB = WK + 4;
    static constexpr int STAGE_DOUBLES = WK * LDA + TN * LDB;
#ifdef SANM_HOSTEMU
    // the emulated tile keeps its product in the block's scratch
    static constexpr size_t SMEM = (size_t)WT * WT * sizeof(double);
#else
    static constexpr size_t SMEM = (size_t)STAGES * STAGE_DOUBLES * sizeof(double);
#endif
    static_assert(TM <= WT && TN <= WT && WK % 4 == 0 && (LDB % 16) == 4 &&
                          (LDA % 16) == 4,
                  "tile shape");
};
//! panels (128 x 128 outputs, K = 128): 8 warps of 64 x 32, one CTA per SM
using WidePanelCfg = WideCfgT<2, 4, 8, 4, 32, 2, 1>;
/*! trailing update, selectable at run time (SANM_LU_WIDE_CFG).  Measured on the
 * 1.08 M-tet cuboid, trailing updates per factorisation, all with the fast
 * interior tiles below (profiles/r2_x_widegemm_fast_ncu.txt):
 *   persist (default) 128 x 128, one persistent CTA per SM, next tile's C
 *           and first stage prefetched under the DMMAs (WideGemmPersistKernel) 1.032 s
 *   big     128 x 128, 8 warps of 64 x 32, one CTA per SM, two 32-deep stages   1.134 s
 *   big3    the same with three stages (a larger prologue per tile)            1.385 s
 *   split   64 x 128, 8 warps of 32 x 32, two CTAs per SM                       1.258 s
 *   halfn   128 x 64, 4 warps of 64 x 32, two CTAs per SM                       1.303 s
 * Two CTAs per SM overlap one tile's loads and stores with the other's DMMAs,
 * but move 25 % more bytes per flop through L2 (0.152 vs 0.122 B/flop) and
 * lose more than the overlap gives. */
using WideGemmBig = WideCfgT<2, 4, 8, 4, 32, 2, 1>;
using WideGemmSplit = WideCfgT<2, 4, 4, 4, 32, 2, 2>;
//! Big with three stages (212 KB): a whole chunk of slack for the copies
using WideGemmBig3 = WideCfgT<2, 4, 8, 4, 32, 3, 1>;
//! HalfN: 128 x 64 tiles, FOUR warps of 64 x 32 (the warp tile and fragment-load
//! ratio of Big), two 128-thread CTAs per SM that run out of phase
using WideGemmHalfN = WideCfgT<2, 2, 8, 4, 32, 2, 2>;
//! which of the three runs: SANM_LU_WIDE_CFG = big | split | halfn
inline int wide_gemm_cfg() {
    // read per call: tests switch it inside one process
    const char* e = getenv("SANM_LU_WIDE_CFG");
    if (e && !strcmp(e, "big")) return 0;
    if (e && !strcmp(e, "split")) return 1;
    if (e && !strcmp(e, "halfn")) return 2;
    if (e && !strcmp(e, "big3")) return 3;
    if (e && !strcmp(e, "persist")) return 4;
    return SANM_WIDE_GEMM_DEFAULT;
}

enum WideMode { W_SUB = 0, W_SET = 1, W_SETNEG = 2 };

/*! One CTA: C(m x n) {-=, =, = -} A(m x k) B(k x n), all column-major.
 *   A(i, q) = Ap[q * lda + i]
 *   B(q, j) = Bp[j * ldb + (kperm ? kperm[q] : q)]
 *   C(i, j) = Cp[j * ldc + i]
 * m <= G::TM, n <= G::TN, k <= WB.  Entries with local row in [rs0, rs1) or local column
 * in [cs0, cs1) are not stored (the pivot rows / columns of the step).  C may
 * alias A or B (the panels are updated in place): every operand element is in
 * shared memory before the first store. */
template <class G, int MODE>
SANM_D inline void wide_gemm_tile(const rt::BlockCtx& c, const double* Ap,
                                  uint64_t lda, const double* Bp, uint64_t ldb,
                                  const uint32_t* kperm, double* Cp, uint64_t ldc,
                                  uint32_t m, uint32_t n, uint32_t k, uint32_t rs0,
                                  uint32_t rs1, uint32_t cs0, uint32_t cs1) {
#ifdef SANM_HOSTEMU
    // same arithmetic, plain loops (one emulated thread per block)
    double* acc = reinterpret_cast<double*>(c.smem);  // [n][m]
    for (uint32_t j = 0; j < n; ++j) {
        for (uint32_t i = 0; i < m; ++i) {
            double s = 0;
            for (uint32_t q = 0; q < k; ++q) {
                s += Ap[(uint64_t)q * lda + i] *
                     Bp[(uint64_t)j * ldb + (kperm ? kperm[q] : q)];
            }
            acc[(size_t)j * WT + i] = s;
        }
    }
    for (uint32_t j = 0; j < n; ++j) {
        if (j >= cs0 && j < cs1) continue;
        for (uint32_t i = 0; i < m; ++i) {
            if (i >= rs0 && i < rs1) continue;
            double* p = Cp + (uint64_t)j * ldc + i;
            const double s = acc[(size_t)j * WT + i];
            if (MODE == W_SUB) {
                *p -= s;
            } else if (MODE == W_SET) {
                *p = s;
            } else {
                *p = -s;
            }
        }
    }
#else
    constexpr int WK = G::WK, MT = G::MT, NT = G::NT, LDA = G::LDA, LDB = G::LDB;
    double* sm = reinterpret_cast<double*>(c.smem);
    const int tid = c.tid, lane = tid & 31, warp = tid >> 5;
    const int wm = warp % G::WM, wn = warp / G::WM;
    const int g = lane >> 2, t4 = lane & 3;
    double acc[MT][NT][2];

    auto stage_ptr = [&](int s) { return sm + (size_t)s * G::STAGE_DOUBLES; };
    auto load_stage = [&](int s, uint32_t k0) {
        double* As = stage_ptr(s);
        double* Bs = As + WK * LDA;
        const uint32_t kk = k - k0 < (uint32_t)WK ? k - k0 : (uint32_t)WK;
        // A chunk: WK columns (q) of TM rows; consecutive threads -> rows
        for (int e = tid; e < WK * G::TM; e += G::THREADS) {
            const uint32_t q = e / G::TM, i = e % G::TM;
            const bool ok = q < kk && i < m;
            const double* src = ok ? Ap + (uint64_t)(k0 + q) * lda + i : Ap;
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(As + q * LDA + i);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(dst),
                         "l"(src), "r"(ok ? 8 : 0)
                         : "memory");
        }
        // B chunk: TN columns (j) of WK rows (q); consecutive threads -> q
        for (int e = tid; e < G::TN * WK; e += G::THREADS) {
            const uint32_t j = e / WK, q = e % WK;
            const bool ok = q < kk && j < n;
            const uint32_t qq = ok ? (kperm ? kperm[k0 + q] : k0 + q) : 0;
            const double* src = ok ? Bp + (uint64_t)j * ldb + qq : Bp;
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(Bs + j * LDB + q);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(dst),
                         "l"(src), "r"(ok ? 8 : 0)
                         : "memory");
        }
    };
    auto commit = [] { asm volatile("cp.async.commit_group;" ::: "memory"); };

    const uint32_t nchunk = (k + WK - 1) / WK;
    // prologue: STAGES - 1 chunks in flight (empty groups keep the count)
#pragma unroll
    for (int s = 0; s < G::STAGES - 1; ++s) {
        if ((uint32_t)s < nchunk) load_stage(s, s * WK);
        commit();
    }
    // W_SUB: the accumulators start from the C tile and the MMAs run with -A,
    // so that the tile is read once, up front, with all of its loads in flight
    // behind the first operand stages, and the epilogue only stores (a
    // read-modify-write epilogue of 64 dependent round trips per thread cost
    // twice the time of the MMAs: ncu long-scoreboard 57 % of the samples).
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
        const uint32_t row = wm * (MT * 8) + mt * 8 + g;
        const bool rok = MODE == W_SUB && row < m && !(row >= rs0 && row < rs1);
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const uint32_t col = wn * (NT * 8) + nt * 8 + 2 * t4 + h;
                const bool ok = rok && col < n && !(col >= cs0 && col < cs1);
                acc[mt][nt][h] = ok ? Cp[(uint64_t)col * ldc + row] : 0.0;
            }
        }
    }
    for (uint32_t kc = 0; kc < nchunk; ++kc) {
        if (kc + G::STAGES - 1 < nchunk) {
            load_stage((kc + G::STAGES - 1) % G::STAGES, (kc + G::STAGES - 1) * WK);
        }
        commit();
        // all but the newest STAGES - 1 groups are complete: chunk kc is in
        asm volatile("cp.async.wait_group %0;" ::"n"(G::STAGES - 1) : "memory");
        __syncthreads();
        const double* As = stage_ptr(kc % G::STAGES);
        const double* Bs = As + WK * LDA;
#pragma unroll
        for (int ks = 0; ks < WK / 4; ++ks) {
            double a[MT], b[NT];
#pragma unroll
            for (int mt = 0; mt < MT; ++mt) {
                a[mt] = As[(ks * 4 + t4) * LDA + wm * (MT * 8) + mt * 8 + g];
                if (MODE != W_SET) {
                    // -A: flip the sign bit (integer pipe)
                    a[mt] = __longlong_as_double(__double_as_longlong(a[mt]) ^
                                                 (long long)0x8000000000000000ull);
                }
            }
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                b[nt] = Bs[(wn * (NT * 8) + nt * 8 + g) * LDB + ks * 4 + t4];
            }
#pragma unroll
            for (int mt = 0; mt < MT; ++mt) {
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    asm volatile(
                            "mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 "
                            "{%0,%1}, {%2}, {%3}, {%0,%1};"
                            : "+d"(acc[mt][nt][0]), "+d"(acc[mt][nt][1])
                            : "d"(a[mt]), "d"(b[nt]));
                }
            }
        }
        __syncthreads();
    }
    // epilogue: thread holds C(row, col) and C(row, col + 1)
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
        const uint32_t row = wm * (MT * 8) + mt * 8 + g;
        if (row >= m || (row >= rs0 && row < rs1)) continue;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const uint32_t col = wn * (NT * 8) + nt * 8 + 2 * t4 + h;
                if (col >= n || (col >= cs0 && col < cs1)) continue;
                // W_SUB: C - A B ; W_SETNEG: -(A B) ; W_SET: A B
                Cp[(uint64_t)col * ldc + row] = acc[mt][nt][h];
            }
        }
    }
#endif
}

#ifndef SANM_HOSTEMU
/*! Interior tiles of the trailing update: C(TM x TN) -= A(TM x WB) B(WB x TN),
 * full tile, K = WB, no pivot rows or columns inside.  Same arithmetic as
 * wide_gemm_tile<G, W_SUB> bit for bit, with the instruction count per DMMA cut
 * down -- which is what the tile time is made of: with two warps per scheduler
 * a warp's non-DMMA instructions are not hidden by anything (ncu on the
 * generic tile: 5,858 warp instructions per 1,024 DMMAs, 9 cycles per issued
 * instruction, tensor pipe busy 59 % = 21.9 TFLOP/s; per tile
 * 2 x 1,024 x 16 cycles of DMMA + 4,834 x 4.7 cycles of everything else).
 *   - operand stages: every thread owns one row (A) / one k index (B) and walks
 *     the columns with pointer increments: no index arithmetic, no predicates;
 *   - the accumulators start from -C and the MMAs run with +A (no sign flip
 *     per fragment; -(-C + A B) = C - A B exactly, rounding is symmetric);
 *   - C is read and written through eight column pointers per thread with
 *     immediate offsets, as global (not generic) accesses. */
template <class G>
SANM_D inline void wide_gemm_tile_fast(const rt::BlockCtx& c,
                                       const double* __restrict__ Ap, uint64_t lda,
                                       const double* __restrict__ Bp, uint64_t ldb,
                                       double* __restrict__ Cp, uint64_t ldc) {
    constexpr int WK = G::WK, MT = G::MT, NT = G::NT, LDA = G::LDA, LDB = G::LDB;
    constexpr int TM = G::TM, TN = G::TN, THREADS = G::THREADS;
    constexpr int NCHUNK = WB / WK;
    static_assert(THREADS % TM == 0 && THREADS % WK == 0 && WB % WK == 0, "stage walk");
    constexpr int A_QSTEP = THREADS / TM, A_IT = WK / A_QSTEP;
    constexpr int B_JSTEP = THREADS / WK, B_IT = TN / B_JSTEP;
    constexpr uint32_t STAGE_BYTES = G::STAGE_DOUBLES * sizeof(double);
    const int tid = c.tid, lane = tid & 31, warp = tid >> 5;
    const int wm = warp % G::WM, wn = warp / G::WM;
    const int g = lane >> 2, t4 = lane & 3;
    double* sm = reinterpret_cast<double*>(c.smem);
    const uint32_t sm0 = (uint32_t)__cvta_generic_to_shared(sm);

    const int ai = tid % TM, aq = tid / TM;
    const int bq = tid % WK, bj = tid / WK;
    const double* asrc = Ap + (uint64_t)aq * lda + ai;
    const double* bsrc = Bp + (uint64_t)bj * ldb + bq;
    const uint32_t adst0 = sm0 + (uint32_t)(aq * LDA + ai) * 8u;
    const uint32_t bdst0 = sm0 + (uint32_t)(WK * LDA + bj * LDB + bq) * 8u;
    // The copies of a stage are issued in KS = WK / 4 parts, one part after the
    // DMMAs of each k-step of the chunk before: 32 copies in a row per thread
    // fill the load/store queue and stall the warp -- one of only two DMMA
    // issuers of its scheduler (ncu: lg_throttle 17 % of the samples).
    constexpr int KS = WK / 4;
    static_assert(A_IT % KS == 0 && B_IT % KS == 0, "stage parts");
    constexpr int A_PART = A_IT / KS, B_PART = B_IT / KS;
    const double* a_run = asrc;
    const double* b_run = bsrc;
    uint32_t ad_run = adst0, bd_run = bdst0;
    auto stage_begin = [&](int s, int kc) {
        a_run = asrc + (uint64_t)(kc * WK) * lda;
        b_run = bsrc + kc * WK;
        ad_run = adst0 + (uint32_t)s * STAGE_BYTES;
        bd_run = bdst0 + (uint32_t)s * STAGE_BYTES;
    };
    auto stage_part = [&]() {
#pragma unroll
        for (int it = 0; it < A_PART; ++it) {
            asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(ad_run),
                         "l"(a_run)
                         : "memory");
            a_run += (uint64_t)A_QSTEP * lda;
            ad_run += (uint32_t)(A_QSTEP * LDA) * 8u;
        }
#pragma unroll
        for (int it = 0; it < B_PART; ++it) {
            asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(bd_run),
                         "l"(b_run)
                         : "memory");
            b_run += (uint64_t)B_JSTEP * ldb;
            bd_run += (uint32_t)(B_JSTEP * LDB) * 8u;
        }
    };
    auto commit = [] { asm volatile("cp.async.commit_group;" ::: "memory"); };

#pragma unroll
    for (int s = 0; s < G::STAGES - 1; ++s) {
        if (s < NCHUNK) {
            stage_begin(s, s);
#pragma unroll
            for (int part = 0; part < KS; ++part) stage_part();
        }
        commit();
    }
    double acc[MT][NT][2];
    double* cbase = Cp + (uint64_t)(wn * (NT * 8) + 2 * t4) * ldc + wm * (MT * 8) + g;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const double* col = cbase + (uint64_t)(nt * 8 + h) * ldc;
#pragma unroll
            for (int mt = 0; mt < MT; ++mt) acc[mt][nt][h] = -__ldcg(col + mt * 8);
        }
    }
#pragma unroll 1
    for (int kc = 0; kc < NCHUNK; ++kc) {
        // chunk kc is the oldest group in flight: wait for it alone
        asm volatile("cp.async.wait_group %0;" ::"n"(G::STAGES - 2) : "memory");
        __syncthreads();  // ... and every warp is done with the buffer refilled below
        const bool more = kc + G::STAGES - 1 < NCHUNK;
        if (more) stage_begin((kc + G::STAGES - 1) % G::STAGES, kc + G::STAGES - 1);
        const double* As = sm + (size_t)(kc % G::STAGES) * G::STAGE_DOUBLES +
                           wm * (MT * 8) + g + t4 * LDA;
        const double* Bs = sm + (size_t)(kc % G::STAGES) * G::STAGE_DOUBLES + WK * LDA +
                           (wn * (NT * 8) + g) * LDB + t4;
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
            double a[MT], b[NT];
#pragma unroll
            for (int mt = 0; mt < MT; ++mt) a[mt] = As[ks * 4 * LDA + mt * 8];
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) b[nt] = Bs[nt * 8 * LDB + ks * 4];
#pragma unroll
            for (int mt = 0; mt < MT; ++mt) {
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    asm volatile(
                            "mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 "
                            "{%0,%1}, {%2}, {%3}, {%0,%1};"
                            : "+d"(acc[mt][nt][0]), "+d"(acc[mt][nt][1])
                            : "d"(a[mt]), "d"(b[nt]));
                }
            }
            if (more) stage_part();
        }
        commit();
    }
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            double* col = cbase + (uint64_t)(nt * 8 + h) * ldc;
#pragma unroll
            for (int mt = 0; mt < MT; ++mt) __stcg(col + mt * 8, -acc[mt][nt][h]);
        }
    }
}
#endif

//! [a0, a1) intersected with the tile [t0, t0 + w), in tile-local indices
SANM_HD inline void wide_clip(uint32_t a0, uint32_t a1, uint32_t t0, uint32_t* lo,
                              uint32_t* hi, uint32_t w = WT) {
    const uint32_t t1 = t0 + w;
    const uint32_t b0 = a0 > t0 ? a0 : t0, b1 = a1 < t1 ? a1 : t1;
    if (b0 < b1) {
        *lo = b0 - t0;
        *hi = b1 - t0;
    } else {
        *lo = *hi = 0;
    }
}

/*! D = (P_r F[P,P])^-1 in shared memory, one CTA per front; records the row
 * permutation (lperm: position i of the block holds original row lperm[i]).
 *
 * CUDA build: blocked Gauss-Jordan inside shared memory, 32 pivots at a time:
 * the 32 x 32 diagonal sub-block is inverted by the latency-tuned block_gj32
 * (gj32.cuh, warps 0-3, partial pivoting inside the sub-block), then all 16
 * warps apply it to the 128 x 128 block -- row panel R' = D_b (P R), rank-32
 * update T -= C R' and column panel C' = -C D_b, each thread a 4 x 4 register
 * block.  ~65 us per 128 pivots instead of ~390 us for the unblocked sweep
 * (five block-wide barriers and a rank-1 update per pivot), which was 8 % of a
 * large factorisation and kept fronts below 1024 rows off this path.
 * Host-emulation build: the plain sweep (pivot search over the whole block). */
struct WideDiagKernel {
    const FrontD* fronts;
    const uint32_t* list;
    uint32_t kb;
    const double* amax;
    int* nr_perturbed;
    SANM_D FrontD pre(const rt::BlockCtx& c) const { return fronts[list[c.bid]]; }
    SANM_D void operator()(const rt::BlockCtx& c, const FrontD& f) const {
        if (kb >= f.nc) return;
        const uint32_t wb = f.nc - kb < (uint32_t)WB ? f.nc - kb : (uint32_t)WB;
        double* M = reinterpret_cast<double*>(c.smem);  // [wb][WB]
        double* G = f.F + (uint64_t)kb * f.nr + kb;
        for (uint32_t e = c.tid; e < wb * wb; e += c.nthr) {
            const uint32_t i = e % wb, j = e / wb;
            M[j * WB + i] = G[(uint64_t)j * f.nr + i];
        }
#ifdef SANM_HOSTEMU
        uint32_t* lperm = reinterpret_cast<uint32_t*>(M + (size_t)WB * WB);
        int* sh = reinterpret_cast<int*>(lperm + WB);
        for (uint32_t i = c.tid; i < wb; i += c.nthr) lperm[i] = i;
        c.sync();
        gj_steps(c, M, WB, wb, 0, wb, wb, lperm, sh, amax[0] * 1e-13, nr_perturbed);
#else
        constexpr int SB = 32;
        double* D32 = M + (size_t)WB * WB;   // [32][32] inverted sub-block
        double* Rb = D32 + SB * SB;          // [WB cols][32]: R'(m, j) at Rb[j * 32 + m]
        double* Cb = Rb + (size_t)WB * SB;   // [32][WB rows]: C'(i, m) at Cb[m * WB + i]
        double* acol = Cb + (size_t)WB * SB; // [64]
        uint32_t* lp32 = reinterpret_cast<uint32_t*>(acol + 64);  // [32]
        uint32_t* lperm = lp32 + SB;                              // [WB]
        const double thresh = amax[0] * 1e-13;
        const int tid = c.tid;
        for (uint32_t o = 0; o < wb; o += SB) {
            const uint32_t nb = wb - o < (uint32_t)SB ? wb - o : (uint32_t)SB;
            __syncthreads();
            if (tid < 128) {
                // sub-block -> D32 (zero padded), inverted in place
                const uint32_t lane = tid & 31, j0 = (tid >> 5) * 8;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    D32[(j0 + j) * SB + lane] =
                            (lane < nb && j0 + j < nb) ? M[(o + j0 + j) * WB + o + lane] : 0.0;
                }
                gj32_sync();
                block_gj32(D32, nb, lp32, acol, thresh, nr_perturbed);
            }
            __syncthreads();
            // row panel: R'(m, j) = sum_q D32(m, q) M(o + lp32[q], j), j outside the sub-block
            // 32 x wb outputs, each thread 4 rows (m) x 1 column
            for (uint32_t e = tid; e < 8 * wb; e += c.nthr) {
                const uint32_t j = e / 8, m4 = (e % 8) * 4;
                double acc[4] = {0, 0, 0, 0};
                if (j < o || j >= o + nb) {
                    for (uint32_t q = 0; q < nb; ++q) {
                        const double r = M[j * WB + o + lp32[q]];
#pragma unroll
                        for (int u = 0; u < 4; ++u) acc[u] += D32[q * SB + m4 + u] * r;
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) Rb[j * SB + m4 + u] = acc[u];
            }
            // column panel: C'(i, m) = -sum_q M(i, o + q) D32(q, m), i outside the sub-block
            for (uint32_t e = tid; e < 8 * wb; e += c.nthr) {
                const uint32_t i = e % wb, m4 = (e / wb) * 4;
                double acc[4] = {0, 0, 0, 0};
                if (i < o || i >= o + nb) {
                    for (uint32_t q = 0; q < nb; ++q) {
                        const double cv = M[(o + q) * WB + i];
#pragma unroll
                        for (int u = 0; u < 4; ++u) acc[u] += cv * D32[(m4 + u) * SB + q];
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) Cb[(m4 + u) * WB + i] = -acc[u];
            }
            __syncthreads();
            // trailing update: M(i, j) -= sum_m M(i, o + m) R'(m, j), i and j outside;
            // 4 x 4 register blocks over the wb x wb block
            const uint32_t nt = (wb + 3) / 4;
            for (uint32_t e = tid; e < nt * nt; e += c.nthr) {
                const uint32_t i4 = (e % nt) * 4, j4 = (e / nt) * 4;
                double acc[4][4] = {};
                for (uint32_t m = 0; m < nb; ++m) {
                    double a[4], b[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) a[u] = i4 + u < wb ? M[(o + m) * WB + i4 + u] : 0.0;
#pragma unroll
                    for (int v = 0; v < 4; ++v) b[v] = j4 + v < wb ? Rb[(j4 + v) * SB + m] : 0.0;
#pragma unroll
                    for (int u = 0; u < 4; ++u)
#pragma unroll
                        for (int v = 0; v < 4; ++v) acc[u][v] += a[u] * b[v];
                }
#pragma unroll
                for (int v = 0; v < 4; ++v) {
                    const uint32_t j = j4 + v;
                    if (j >= wb || (j >= o && j < o + nb)) continue;
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const uint32_t i = i4 + u;
                        if (i >= wb || (i >= o && i < o + nb)) continue;
                        M[j * WB + i] -= acc[u][v];
                    }
                }
            }
            __syncthreads();
            // write the three finished pieces back into M
            for (uint32_t e = tid; e < nb * wb; e += c.nthr) {
                const uint32_t m = e % nb, j = e / nb;
                if (j < o || j >= o + nb) {
                    M[j * WB + o + m] = Rb[j * SB + m];
                } else {
                    M[j * WB + o + m] = D32[(j - o) * SB + m];
                }
            }
            for (uint32_t e = tid; e < nb * wb; e += c.nthr) {
                const uint32_t i = e % wb, m = e / wb;
                if (i < o || i >= o + nb) M[(o + m) * WB + i] = Cb[m * WB + i];
            }
            if (tid < (int)nb) lperm[o + tid] = o + lp32[tid];
        }
        __syncthreads();
#endif
        for (uint32_t e = c.tid; e < wb * wb; e += c.nthr) {
            const uint32_t i = e % wb, j = e / wb;
            G[(uint64_t)j * f.nr + i] = M[j * WB + i];
        }
        for (uint32_t i = c.tid; i < wb; i += c.nthr) {
            f.lperm[i] = lperm[i];
            f.rowperm[kb + i] = kb + lperm[i];
        }
    }
};
#ifdef SANM_HOSTEMU
constexpr size_t W_DIAG_SMEM =
        (size_t)WB * WB * sizeof(double) + WB * sizeof(uint32_t) + 16;
#else
constexpr size_t W_DIAG_SMEM =
        ((size_t)WB * WB + 32 * 32 + 2 * (size_t)WB * 32 + 64) * sizeof(double) +
        (32 + WB) * sizeof(uint32_t);
#endif

#ifndef SANM_HOSTEMU
/*! Persistent form of the trailing update: one CTA per SM walks the tiles of
 * the launch (t = bid, bid + grid, ...) and, while the DMMAs of one interior
 * tile run, brings in the NEXT tile's C (128 KB, into a shared-memory buffer)
 * and its first operand stage.  What a tile costs beyond its DMMAs is the data
 * that has to arrive before the first of them (C + one stage; a third stage in
 * the prologue made the plain kernel 22 % slower) -- here that wait belongs to
 * the previous tile.  16-deep stages leave room for the C buffer:
 * 2 x 36.5 KB + 130 KB.  Edge tiles and tiles that contain pivot rows or
 * columns go through the generic tile function (nothing prefetched across
 * them).  Arithmetic is that of wide_gemm_tile_fast, bit for bit.
 * The copies are spread over the k-steps (8 A, 8 B, 4 C, 4 C per thread and
 * chunk): issuing a chunk's 16 stage copies in one burst cost 12 % (the
 * load/store queue fills and stalls one of the two DMMA issuers). */
struct WidePersist {
    static constexpr int WK = 16, KS = WK / 4, NCHUNK = WB / WK, THREADS = 256;
    static constexpr int WM = 2, WN = 4, MT = 8, NT = 4;
    static constexpr int LDA = WT + 4, LDB = WK + 4, LDC = WT + 2;
    static constexpr int STAGE_DOUBLES = WK * LDA + WT * LDB;
    static constexpr int CBUF_DOUBLES = WT * LDC;
    static constexpr size_t SMEM =
            (size_t)(2 * STAGE_DOUBLES + CBUF_DOUBLES) * sizeof(double);
    static constexpr int A_IT = WK * WT / THREADS, B_IT = WT * WK / THREADS;  // 8, 8
    static constexpr int C_IT = WT * WT / THREADS;                            // 64
    static_assert(SMEM >= WideGemmBig::SMEM && (LDB % 16) == 4 && (NCHUNK % 2) == 0, "layout");
};

struct WideGemmPersistKernel {
    using P = WidePersist;
    const FrontD* fronts;
    const uint32_t* list;
    uint32_t kb, tiles, nfront;
    //! the tile of the NEXT pivot block was updated by WidePivotTileKernel and is
    //! being inverted on the side stream while this kernel runs: not touched here
    int skip_next_pivot;

    struct Tile {
        int kind;  //!< 0 nothing to do, 1 interior (fast), 2 generic
        double* F;
        uint32_t nr, nc, i0, j0;
    };
    SANM_D Tile describe(uint64_t t, uint32_t& cached_fi, FrontD& cached) const {
        const uint64_t per = (uint64_t)tiles * tiles;
        const uint32_t fi = (uint32_t)(t / per), idx = (uint32_t)(t % per);
        if (fi != cached_fi) {
            cached = fronts[list[fi]];
            cached_fi = fi;
        }
        Tile d;
        d.F = cached.F;
        d.nr = cached.nr;
        d.nc = cached.nc;
        d.i0 = (idx % tiles) * WT;
        d.j0 = (idx / tiles) * WT;
        d.kind = 0;
        if (kb >= d.nc || d.i0 >= d.nr || d.j0 >= d.nr) return d;
        if (skip_next_pivot && d.i0 == kb + WB && d.j0 == kb + WB && kb + WB < d.nc) return d;
        const uint32_t wb = d.nc - kb < (uint32_t)WB ? d.nc - kb : (uint32_t)WB;
        const uint32_t mi = d.nr - d.i0 < (uint32_t)WT ? d.nr - d.i0 : (uint32_t)WT;
        const uint32_t nj = d.nr - d.j0 < (uint32_t)WT ? d.nr - d.j0 : (uint32_t)WT;
        uint32_t rs0, rs1, cs0, cs1;
        wide_clip(kb, kb + wb, d.i0, &rs0, &rs1);
        wide_clip(kb, kb + wb, d.j0, &cs0, &cs1);
        if (rs1 - rs0 == mi || cs1 - cs0 == nj) return d;
        d.kind = (wb == (uint32_t)WB && mi == (uint32_t)WT && nj == (uint32_t)WT &&
                  rs0 == rs1 && cs0 == cs1)
                         ? 1
                         : 2;
        return d;
    }
    SANM_D void generic(const rt::BlockCtx& c, const Tile& d) const {
        const uint32_t wb = d.nc - kb < (uint32_t)WB ? d.nc - kb : (uint32_t)WB;
        const uint32_t mi = d.nr - d.i0 < (uint32_t)WT ? d.nr - d.i0 : (uint32_t)WT;
        const uint32_t nj = d.nr - d.j0 < (uint32_t)WT ? d.nr - d.j0 : (uint32_t)WT;
        uint32_t rs0, rs1, cs0, cs1;
        wide_clip(kb, kb + wb, d.i0, &rs0, &rs1);
        wide_clip(kb, kb + wb, d.j0, &cs0, &cs1);
        wide_gemm_tile<WideGemmBig, W_SUB>(c, d.F + (uint64_t)kb * d.nr + d.i0, d.nr,
                                           d.F + (uint64_t)d.j0 * d.nr + kb, d.nr, nullptr,
                                           d.F + (uint64_t)d.j0 * d.nr + d.i0, d.nr, mi, nj,
                                           wb, rs0, rs1, cs0, cs1);
    }
    SANM_D void operator()(const rt::BlockCtx& c) const;
};
#endif

#ifndef SANM_HOSTEMU
SANM_D inline void WideGemmPersistKernel::operator()(const rt::BlockCtx& c) const {
    constexpr int WK = P::WK, NCHUNK = P::NCHUNK, MT = P::MT, NT = P::NT;
    constexpr int LDA = P::LDA, LDB = P::LDB, LDC = P::LDC;
    constexpr uint32_t STAGE_BYTES = P::STAGE_DOUBLES * sizeof(double);
    const int tid = c.tid, lane = tid & 31, warp = tid >> 5;
    const int wm = warp % P::WM, wn = warp / P::WM;
    const int g = lane >> 2, t4 = lane & 3;
    double* sm = reinterpret_cast<double*>(c.smem);
    double* cbuf = sm + 2 * P::STAGE_DOUBLES;
    const uint32_t sm0 = (uint32_t)__cvta_generic_to_shared(sm);
    // the copies every thread owns: A row ai, k columns aq, aq + 2, ...; B k index
    // bq, columns bj, bj + 16, ...; C row ai, columns aq, aq + 2, ...
    const int ai = tid % WT, aq = tid / WT;
    const int bq = tid % WK, bj = tid / WK;
    const uint32_t adst0 = sm0 + (uint32_t)(aq * LDA + ai) * 8u;
    const uint32_t bdst0 = sm0 + (uint32_t)(WK * LDA + bj * LDB + bq) * 8u;
    const uint32_t cdst0 = sm0 + 2u * STAGE_BYTES + (uint32_t)(aq * LDC + ai) * 8u;
    const double* As0 = sm + wm * (MT * 8) + g + t4 * LDA;
    const double* Bs0 = sm + WK * LDA + (wn * (NT * 8) + g) * LDB + t4;
    const double* Cs0 = cbuf + (wn * (NT * 8) + 2 * t4) * LDC + wm * (MT * 8) + g;
    auto cp8 = [](uint32_t dst, const double* src) {
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src)
                     : "memory");
    };
    auto commit = [] { asm volatile("cp.async.commit_group;" ::: "memory"); };
    auto wait_all = [] { asm volatile("cp.async.wait_group 0;" ::: "memory"); };

    const uint64_t total = (uint64_t)nfront * tiles * tiles;
    uint32_t cfi = ~0u;
    FrontD cf;
    uint64_t t = c.bid;
    Tile cur{};
    for (; t < total; t += c.nblk) {
        cur = describe(t, cfi, cf);
        if (cur.kind) break;
    }
    bool primed = false;
    while (t < total) {
        uint64_t tn = t + c.nblk;
        Tile nxt{};
        for (; tn < total; tn += c.nblk) {
            nxt = describe(tn, cfi, cf);
            if (nxt.kind) break;
        }
        if (cur.kind == 2) {
            // nothing is in flight here: the tile before did not prefetch into a
            // generic tile, and its own copies were all waited for
            __syncthreads();
            generic(c, cur);
            __syncthreads();
            primed = false;
        } else {
            const uint64_t ld = cur.nr;
            const double* a_cur = cur.F + (uint64_t)(kb + aq) * ld + cur.i0 + ai;
            const double* b_cur = cur.F + (uint64_t)(cur.j0 + bj) * ld + kb + bq;
            double* c_cur = cur.F + (uint64_t)cur.j0 * ld + cur.i0;
            const bool pf = tn < total && nxt.kind == 1;
            const uint64_t ldn = nxt.nr;
            const double* a_nxt = nxt.F + (uint64_t)(kb + aq) * ldn + nxt.i0 + ai;
            const double* b_nxt = nxt.F + (uint64_t)(nxt.j0 + bj) * ldn + kb + bq;
            const double* c_nxt = nxt.F + (uint64_t)(nxt.j0 + aq) * ldn + nxt.i0 + ai;
            if (!primed) {
                const double* cs = c_cur + (uint64_t)aq * ld + ai;
#pragma unroll 8
                for (int it = 0; it < P::C_IT; ++it) {
                    cp8(cdst0 + (uint32_t)(it * 2 * LDC) * 8u, cs);
                    cs += 2 * ld;
                }
                const double* a = a_cur;
#pragma unroll
                for (int it = 0; it < P::A_IT; ++it) {
                    cp8(adst0 + (uint32_t)(it * 2 * LDA) * 8u, a);
                    a += 2 * ld;
                }
                const double* b = b_cur;
#pragma unroll
                for (int it = 0; it < P::B_IT; ++it) {
                    cp8(bdst0 + (uint32_t)(it * 16 * LDB) * 8u, b);
                    b += 16 * ld;
                }
                commit();
            }
            wait_all();
            __syncthreads();
            double acc[MT][NT][2];
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int mt = 0; mt < MT; ++mt)
                        acc[mt][nt][h] = -Cs0[(nt * 8 + h) * LDC + mt * 8];
            __syncthreads();  // the C buffer may be refilled from here on
            uint32_t cn_dst = cdst0;
#pragma unroll 1
            for (int kc = 0; kc < NCHUNK; ++kc) {
                if (kc) {
                    wait_all();
                    __syncthreads();
                }
                // what is staged during this chunk: chunk kc + 1 of this tile, or
                // chunk 0 of the next one
                const bool last = kc + 1 == NCHUNK;
                const bool do_stage = !last || pf;
                const uint64_t sld = last ? ldn : ld;
                const double* sa = last ? a_nxt : a_cur + (uint64_t)((kc + 1) * WK) * ld;
                const double* sb = last ? b_nxt : b_cur + (kc + 1) * WK;
                const uint32_t sbuf = (uint32_t)((kc + 1) & 1) * STAGE_BYTES;
                const double* As = As0 + (size_t)(kc & 1) * P::STAGE_DOUBLES;
                const double* Bs = Bs0 + (size_t)(kc & 1) * P::STAGE_DOUBLES;
#pragma unroll
                for (int ks = 0; ks < P::KS; ++ks) {
                    double a[MT], b[NT];
#pragma unroll
                    for (int mt = 0; mt < MT; ++mt) a[mt] = As[ks * 4 * LDA + mt * 8];
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt) b[nt] = Bs[nt * 8 * LDB + ks * 4];
#pragma unroll
                    for (int mt = 0; mt < MT; ++mt) {
#pragma unroll
                        for (int nt = 0; nt < NT; ++nt) {
                            asm volatile(
                                    "mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 "
                                    "{%0,%1}, {%2}, {%3}, {%0,%1};"
                                    : "+d"(acc[mt][nt][0]), "+d"(acc[mt][nt][1])
                                    : "d"(a[mt]), "d"(b[nt]));
                        }
                    }
                    if (ks == 0) {
                        if (do_stage) {
#pragma unroll
                            for (int it = 0; it < P::A_IT; ++it) {
                                cp8(adst0 + sbuf + (uint32_t)(it * 2 * LDA) * 8u, sa);
                                sa += 2 * sld;
                            }
                        }
                    } else if (ks == 1) {
                        if (do_stage) {
#pragma unroll
                            for (int it = 0; it < P::B_IT; ++it) {
                                cp8(bdst0 + sbuf + (uint32_t)(it * 16 * LDB) * 8u, sb);
                                sb += 16 * sld;
                            }
                        }
                        // this group: the stage above + the C copies of the chunk
                        // before, which so get a whole chunk to arrive
                        commit();
                    } else if (pf) {
                        // next tile's C: 64 copies per thread, 4 per slot
#pragma unroll
                        for (int it = 0; it < P::C_IT / (2 * NCHUNK); ++it) {
                            cp8(cn_dst, c_nxt);
                            c_nxt += 2 * ldn;
                            cn_dst += (uint32_t)(2 * LDC) * 8u;
                        }
                    }
                }
            }
            commit();
            double* cst = c_cur + (uint64_t)(wn * (NT * 8) + 2 * t4) * ld + wm * (MT * 8) + g;
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    double* col = cst + (uint64_t)(nt * 8 + h) * ld;
#pragma unroll
                    for (int mt = 0; mt < MT; ++mt) __stcg(col + mt * 8, -acc[mt][nt][h]);
                }
            primed = pf;
        }
        t = tn;
        cur = nxt;
    }
}
#endif

/*! Look-ahead of the pivot-block inverse: the one trailing tile of step kb that
 * holds the NEXT pivot block, updated ahead of the rest of the trailing
 * update, so that WideDiagKernel(kb + WB) -- a single CTA per front, 0.19 ms,
 * during which the other 147 SMs used to idle -- can run on a side stream
 * beside the trailing update instead of after it.  Same tile function and
 * operands as the trailing update: results do not change. */
struct WidePivotTileKernel {
    const FrontD* fronts;
    const uint32_t* list;
    uint32_t kb;
    SANM_D FrontD pre(const rt::BlockCtx& c) const { return fronts[list[c.bid]]; }
    SANM_D void operator()(const rt::BlockCtx& c, const FrontD& f) const {
        const uint32_t p0 = kb + WB;
        if (p0 >= f.nc) return;  // no next step for this front
        const uint32_t ext = f.nr - p0 < (uint32_t)WT ? f.nr - p0 : (uint32_t)WT;
        wide_gemm_tile<WideGemmBig, W_SUB>(c, f.F + (uint64_t)kb * f.nr + p0, f.nr,
                                           f.F + (uint64_t)p0 * f.nr + kb, f.nr, nullptr,
                                           f.F + (uint64_t)p0 * f.nr + p0, f.nr, ext, ext,
                                           (uint32_t)WB, 0, 0, 0, 0);
    }
};

//! ROW: R' = D (P_r R) on one 128-column tile; COL: C' = -C D on one 128-row tile
template <bool ROW>
struct WidePanelKernel {
    const FrontD* fronts;
    const uint32_t* list;
    uint32_t kb, tiles;
    SANM_D FrontD pre(const rt::BlockCtx& c) const { return fronts[list[c.bid / tiles]]; }
    SANM_D void operator()(const rt::BlockCtx& c, const FrontD& f) const {
        const uint32_t t0 = (uint32_t)(c.bid % tiles) * WT;
        if (kb >= f.nc || t0 >= f.nr) return;
        const uint32_t wb = f.nc - kb < (uint32_t)WB ? f.nc - kb : (uint32_t)WB;
        const uint32_t ext = f.nr - t0 < (uint32_t)WT ? f.nr - t0 : (uint32_t)WT;
        uint32_t s0, s1;
        wide_clip(kb, kb + wb, t0, &s0, &s1);
        if (s1 - s0 == ext) return;  // the tile is the pivot block itself
        double* D = f.F + (uint64_t)kb * f.nr + kb;
        if (ROW) {
            double* R = f.F + (uint64_t)t0 * f.nr + kb;  // rows kb.., columns t0..
            wide_gemm_tile<WidePanelCfg, W_SET>(c, D, f.nr, R, f.nr, f.lperm, R, f.nr, wb, ext, wb,
                                  0, 0, s0, s1);
        } else {
            double* C = f.F + (uint64_t)kb * f.nr + t0;  // rows t0.., columns kb..
            wide_gemm_tile<WidePanelCfg, W_SETNEG>(c, C, f.nr, D, f.nr, nullptr, C, f.nr, ext, wb,
                                     wb, s0, s1, 0, 0);
        }
    }
};

//! F[Q,Q] -= F[Q,P] R' on one TM x TN tile (tiles_m x tiles_n tiles per front)
template <class G>
struct WideGemmKernelT {
    const FrontD* fronts;
    const uint32_t* list;
    uint32_t kb, tiles_m, tiles_n;
    int wide_fast;  //!< interior tiles take wide_gemm_tile_fast (SANM_LU_WIDE_SLOW=1: off)
    SANM_D FrontD pre(const rt::BlockCtx& c) const {
        return fronts[list[c.bid / ((size_t)tiles_m * tiles_n)]];
    }
    SANM_D void operator()(const rt::BlockCtx& c, const FrontD& f) const {
        const uint32_t t = (uint32_t)(c.bid % ((size_t)tiles_m * tiles_n));
        const uint32_t i0 = (t % tiles_m) * G::TM, j0 = (t / tiles_m) * G::TN;
        if (kb >= f.nc || i0 >= f.nr || j0 >= f.nr) return;
        const uint32_t wb = f.nc - kb < (uint32_t)WB ? f.nc - kb : (uint32_t)WB;
        const uint32_t mi = f.nr - i0 < (uint32_t)G::TM ? f.nr - i0 : (uint32_t)G::TM;
        const uint32_t nj = f.nr - j0 < (uint32_t)G::TN ? f.nr - j0 : (uint32_t)G::TN;
        uint32_t rs0, rs1, cs0, cs1;
        wide_clip(kb, kb + wb, i0, &rs0, &rs1, G::TM);
        wide_clip(kb, kb + wb, j0, &cs0, &cs1, G::TN);
        if (rs1 - rs0 == mi || cs1 - cs0 == nj) return;  // nothing to store
#ifndef SANM_HOSTEMU
        if (wide_fast && wb == (uint32_t)WB && mi == (uint32_t)G::TM &&
            nj == (uint32_t)G::TN && rs0 == rs1 && cs0 == cs1) {
            wide_gemm_tile_fast<G>(c, f.F + (uint64_t)kb * f.nr + i0, f.nr,
                                   f.F + (uint64_t)j0 * f.nr + kb, f.nr,
                                   f.F + (uint64_t)j0 * f.nr + i0, f.nr);
            return;
        }
#endif
        wide_gemm_tile<G, W_SUB>(c, f.F + (uint64_t)kb * f.nr + i0, f.nr,
                                 f.F + (uint64_t)j0 * f.nr + kb, f.nr, nullptr,
                                 f.F + (uint64_t)j0 * f.nr + i0, f.nr, mi, nj, wb, rs0,
                                 rs1, cs0, cs1);
    }
};
