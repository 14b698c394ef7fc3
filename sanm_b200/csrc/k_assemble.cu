/* sanm_b200/csrc/k_assemble.cu -- element-wise assembly of the tangent matrix
 * K = R_out blockdiag(J_e) R_in (libsanm/anm.cpp:362-438, 520-608) for the
 * mesh remaps of `fea` (fused_io.h).
 *
 * Two kernels, no atomics, bitwise reproducible:
 *   1. one thread per tet forms its element matrix
 *        Ke[(d,c)][col] = sum_j n_d[j] sum_q J[(c,j)][q] coef(q -> col)
 *      (12 row slots x <= 13 distinct unknowns) and stores it structure-of-
 *      arrays into a contribution buffer [156][npad];
 *   2. one thread per CSR entry of K (and per entry of grad_t, the column of
 *      the extra unknown t of the implicit solver) adds its contributions in
 *      ascending tet order through a precomputed source list.
 * The generic row-wise kernel (anm.cu AssembleK) remains for every other pair
 * of descriptors.
 */
#include "anm.h"
#include "prof.h"

#include <algorithm>
#include <omp.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>

namespace sanm_b200 {

namespace {

constexpr int KE_COLS = 13;
constexpr int KE_ENTRIES = 12 * KE_COLS;
constexpr uint32_t NONE32 = 0xFFFFFFFFu;

struct ElementK {
    const uint8_t* tab;   // [27][npad]: col slot | code << 4
    const double* gcoef;  // [9][npad] or null
    const double* nrm;    // [12][npad]
    const double* jac;    // [81][pitch]
    double* cbuf;         // [156][npad]
    int64_t npad, pitch;
    SANM_D void operator()(size_t e) const {
        uint8_t t[27];
#pragma unroll
        for (int i = 0; i < 27; ++i) t[i] = tab[(int64_t)i * npad + e];
        double gc[9];
#pragma unroll
        for (int q = 0; q < 9; ++q) gc[q] = gcoef ? gcoef[(int64_t)q * npad + e] : 0.0;
        for (int c = 0; c < 3; ++c) {
            double Jc[27];
#pragma unroll
            for (int i = 0; i < 27; ++i) {
                Jc[i] = jac[(int64_t)(c * 27 + i) * pitch + e];
            }
            for (int d = 0; d < 4; ++d) {
                const double n0 = nrm[(int64_t)(d * 3 + 0) * npad + e];
                const double n1 = nrm[(int64_t)(d * 3 + 1) * npad + e];
                const double n2 = nrm[(int64_t)(d * 3 + 2) * npad + e];
                double vec[KE_COLS];
#pragma unroll
                for (int s = 0; s < KE_COLS; ++s) vec[s] = 0;
                for (int q = 0; q < 9; ++q) {
                    const double a = n0 * Jc[q] + n1 * Jc[9 + q] + n2 * Jc[18 + q];
                    for (int w = 0; w < 3; ++w) {
                        const int code = t[q * 3 + w] >> 4;
                        if (!code) continue;
                        const double cf = code == 1 ? 1.0 : code == 2 ? -1.0 : gc[q];
                        vec[t[q * 3 + w] & 15] += a * cf;
                    }
                }
                double* dst = cbuf + (int64_t)((d * 3 + c) * KE_COLS) * npad + e;
#pragma unroll
                for (int s = 0; s < KE_COLS; ++s) dst[(int64_t)s * npad] = vec[s];
            }
        }
    }
};

struct GatherK {
    const uint32_t* kptr;
    const uint32_t* ksrc;
    const double* cbuf;
    double* vals;
    double* grad_t;
    uint64_t nnz;
    SANM_D void operator()(size_t p) const {
        double s = 0;
        for (uint32_t i = kptr[p], ie = kptr[p + 1]; i < ie; ++i) s += cbuf[ksrc[i]];
        if (p < nnz) {
            vals[p] = s;
        } else if (grad_t) {
            grad_t[p - nnz] = s;
        }
    }
};

}  // namespace

std::unique_ptr<FusedKAssembly> FusedKAssembly::build(
        const FusedIO& io, uint64_t n, const std::vector<uint32_t>& k_rowptr,
        const std::vector<uint32_t>& k_colidx) {
    return finish(build_host(io, n, k_rowptr, k_colidx));
}

std::unique_ptr<FusedKAssembly> FusedKAssembly::finish(HostLists&& h) {
    if (!h.ok) return nullptr;
    auto ret = std::unique_ptr<FusedKAssembly>(new FusedKAssembly);
    ret->m_n_tets = h.n_tets;
    ret->m_npad = h.npad;
    ret->m_nnz = h.nnz;
    ret->m_n = h.n;
    ret->m_tab.upload(h.tab);
    ret->m_kptr.upload(h.kptr);
    ret->m_ksrc.upload(h.ksrc);
    ret->m_cbuf.alloc((size_t)KE_ENTRIES * h.npad);
    return ret;
}

FusedKAssembly::HostLists FusedKAssembly::build_host(
        const FusedIO& io, uint64_t n, const std::vector<uint32_t>& k_rowptr,
        const std::vector<uint32_t>& k_colidx) {
    HostLists out;
    const uint64_t N = io.n_tets;
    const int64_t npad = io.npad;
    if ((uint64_t)KE_ENTRIES * npad >= (1ull << 32)) return out;
    const uint64_t nnz = k_rowptr[n];
    if (nnz + n >= (1ull << 32) || io.nout != n) return out;
    const bool dbg = getenv("SANM_LU_DEBUG") != nullptr;
    auto tnow = [] { return std::chrono::steady_clock::now(); };
    auto t_start = tnow();
    auto lap = [&](const char* what) {
        if (!dbg) return;
        auto t1 = tnow();
        printf("[k-lists] %-28s %8.2f ms\n", what,
               std::chrono::duration<double, std::milli>(t1 - t_start).count());
        t_start = t1;
    };
    std::vector<uint8_t> tab((size_t)27 * npad, 0);
    std::vector<uint32_t> dloc((size_t)KE_ENTRIES * N, NONE32);
    bool bad = false;
#pragma omp parallel for num_threads(::sanm_b200::rt::host_threads()) reduction(|| : bad)
    for (int64_t e = 0; e < (int64_t)N; ++e) {
        uint32_t cols[KE_COLS];
        int nc = 0;
        bool fail = false;
        auto slot_of = [&](uint32_t col) {
            for (int s = 0; s < nc; ++s) {
                if (cols[s] == col) return s;
            }
            if (nc == KE_COLS) {
                fail = true;
                return 0;
            }
            cols[nc] = col;
            return nc++;
        };
        for (int q = 0; q < 9; ++q) {
            for (int w = 0; w < io.wu; ++w) {
                const uint32_t u = io.h_uidx[((size_t)w * 9 + q) * npad + e];
                if (u == NONE32) continue;
                tab[(size_t)(q * 3 + w) * npad + e] =
                        (uint8_t)(slot_of(u >> 1) | ((1 + (u & 1)) << 4));
            }
            if (io.has_general) {
                const uint32_t g = io.h_gidx[(size_t)q * npad + e];
                if (g != NONE32) {
                    tab[(size_t)(q * 3 + 2) * npad + e] =
                            (uint8_t)(slot_of(g) | (3 << 4));
                }
            }
        }
        if (fail) {
            bad = true;
            continue;
        }
        for (int s = 0; s < 12; ++s) {
            const uint32_t row = io.h_grow[(size_t)s * npad + e];
            if (row == NONE32) continue;
            const uint32_t* rb = k_colidx.data() + k_rowptr[row];
            const uint32_t* re = k_colidx.data() + k_rowptr[row + 1];
            for (int cs = 0; cs < nc; ++cs) {
                uint32_t dest;
                if (cols[cs] >= n) {
                    dest = (uint32_t)(nnz + row);  // column of t -> grad_t
                } else {
                    const uint32_t* it = std::lower_bound(rb, re, cols[cs]);
                    if (it == re || *it != cols[cs]) {
                        bad = true;
                        continue;
                    }
                    dest = (uint32_t)(it - k_colidx.data());
                }
                dloc[(size_t)e * KE_ENTRIES + s * KE_COLS + cs] = dest;
            }
        }
    }
    lap("dloc (parallel)");
    if (bad) return out;
    // source lists per destination, ascending (tet, local entry).  Bucketed
    // counting sort: every thread walks its own tet range once and files its
    // (destination, source) pairs by destination range; then every range is
    // counted and filled from the files in thread order = tet order.  The order
    // inside a list is the tet order whatever the thread count.
    const uint64_t nd = nnz + n;
    std::vector<uint32_t> kptr(nd + 1, 0);
    const int T = rt::host_threads();
    struct Pair {
        uint32_t d, src;
    };
    std::vector<std::vector<Pair>> files((size_t)T * T);  // [walker * T + range]
    auto range_of = [nd, T](uint32_t d) { return (int)((uint64_t)d * T / nd); };
#pragma omp parallel num_threads(T)
    {
        const int t = omp_get_thread_num(), nt = omp_get_num_threads();
        // (nt == T: the team size was requested explicitly; fewer threads still
        // work, each one walks several tet ranges)
        for (int w = t; w < T; w += nt) {
            const uint64_t e0 = N * w / T, e1 = N * (w + 1) / T;
            std::vector<Pair>* mine = files.data() + (size_t)w * T;
            for (int r = 0; r < T; ++r) mine[r].reserve((e1 - e0) * 40 / T + 16);
            for (uint64_t e = e0; e < e1; ++e) {
                const uint32_t* row = dloc.data() + (size_t)e * KE_ENTRIES;
                for (int l = 0; l < KE_ENTRIES; ++l) {
                    const uint32_t d = row[l];
                    if (d == NONE32) continue;
                    mine[range_of(d)].push_back(
                            Pair{d, (uint32_t)((uint64_t)l * npad + e)});
                }
            }
        }
    }
    lap("file by range");
#pragma omp parallel num_threads(T)
    {
        const int t = omp_get_thread_num(), nt = omp_get_num_threads();
        for (int r = t; r < T; r += nt) {
            for (int w = 0; w < T; ++w) {
                for (const Pair& p : files[(size_t)w * T + r]) ++kptr[p.d + 1];
            }
        }
    }
    lap("count");
    for (uint64_t p = 0; p < nd; ++p) kptr[p + 1] += kptr[p];
    std::vector<uint32_t> ksrc(kptr[nd]);
    {
        std::vector<uint32_t> fill(kptr.begin(), kptr.end() - 1);
#pragma omp parallel num_threads(T)
        {
            const int t = omp_get_thread_num(), nt = omp_get_num_threads();
            for (int r = t; r < T; r += nt) {
                for (int w = 0; w < T; ++w) {
                    for (const Pair& p : files[(size_t)w * T + r]) ksrc[fill[p.d]++] = p.src;
                }
            }
        }
    }
    lap("fill");
    out.ok = true;
    out.n_tets = N;
    out.npad = npad;
    out.nnz = nnz;
    out.n = n;
    out.tab = std::move(tab);
    out.kptr = std::move(kptr);
    out.ksrc = std::move(ksrc);
    return out;
}

void FusedKAssembly::assemble(double* vals, double* grad_t, const double* jac,
                              int64_t jac_pitch, const FusedIO& io) const {
    sanm_check(io.npad == m_npad && jac_pitch > 0);
    rt::for_each(m_n_tets, ElementK{m_tab.ptr(), io.gcoef.ptr(), io.nrm.ptr(),
                                    jac, m_cbuf.ptr(), m_npad, jac_pitch});
    rt::for_each(m_nnz + m_n, GatherK{m_kptr.ptr(), m_ksrc.ptr(), m_cbuf.ptr(),
                                      vals, grad_t, m_nnz});
}

}  // namespace sanm_b200
