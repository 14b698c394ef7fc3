/* sanm_b200/csrc/fused_io.cu -- see fused_io.h. */
#include "fused_io.h"

#include <algorithm>
#include <cstring>

namespace sanm_b200 {

using fused::TILE;

namespace {

struct FinalizeKernel {
    const uint32_t* fin_ptr;
    const uint32_t* fin_idx;
    const double* partial;
    double* b;
    uint64_t row0;
    SANM_D void operator()(size_t i) const {
        const size_t row = row0 + i;
        double s = 0;
        for (uint32_t p = fin_ptr[row], pe = fin_ptr[row + 1]; p < pe; ++p) {
            s += partial[fin_idx[p]];
        }
        b[row] = s;
    }
};

constexpr uint32_t NONE32 = 0xFFFFFFFFu;
constexpr uint16_t NONE16 = 0xFFFFu;

}  // namespace

void FusedIO::finalize(double* b, bool local_only) const {
    const uint64_t r0 = local_only ? row_lo : 0;
    const uint64_t r1 = local_only ? row_hi : nout;
    if (r1 > r0) {
        rt::for_each(r1 - r0, FinalizeKernel{fin_ptr.ptr(), fin_idx.ptr(),
                                             partial.ptr(), b, r0});
    }
}

uint64_t FusedIO::table_bytes() const {
    return uidx.size() * 4 + gidx.size() * 4 + gcoef.size() * 8 +
           nrm.size() * 8 + lrow.size() * 2 + con_idx.size() * 2 +
           con_ptr.size() * 4;
}

std::unique_ptr<FusedIO> FusedIO::analyse(const Remap& rin, const Remap& rout,
                                          uint64_t n_tets) {
    const uint64_t N = n_tets;
    if (rin.nr_out() != N * 9 || rout.nr_in() != N * 9) return nullptr;
    if (rin.nr_in() >= (1ull << 31) || rout.nr_out() >= (1ull << 31)) {
        return nullptr;
    }
    auto io = std::make_unique<FusedIO>();
    io->n_tets = N;
    io->nin = rin.nr_in();
    io->nout = rout.nr_out();
    const int64_t npad = (int64_t)((N + TILE - 1) / TILE * TILE);
    io->npad = npad;
    const size_t ntiles = (size_t)npad / TILE;

    /* ---------------- gather: signed unit references + one general */
    {
        const auto& rp = rin.h_rowptr();
        const auto& ix = rin.h_idx();
        const auto& cf = rin.h_coeff();
        int wu = 0;
        bool bad = false, general = false;
#pragma omp parallel for num_threads(::sanm_b200::rt::host_threads()) reduction(max : wu) reduction(|| : bad, general)
        for (int64_t r = 0; r < (int64_t)(N * 9); ++r) {
            int nu = 0, ng = 0;
            bool seen_general = false;
            for (uint32_t p = rp[r]; p < rp[r + 1]; ++p) {
                if (cf[p] == 1.0 || cf[p] == -1.0) {
                    // the kernel adds unit references before the general one;
                    // keep the reference's summation order
                    if (seen_general) bad = true;
                    ++nu;
                } else {
                    seen_general = true;
                    ++ng;
                }
            }
            if (nu > 2 || ng > 1) bad = true;
            if (ng) general = true;
            wu = std::max(wu, nu);
        }
        if (bad) return nullptr;
        io->wu = wu;
        io->has_general = general;
        std::vector<uint32_t> uidx((size_t)wu * 9 * npad, NONE32);
        std::vector<uint32_t> gidx;
        std::vector<double> gcoef;
        if (general) {
            gidx.assign((size_t)9 * npad, NONE32);
            gcoef.assign((size_t)9 * npad, 0.0);
        }
#pragma omp parallel for num_threads(::sanm_b200::rt::host_threads())
        for (int64_t e = 0; e < (int64_t)N; ++e) {
            for (int q = 0; q < 9; ++q) {
                const uint64_t r = (uint64_t)e * 9 + q;
                int w = 0;
                for (uint32_t p = rp[r]; p < rp[r + 1]; ++p) {
                    if (cf[p] == 1.0 || cf[p] == -1.0) {
                        uidx[((size_t)w * 9 + q) * npad + e] =
                                (ix[p] << 1) | (cf[p] < 0 ? 1u : 0u);
                        ++w;
                    } else {
                        gidx[(size_t)q * npad + e] = ix[p];
                        gcoef[(size_t)q * npad + e] = cf[p];
                    }
                }
            }
        }
        io->uidx.upload(uidx);
        if (general) {
            io->gidx.upload(gidx);
            io->gcoef.upload(gcoef);
        }
        io->h_uidx = std::move(uidx);
        io->h_gidx = std::move(gidx);
    }

    /* ---------------- assembly */
    const auto& rp = rout.h_rowptr();
    const auto& ix = rout.h_idx();
    const auto& cf = rout.h_coeff();
    const uint64_t nout = io->nout;
    const uint64_t nnz = rp[nout];
    // transpose to per-tet lists (counting sort by tet)
    std::vector<uint32_t> tptr(N + 1, 0);
    for (uint64_t p = 0; p < nnz; ++p) ++tptr[ix[p] / 9 + 1];
    for (uint64_t e = 0; e < N; ++e) {
        if (tptr[e + 1] > 36) return nullptr;
        tptr[e + 1] += tptr[e];
    }
    std::vector<uint32_t> t_row(nnz);
    std::vector<uint8_t> t_q(nnz);
    std::vector<double> t_cf(nnz);
    {
        std::vector<uint32_t> fill(tptr.begin(), tptr.end() - 1);
        for (uint64_t r = 0; r < nout; ++r) {
            for (uint32_t p = rp[r]; p < rp[r + 1]; ++p) {
                const uint32_t e = ix[p] / 9;
                const uint32_t dst = fill[e]++;
                t_row[dst] = (uint32_t)r;
                t_q[dst] = (uint8_t)(ix[p] % 9);
                t_cf[dst] = cf[p];
            }
        }
    }
    // per tet: rows -> (c, n[3]); rows with equal n form one vertex slot
    std::vector<double> nrm((size_t)12 * npad, 0.0);
    std::vector<uint32_t> grow((size_t)12 * npad, NONE32);  // global row of (d, c)
    bool bad = false;
#pragma omp parallel for num_threads(::sanm_b200::rt::host_threads()) reduction(|| : bad)
    for (int64_t e = 0; e < (int64_t)N; ++e) {
        struct RowInfo {
            uint32_t row;
            int c;
            double n[3];
            bool has[3];
        } rows[12];
        int nr = 0;
        bool fail = false;
        for (uint32_t p = tptr[e]; p < tptr[e + 1] && !fail; ++p) {
            const int c = t_q[p] / 3, j = t_q[p] % 3;
            int k = 0;
            while (k < nr && rows[k].row != t_row[p]) ++k;
            if (k == nr) {
                if (nr == 12) {
                    fail = true;
                    break;
                }
                rows[nr] = RowInfo{t_row[p], c, {0, 0, 0}, {false, false, false}};
                ++nr;
            }
            if (rows[k].c != c || rows[k].has[j]) {
                fail = true;
                break;
            }
            rows[k].n[j] = t_cf[p];
            rows[k].has[j] = true;
        }
        int nslot = 0;
        double sn[4][3];
        bool used[4][3];
        for (int k = 0; k < nr && !fail; ++k) {
            int d = 0;
            for (; d < nslot; ++d) {
                if (!used[d][rows[k].c] && sn[d][0] == rows[k].n[0] &&
                    sn[d][1] == rows[k].n[1] && sn[d][2] == rows[k].n[2]) {
                    break;
                }
            }
            if (d == nslot) {
                if (nslot == 4) {
                    fail = true;
                    break;
                }
                for (int j = 0; j < 3; ++j) {
                    sn[d][j] = rows[k].n[j];
                    used[d][j] = false;
                    nrm[((size_t)d * 3 + j) * npad + e] = rows[k].n[j];
                }
                ++nslot;
            }
            used[d][rows[k].c] = true;
            grow[((size_t)d * 3 + rows[k].c) * npad + e] = rows[k].row;
        }
        if (fail) bad = true;
    }
    if (bad) return nullptr;
    t_row = {};
    t_q = {};
    t_cf = {};

    // per block: distinct rows -> local ids, contributor lists
    std::vector<uint32_t> blk_cnt(ntiles + 1, 0);
    std::vector<std::vector<uint32_t>> blk_rows(ntiles);
#pragma omp parallel for num_threads(::sanm_b200::rt::host_threads())
    for (int64_t b = 0; b < (int64_t)ntiles; ++b) {
        auto& v = blk_rows[b];
        for (int s = 0; s < 12; ++s) {
            for (int l = 0; l < TILE; ++l) {
                uint32_t r = grow[(size_t)s * npad + b * TILE + l];
                if (r != NONE32) v.push_back(r);
            }
        }
        std::sort(v.begin(), v.end());
        v.erase(std::unique(v.begin(), v.end()), v.end());
        blk_cnt[b + 1] = (uint32_t)v.size();
    }
    for (size_t b = 0; b < ntiles; ++b) blk_cnt[b + 1] += blk_cnt[b];
    const uint64_t R = blk_cnt[ntiles];
    io->nr_partial = R;
    std::vector<uint16_t> lrow((size_t)12 * npad, NONE16);
    std::vector<uint32_t> con_ptr(R + 1, 0);
    std::vector<uint32_t> part_row(R);
#pragma omp parallel for num_threads(::sanm_b200::rt::host_threads())
    for (int64_t b = 0; b < (int64_t)ntiles; ++b) {
        const auto& v = blk_rows[b];
        const uint32_t base = blk_cnt[b];
        for (size_t k = 0; k < v.size(); ++k) part_row[base + k] = v[k];
        for (int s = 0; s < 12; ++s) {
            for (int l = 0; l < TILE; ++l) {
                const size_t a = (size_t)s * npad + b * TILE + l;
                if (grow[a] == NONE32) continue;
                const uint32_t lr = (uint32_t)(
                        std::lower_bound(v.begin(), v.end(), grow[a]) - v.begin());
                lrow[a] = (uint16_t)lr;
                ++con_ptr[base + lr + 1];
            }
        }
    }
    for (uint64_t g = 0; g < R; ++g) con_ptr[g + 1] += con_ptr[g];
    std::vector<uint16_t> con_idx(con_ptr[R]);
#pragma omp parallel for num_threads(::sanm_b200::rt::host_threads())
    for (int64_t b = 0; b < (int64_t)ntiles; ++b) {
        const uint32_t base = blk_cnt[b];
        const uint32_t nloc = blk_cnt[b + 1] - base;
        std::vector<uint32_t> fill(con_ptr.begin() + base,
                                   con_ptr.begin() + base + nloc);
        // contributors in ascending (lane, slot) order: tets in index order as
        // the reference's reverse list (fea/tetrahedral_mesh.cpp:31-91)
        for (int l = 0; l < TILE; ++l) {
            for (int s = 0; s < 12; ++s) {
                const size_t a = (size_t)s * npad + b * TILE + l;
                if (lrow[a] == NONE16) continue;
                con_idx[fill[lrow[a]]++] = (uint16_t)(s * TILE + l);
            }
        }
    }
    // final reduction lists: partial ids per output row, ascending block order
    std::vector<uint32_t> fin_ptr(nout + 1, 0);
    for (uint64_t g = 0; g < R; ++g) ++fin_ptr[part_row[g] + 1];
    for (uint64_t r = 0; r < nout; ++r) fin_ptr[r + 1] += fin_ptr[r];
    std::vector<uint32_t> fin_idx(R);
    {
        std::vector<uint32_t> fill(fin_ptr.begin(), fin_ptr.end() - 1);
        for (uint64_t g = 0; g < R; ++g) fin_idx[fill[part_row[g]]++] = (uint32_t)g;
    }
    io->row_lo = nout;
    io->row_hi = 0;
    for (uint64_t r = 0; r < nout; ++r) {
        if (fin_ptr[r + 1] > fin_ptr[r]) {
            io->row_lo = std::min(io->row_lo, r);
            io->row_hi = r + 1;
        }
    }
    if (io->row_lo >= io->row_hi) io->row_lo = io->row_hi = 0;
    io->h_grow = std::move(grow);
    io->nrm.upload(nrm);
    io->lrow.upload(lrow);
    io->blk_ptr.upload(blk_cnt);
    io->con_ptr.upload(con_ptr);
    io->con_idx.upload(con_idx);
    io->fin_ptr.upload(fin_ptr);
    io->fin_idx.upload(fin_idx);
    io->partial.alloc(R ? R : 1);
    return io;
}

}  // namespace sanm_b200
