/* sanm_b200/csrc/sparse_solver.cu -- see sparse_solver.h. */
#include "sparse_solver.h"

#include <algorithm>
#include <cmath>
#include <numeric>

namespace sanm_b200 {

namespace {
struct SpmvT {
    // y = A x, plain CSR
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

void coo_to_csr(uint64_t n, const std::vector<uint32_t>& rows,
                const std::vector<uint32_t>& cols,
                const std::vector<double>& vals, std::vector<uint32_t>& rowptr,
                std::vector<uint32_t>& colidx, std::vector<double>& a) {
    const size_t nnz = rows.size();
    std::vector<uint32_t> cnt(n + 1, 0);
    for (size_t i = 0; i < nnz; ++i) ++cnt[rows[i] + 1];
    for (uint64_t i = 0; i < n; ++i) cnt[i + 1] += cnt[i];
    std::vector<uint32_t> order(nnz);
    {
        std::vector<uint32_t> pos(cnt.begin(), cnt.end() - 1);
        for (size_t i = 0; i < nnz; ++i) order[pos[rows[i]]++] = (uint32_t)i;
    }
    rowptr.assign(n + 1, 0);
    colidx.clear();
    a.clear();
    colidx.reserve(nnz);
    a.reserve(nnz);
    for (uint64_t r = 0; r < n; ++r) {
        auto b = order.begin() + cnt[r], e = order.begin() + cnt[r + 1];
        // stable: equal columns keep insertion order, so duplicate sums are
        // accumulated in the order the caller added them
        std::stable_sort(b, e, [&](uint32_t x, uint32_t y) { return cols[x] < cols[y]; });
        for (auto it = b; it != e;) {
            uint32_t c = cols[*it];
            double v = 0;
            while (it != e && cols[*it] == c) {
                v += vals[*it];
                ++it;
            }
            colidx.push_back(c);
            a.push_back(v);
        }
        rowptr[r + 1] = (uint32_t)colidx.size();
    }
}

SparseSolver::SparseSolver(uint64_t n) : m_n{n} {
    sanm_check(n > 0 && n < (1ull << 31));
}

void SparseSolver::add_constraints(uint64_t nnz, const uint64_t* cidx,
                                   const uint64_t* xidx, const double* coeff) {
    // libsanm/sparse_solver.cpp:284-305
    sanm_check(!m_prepared, "add_constraint after prepare");
    for (uint64_t i = 0; i < nnz; ++i) {
        sanm_check(std::isfinite(coeff[i]), "coeff[%zu,%zu]=%g", (size_t)cidx[i],
                   (size_t)xidx[i], coeff[i]);
        if (std::fabs(coeff[i]) < 1e-9) continue;
        sanm_check(cidx[i] < m_n && xidx[i] < m_n);
        if (m_any) {
            sanm_check(cidx[i] == m_prev_row || cidx[i] == m_prev_row + 1,
                       "constraints not added in order: prev=%zu cur=%zu",
                       (size_t)m_prev_row, (size_t)cidx[i]);
        }
        m_any = true;
        m_prev_row = cidx[i];
        m_rows.push_back((uint32_t)cidx[i]);
        m_cols.push_back((uint32_t)xidx[i]);
        m_vals.push_back(coeff[i]);
    }
}

void SparseSolver::prepare(double l2_penalty) {
    sanm_check(!m_prepared);
    rt::check_device();
    m_prepared = true;
    m_l2 = l2_penalty;
    coo_to_csr(m_n, m_rows, m_cols, m_vals, m_rowptr, m_colidx, m_a);
    for (uint64_t r = 0; r < m_n; ++r) {
        sanm_check(m_rowptr[r + 1] > m_rowptr[r], "empty row %zu", (size_t)r);
    }
    m_rows.clear();
    m_cols.clear();
    m_vals.clear();
    m_d_rowptr.upload(m_rowptr);
    m_d_colidx.upload(m_colidx);
    m_a_dev.upload(m_a);
    m_dx.alloc(m_n);
    m_db.alloc(m_n);
    if (!l2_penalty) {
        m_lu.analyze(m_n, m_rowptr.data(), m_colidx.data());
        m_lu.factor(m_a_dev.ptr());
        return;
    }
    // Tikhonov: (A'A + p I) x = A'b  (libsanm/sparse_solver.cpp:366-395,163-175)
    std::vector<uint32_t> r2, c2;
    std::vector<double> v2;
    // A'A = sum over rows r of outer(a_r, a_r)
    for (uint64_t r = 0; r < m_n; ++r) {
        for (uint32_t p = m_rowptr[r]; p < m_rowptr[r + 1]; ++p) {
            for (uint32_t q = m_rowptr[r]; q < m_rowptr[r + 1]; ++q) {
                r2.push_back(m_colidx[p]);
                c2.push_back(m_colidx[q]);
                v2.push_back(m_a[p] * m_a[q]);
            }
        }
    }
    for (uint64_t i = 0; i < m_n; ++i) {
        r2.push_back((uint32_t)i);
        c2.push_back((uint32_t)i);
        v2.push_back(l2_penalty);
    }
    std::vector<uint32_t> rp, ci;
    std::vector<double> av;
    coo_to_csr(m_n, r2, c2, v2, rp, ci, av);
    rt::DBuf<double> av_dev;
    av_dev.upload(av);
    m_lu.analyze(m_n, rp.data(), ci.data());
    m_lu.factor(av_dev.ptr());
    rt::sync();
}

void SparseSolver::solve_host(const double* b, double* x) {
    sanm_check(m_prepared, "solve before prepare");
    for (uint64_t i = 0; i < m_n; ++i) {
        sanm_check(std::isfinite(b[i]), "b[%zu]=%g", (size_t)i, b[i]);
    }
    if (m_l2) {
        // A'b on the host (debug-scale path)
        std::vector<double> atb(m_n, 0.0);
        for (uint64_t r = 0; r < m_n; ++r) {
            for (uint32_t p = m_rowptr[r]; p < m_rowptr[r + 1]; ++p) {
                atb[m_colidx[p]] += m_a[p] * b[r];
            }
        }
        m_db.upload(atb.data(), m_n);
    } else {
        m_db.upload(b, m_n);
    }
    m_lu.solve(m_dx.ptr(), m_db.ptr());
    m_dx.download(x, m_n);
}

void SparseSolver::apply_host(const double* x, double* y) {
    sanm_check(m_prepared, "apply before prepare");
    m_dx.upload(x, m_n);
    rt::for_each(m_n, SpmvT{m_d_rowptr.ptr(), m_d_colidx.ptr(), m_a_dev.ptr(),
                            m_dx.ptr(), m_db.ptr()});
    m_db.download(y, m_n);
}

double SparseSolver::coeff_l2() const {
    double s = 0;
    for (double v : m_a) s += v * v;
    return std::sqrt(s);
}

}  // namespace sanm_b200
