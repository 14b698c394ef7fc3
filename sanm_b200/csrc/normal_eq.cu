/* sanm_b200/csrc/normal_eq.cu -- see normal_eq.h. */
#include "normal_eq.h"

#include <algorithm>

namespace sanm_b200 {

namespace {

//! (K'K)[i,j] = sum over rows r present in both columns: merge of two sorted
//! row lists, ascending r (fixed order)
struct FormKernel {
    const uint32_t* colptr;
    const uint32_t* rowidx;
    const uint32_t* pos;
    const uint32_t* ri2;
    const uint32_t* ci2;
    const double* kvals;
    double p;
    double* out;
    SANM_D void operator()(size_t e) const {
        const uint32_t i = ri2[e], j = ci2[e];
        uint32_t a = colptr[i], ae = colptr[i + 1];
        uint32_t b = colptr[j], be = colptr[j + 1];
        double acc = 0;
        while (a < ae && b < be) {
            const uint32_t ra = rowidx[a], rb = rowidx[b];
            if (ra == rb) {
                acc += kvals[pos[a]] * kvals[pos[b]];
                ++a;
                ++b;
            } else if (ra < rb) {
                ++a;
            } else {
                ++b;
            }
        }
        out[e] = i == j ? acc + p : acc;
    }
};

struct ApplyTKernel {
    const uint32_t* colptr;
    const uint32_t* rowidx;
    const uint32_t* pos;
    const double* kvals;
    const double* b;
    double* y;
    SANM_D void operator()(size_t i) const {
        double acc = 0;
        for (uint32_t a = colptr[i]; a < colptr[i + 1]; ++a) {
            acc += kvals[pos[a]] * b[rowidx[a]];
        }
        y[i] = acc;
    }
};

}  // namespace

void NormalEquations::analyze(uint64_t n, const std::vector<uint32_t>& rowptr,
                              const std::vector<uint32_t>& colidx) {
    m_n = n;
    const uint32_t nnz = rowptr[n];
    // CSC of K; rows inside a column come out ascending
    std::vector<uint32_t> colptr(n + 1, 0), rowidx(nnz), pos(nnz);
    for (uint32_t p = 0; p < nnz; ++p) ++colptr[colidx[p] + 1];
    for (uint64_t i = 0; i < n; ++i) colptr[i + 1] += colptr[i];
    {
        std::vector<uint32_t> fill(colptr.begin(), colptr.end() - 1);
        for (uint32_t r = 0; r < n; ++r) {
            for (uint32_t p = rowptr[r]; p < rowptr[r + 1]; ++p) {
                const uint32_t q = fill[colidx[p]]++;
                rowidx[q] = r;
                pos[q] = p;
            }
        }
    }
    // pattern of K'K: row i = union of the rows of K that have an entry in
    // column i; the diagonal is always present (p I)
    std::vector<std::vector<uint32_t>> rows(n);
#pragma omp parallel for num_threads(::sanm_b200::rt::host_threads()) schedule(dynamic, 64)
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        std::vector<uint32_t>& out = rows[i];
        out.push_back((uint32_t)i);
        for (uint32_t a = colptr[i]; a < colptr[i + 1]; ++a) {
            const uint32_t r = rowidx[a];
            out.insert(out.end(), colidx.begin() + rowptr[r],
                       colidx.begin() + rowptr[r + 1]);
        }
        std::sort(out.begin(), out.end());
        out.erase(std::unique(out.begin(), out.end()), out.end());
    }
    m_rowptr2.assign(n + 1, 0);
    for (uint64_t i = 0; i < n; ++i) {
        m_rowptr2[i + 1] = m_rowptr2[i] + (uint32_t)rows[i].size();
    }
    sanm_check((uint64_t)m_rowptr2[n] < (1ull << 31), "K'K pattern too large");
    m_colidx2.resize(m_rowptr2[n]);
    std::vector<uint32_t> ri2(m_rowptr2[n]);
    for (uint64_t i = 0; i < n; ++i) {
        std::copy(rows[i].begin(), rows[i].end(), m_colidx2.begin() + m_rowptr2[i]);
        std::fill(ri2.begin() + m_rowptr2[i], ri2.begin() + m_rowptr2[i + 1],
                  (uint32_t)i);
    }
    m_colptr.upload(colptr);
    m_rowidx.upload(rowidx);
    m_pos.upload(pos);
    m_ri2.upload(ri2);
    m_ci2.upload(m_colidx2);
}

void NormalEquations::form(const double* kvals, double p, double* out) const {
    rt::for_each(nnz2(), FormKernel{m_colptr.ptr(), m_rowidx.ptr(), m_pos.ptr(),
                                    m_ri2.ptr(), m_ci2.ptr(), kvals, p, out});
}

void NormalEquations::apply_transposed(const double* kvals, const double* b,
                                       double* y) const {
    rt::for_each(m_n, ApplyTKernel{m_colptr.ptr(), m_rowidx.ptr(), m_pos.ptr(),
                                   kvals, b, y});
}

}  // namespace sanm_b200
