/* sanm_b200/csrc/remap.cu -- see remap.h. */
#include "remap.h"

namespace sanm_b200 {

namespace {
struct ApplyKernel {
    const uint32_t* rowptr;
    const uint32_t* idx;
    const double* coeff;
    const double* x;
    double* y;
    VecLayout xl, yl;
    uint64_t nout;
    int ye;
    SANM_D void operator()(size_t i) const {
        // enumerate outputs in their storage order so that stores coalesce
        size_t row = i;
        if (ye) {
            size_t n = nout / ye;
            size_t c = i / n, t = i % n;
            row = t * ye + c;
        }
        double s = 0;
        for (uint32_t p = rowptr[row], pe = rowptr[row + 1]; p < pe; ++p) {
            s += x[xl.addr(idx[p])] * coeff[p];
        }
        y[yl.addr(row)] = s;
    }
};
}  // namespace

Remap::Remap(const Shape& out_shape, const Shape& in_shape,
             const uint64_t* rowptr, const uint64_t* idx, const double* coeff)
        : m_out{out_shape}, m_in{in_shape} {
    const uint64_t n = out_shape.total(), nin = in_shape.total();
    sanm_check(n > 0 && nin > 0);
    const uint64_t nnz = rowptr[n];
    sanm_check(nnz < (1ull << 32) && nin < (1ull << 32),
               "remap too large for 32-bit indices");
    m_h_rowptr.resize(n + 1);
    for (uint64_t i = 0; i <= n; ++i) {
        sanm_check(i == 0 || rowptr[i] >= rowptr[i - 1], "rowptr not monotone");
        m_h_rowptr[i] = (uint32_t)rowptr[i];
    }
    m_h_idx.resize(nnz);
    for (uint64_t i = 0; i < nnz; ++i) {
        sanm_check(idx[i] < nin, "remap index %zu out of range %zu",
                   (size_t)idx[i], (size_t)nin);
        m_h_idx[i] = (uint32_t)idx[i];
    }
    m_h_coeff.assign(coeff, coeff + nnz);
    m_hash_struct = hash_vec(m_h_idx, hash_vec(m_h_rowptr, n * 31 + nin));
    m_hash_full = hash_vec(m_h_coeff, m_hash_struct);
    m_rowptr.upload(m_h_rowptr);
    m_idx.upload(m_h_idx);
    m_coeff.upload(m_h_coeff);
}

std::shared_ptr<Remap> Remap::make_identity(const Shape& in_shape,
                                            const Shape& out_shape) {
    // libsanm/anm.cpp:19-49,77-81
    const uint64_t n = out_shape.total();
    sanm_check(n <= in_shape.total());
    std::vector<uint64_t> rowptr(n + 1), idx(n);
    std::vector<double> coeff(n, 1.0);
    for (uint64_t i = 0; i < n; ++i) {
        rowptr[i] = i;
        idx[i] = i;
    }
    rowptr[n] = n;
    return std::make_shared<Remap>(out_shape, in_shape, rowptr.data(),
                                   idx.data(), coeff.data());
}

void Remap::apply(double* y, VecLayout ylay, const double* x,
                  VecLayout xlay) const {
    rt::for_each(nr_out(), ApplyKernel{m_rowptr.ptr(), m_idx.ptr(),
                                       m_coeff.ptr(), x, y, xlay, ylay,
                                       nr_out(), ylay.e});
}

}  // namespace sanm_b200
