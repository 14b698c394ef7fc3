/* sanm_b200/csrc/normal_eq.h -- Tikhonov-regularised normal equations of the
 * tangent system: (K'K + p I) x = K'b, the xcoeff_l2_penalty path of the
 * reference (libsanm/sparse_solver.cpp:366-395 forms A'A with mkl_sparse_syrk
 * and factors it; :163-175 multiplies the right-hand side by A').  The pattern
 * work is done once per topology on the host; values and right-hand sides are
 * formed on the device with fixed summation orders (deterministic). */
#pragma once

#include "rt.h"

#include <vector>

namespace sanm_b200 {

class NormalEquations {
public:
    //! K: n x n CSR pattern (sorted columns)
    void analyze(uint64_t n, const std::vector<uint32_t>& rowptr,
                 const std::vector<uint32_t>& colidx);
    //! CSR pattern of K'K (+ diagonal)
    const std::vector<uint32_t>& rowptr2() const { return m_rowptr2; }
    const std::vector<uint32_t>& colidx2() const { return m_colidx2; }
    uint64_t nnz2() const { return m_colidx2.size(); }
    //! out[nnz2] = K'K + p I from the CSR values of K (device pointers)
    void form(const double* kvals, double p, double* out) const;
    //! y = K' b (device pointers)
    void apply_transposed(const double* kvals, const double* b, double* y) const;

private:
    uint64_t m_n = 0;
    std::vector<uint32_t> m_rowptr2, m_colidx2;
    rt::DBuf<uint32_t> m_colptr, m_rowidx, m_pos;  //!< CSC view of K
    rt::DBuf<uint32_t> m_ri2, m_ci2;               //!< COO of the K'K pattern
};

}  // namespace sanm_b200
