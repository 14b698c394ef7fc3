/* sanm_b200/csrc/sparse_solver.h -- sanm::SparseSolver semantics
 * (libsanm/sparse_solver.h:17-87) on top of SparseLU: COO builder with the
 * reference's row-order / duplicate-merge / small-entry-drop rules
 * (libsanm/sparse_solver.cpp:229-325), prepare (with the Tikhonov path of
 * :366-395), solve, apply, coeff_l2.
 */
#pragma once

#include "sparse_lu.h"

namespace sanm_b200 {

class SparseSolver {
public:
    explicit SparseSolver(uint64_t n);
    void add_constraints(uint64_t nnz, const uint64_t* cidx,
                         const uint64_t* xidx, const double* coeff);
    void prepare(double l2_penalty);
    void solve_host(const double* b, double* x);
    void apply_host(const double* x, double* y);
    double coeff_l2() const;
    const SparseLUStats& stats() const { return m_lu.stats(); }
    uint64_t n() const { return m_n; }

private:
    uint64_t m_n;
    bool m_prepared = false;
    double m_l2 = 0;
    uint64_t m_prev_row = 0;
    bool m_any = false;
    std::vector<uint32_t> m_rows, m_cols;
    std::vector<double> m_vals;
    // CSR of A (and of A^T when the l2 path is used)
    std::vector<uint32_t> m_rowptr, m_colidx;
    std::vector<double> m_a;
    SparseLU m_lu;      //!< factor of A, or of A'A + l2 I
    rt::DBuf<double> m_dx, m_db, m_a_dev;
    rt::DBuf<uint32_t> m_d_rowptr, m_d_colidx;
};

//! sort + merge duplicates of a COO list into CSR (n rows)
void coo_to_csr(uint64_t n, const std::vector<uint32_t>& rows,
                const std::vector<uint32_t>& cols,
                const std::vector<double>& vals, std::vector<uint32_t>& rowptr,
                std::vector<uint32_t>& colidx, std::vector<double>& a);

}  // namespace sanm_b200
