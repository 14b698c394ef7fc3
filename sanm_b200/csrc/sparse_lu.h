/* sanm_b200/csrc/sparse_lu.h -- GPU sparse direct solver behind
 * sanm::SparseSolver (libsanm/sparse_solver.h:17-87).
 *
 * The reference hands a CSR matrix to Intel MKL PARDISO (mtype 11, real
 * non-symmetric; libsanm/sparse_solver.cpp:71-127,154-180), a closed binary
 * that is not part of the checkout.  This is a from-scratch multifrontal
 * solver for structurally symmetric patterns (the pattern is symmetrised with
 * explicit zeros otherwise):
 *
 *   analyse (host, once per sparsity pattern)
 *     supervariable compression -> nested dissection by BFS level-structure
 *     separators -> elimination tree, postorder, column counts -> relaxed
 *     supernodes -> frontal structures, child->parent index maps, level
 *     schedule.
 *   factor (device, once per ANM iteration)
 *     every front F = [F11 F12; F21 F22] is assembled (original entries +
 *     gathered child Schur complements; gather form => deterministic, no
 *     atomics) and reduced by in-place block Gauss-Jordan with partial
 *     pivoting inside the pivot block:
 *         F11 <- G = (P F11)^-1, F12 <- G P F12, F21 <- -F21 G,
 *         F22 <- F22 - F21 G P F12   (Schur complement for the parent).
 *     Fronts of one level of the supernodal elimination forest run
 *     concurrently: small fronts one CTA each in shared memory, large fronts
 *     with blocked panel / rank-32 GEMM kernels.
 *   solve (device, K times per factorisation)
 *     forward sweep up the levels: upd = w2 + F21' w1 ; backward sweep down:
 *     x1 = G w1 - F12' x2.  Both are dense GEMVs streamed once from HBM
 *     (algorithmic bytes = 8 * stored factor entries + vectors).
 */
#pragma once

#include "rt.h"

#include <memory>

namespace sanm_b200 {

struct SparseLUStats {
    uint64_t n = 0, nnz_a = 0;
    uint64_t factor_entries = 0;  //!< stored entries of G, F12', F21' (= nnz L+U)
    uint64_t front_entries = 0;   //!< sum of nr^2 (HBM held by the fronts)
    double flops = 0;             //!< flops of one numeric factorisation
    uint64_t nr_supernodes = 0, nr_levels = 0, max_front = 0;
    //! tiny pivots replaced in the last factor() (static pivot perturbation,
    //! what PARDISO does with iparm[9] = 13, libsanm/sparse_solver.cpp:116);
    //! valid after the first solve() that follows the factorisation
    uint64_t nr_perturbed = 0;
    //! iterative-refinement steps taken by solve() because of them
    //! (pardisoinit default iparm[7] = 0: refine when pivots were perturbed)
    uint64_t nr_refine_steps = 0;
};

class SparseLU {
public:
    SparseLU();
    ~SparseLU();

    /*!
     * symbolic analysis of an n x n CSR pattern (columns sorted, unique).
     * The pattern is copied; values are supplied to factor().
     * node_keys (host, n values or null): ordering hint, see
     * symbolic_analyze() -- the ANM drivers pass x0, whose entries are the
     * vertex coordinates in `fea`.
     */
    void analyze(uint64_t n, const uint32_t* rowptr, const uint32_t* colidx,
                 const double* node_keys = nullptr);
    bool analyzed() const { return m_analyzed; }

    //! numeric factorisation; vals_dev: device array of CSR values (nnz)
    void factor(const double* vals_dev);
    //! x = A^-1 b on the device (b, x: device vectors of length n; may alias).
    //! When the factorisation had to perturb pivots, two steps of iterative
    //! refinement x += A^-1 (b - A x) follow.
    void solve(double* x_dev, const double* b_dev);
    //! y = A x with the CSR values of the last factor()
    void apply(double* y_dev, const double* x_dev) const;
    //! y = A^T x
    void apply_transposed(double* y_dev, const double* x_dev) const;
    //! Frobenius norm of the CSR values (host sync)
    double coeff_l2() const;

    const SparseLUStats& stats() const { return m_stats; }
    uint64_t n() const { return m_n; }
    uint64_t nnz() const { return m_nnz; }
    const uint32_t* d_rowptr() const;
    const uint32_t* d_colidx() const;

    struct Impl;

private:
    void solve_once(double* x_dev, const double* b_dev);
    std::unique_ptr<Impl> m_impl;
    bool m_analyzed = false;
    uint64_t m_n = 0, m_nnz = 0;
    SparseLUStats m_stats;
};

}  // namespace sanm_b200
