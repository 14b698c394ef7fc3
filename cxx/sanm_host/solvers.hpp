// cxx/sanm_host/solvers.hpp -- sparse maps, continuation drivers, sparse
// direct solver, Pade approximant and scalar polynomial tools of the
// libsanm-compatible host layer.  Interfaces mirrored:
//   libsanm/anm.h:24-85   SparseLinearDesc / SparseLinearDescCompressed
//   libsanm/anm.h:98-305  ANMDriverHelper, ANMSolverVecScale, ANMEqnSolver,
//                         ANMImplicitSolver
//   libsanm/sparse_solver.h:17-87  SparseSolver
//   libsanm/pade.h:20-58  PadeApproximation
//   libsanm/unary_polynomial.h:17-57
//
// The reference's drivers run the order loop on the host and call the
// propagator / sparse solver once per order.  Here a driver object owns a
// `sanm_anm_t`: the whole expansion (gather, Taylor propagation, assembly,
// factorisation, back-solves, scalar recurrences) stays on the device and the
// host classes only keep shapes and fetch results on request.
#pragma once

#include "sanm_host/graph.hpp"

#include <complex>

struct sanm_remap;
struct sanm_sparse;
struct sanm_anm;
struct sanm_pade;

namespace sanm {

class SparseSolver;
class SparseLinearDesc;
using SparseLinearDescPtr = std::shared_ptr<SparseLinearDesc>;

namespace capi {
//! a `sanm_remap_t` with shared ownership
struct RemapHandle : NonCopyable {
    sanm_remap* h = nullptr;
    explicit RemapHandle(sanm_remap* p) : h(p) {}
    ~RemapHandle();
};
}  // namespace capi

//! user-implemented sparse linear map between flattened tensors
class SparseLinearDesc {
public:
    struct InputElem {
        fp_t coeff;
        size_t idx;
        InputElem() = default;
        InputElem(fp_t coeff_, size_t idx_) : coeff{coeff_}, idx{idx_} {}
    };
    using Linear1d = std::span<const InputElem>;

    virtual ~SparseLinearDesc() = default;
    virtual TensorShape out_shape() const = 0;
    virtual TensorShape inp_shape() const = 0;
    virtual void init_multi_thread(size_t nr_thread);
    virtual Linear1d get(size_t dst_index, size_t thread_id) const = 0;

    //! y = map(x), computed on the device from the flattened form
    TensorND apply(const TensorND& x) const;

    static SparseLinearDescPtr make_identity(
            const TensorShape& inp_shape,
            const Maybe<TensorShape>& out_shape = None);

    /*!
     * \brief the map as a device CSR (`sanm_remap_t`)
     *
     * Built once per object by calling get() for every output index -- the
     * same enumeration as the reference's apply() (libsanm/anm.cpp:55-75) --
     * and cached: descriptors are immutable once a solver has seen them.
     */
    sanm_remap* device_remap() const;

protected:
    //! subclasses with a closed form (identity) install their handle directly
    void set_device_remap(sanm_remap* h) const;

private:
    mutable std::shared_ptr<capi::RemapHandle> m_dev;
};

class SparseLinearDescCompressed : public SparseLinearDesc {
protected:
    std::vector<InputElem> m_all_input_elem;
    std::vector<std::pair<leastsize_t, leastsize_t>> m_oidx_input_elem;

public:
    Linear1d get(size_t dst_index, size_t) const override;
};

class PadeApproximation;

class ANMDriverHelper {
public:
    struct HyperParam {
        bool use_pade = false;
        bool sanity_check = true;
        int order = 8;
        fp_t maxr = 1e-6;
        fp_t solution_check_tol = 1e-4;
        fp_t xcoeff_l2_penalty = 0;

        inline HyperParam() noexcept;
    };

    void update_approx();
    fp_t get_t_upper() const;
    fp_t solve_a(fp_t t) const;
    std::pair<TensorND, fp_t> eval(fp_t a) const;
    std::span<const TensorND> xt_coeffs() const;
    size_t get_nr_ieter() const;

    //! the C-ABI solver behind this driver (statistics, profiling)
    sanm_anm* device_handle() const { return m_h; }

protected:
    const HyperParam m_hyper_param;
    symbolic::VarNode* const m_func_def;
    const SparseLinearDescPtr m_remap_inp, m_remap_out;
    const TensorShape m_x_shape;
    const size_t m_nr_unknown;

    ANMDriverHelper(symbolic::VarNode* f, SparseLinearDescPtr remap_inp,
                    SparseLinearDescPtr remap_out, const TensorShape& x_shape,
                    const HyperParam& hyper_param);
    ~ANMDriverHelper();

    static bool verbose_mode();

    sanm_anm* m_h = nullptr;

private:
    mutable TensorArray m_xt_cache;
    mutable size_t m_xt_cache_iter = size_t(-1);
};

//! f(x) + t v = 0
class ANMSolverVecScale : public ANMDriverHelper {
public:
    ANMSolverVecScale(symbolic::VarNode* f, SparseLinearDescPtr remap_inp,
                      SparseLinearDescPtr remap_out, TensorND x0, fp_t t0,
                      TensorND v, const HyperParam& hyper_param = {});
    virtual ~ANMSolverVecScale() = default;

protected:
    struct DeferCreate {};
    //! for ANMEqnSolver, which creates a different device object
    ANMSolverVecScale(DeferCreate, symbolic::VarNode* f,
                      SparseLinearDescPtr remap_inp,
                      SparseLinearDescPtr remap_out, const TensorShape& x_shape,
                      const HyperParam& hyper_param);
};

//! f(x) + y = 0
class ANMEqnSolver final : private ANMSolverVecScale {
public:
    struct HyperParam : public ANMSolverVecScale::HyperParam {
        fp_t converge_rms = 1e-5;
        inline HyperParam() noexcept;
    };
    ANMEqnSolver(symbolic::VarNode* f, SparseLinearDescPtr remap_inp,
                 SparseLinearDescPtr remap_out, TensorND x0, TensorND y,
                 const HyperParam& hyper_param = {});

    fp_t residual_rms() const;
    bool converged() const;
    ANMEqnSolver& next_iter();
    TensorND get_x() const;
    using ANMSolverVecScale::device_handle;
    using ANMSolverVecScale::get_nr_ieter;
};

//! F(x, t) = F(x0, t0)
class ANMImplicitSolver final : public ANMDriverHelper {
public:
    ANMImplicitSolver(symbolic::VarNode* f, SparseLinearDescPtr remap_inp,
                      SparseLinearDescPtr remap_out, const TensorND& x0,
                      fp_t t0, const HyperParam& hyper_param = {});
    const TensorND& fx0() const;

private:
    mutable TensorND m_fx0;
    mutable size_t m_fx0_iter = size_t(-1);
};

ANMDriverHelper::HyperParam::HyperParam() noexcept = default;
ANMEqnSolver::HyperParam::HyperParam() noexcept = default;

class SparseSolver {
    class Impl;
    std::unique_ptr<Impl> m_pimpl;

public:
    explicit SparseSolver(size_t nr_xs);
    ~SparseSolver();

    static void set_num_threads(int nr);
    static int get_num_threads();

    class SparseMatBuilder {
    protected:
        ~SparseMatBuilder() = default;

    public:
        virtual void add_constraint(size_t cidx, size_t xidx, fp_t coeff) = 0;
    };

    static void set_verbosity(int verbosity);
    SparseMatBuilder* make_builder(size_t cidx_offset = 0);
    SparseSolver& prepare(fp_t l2_penalty = 0);
    TensorND solve(const TensorND& b) const;
    void dump(const TensorND& b, FILE* fout = stdout) const;
    TensorND apply(const TensorND& x) const;
    fp_t coeff_l2() const;
};

class PadeApproximation {
public:
    explicit PadeApproximation(std::span<const TensorND> xs, bool anm_cond,
                               bool sanity_check);
    PadeApproximation(const PadeApproximation&) = delete;
    PadeApproximation& operator=(const PadeApproximation&) = delete;
    ~PadeApproximation();

    bool estimate_valid_range(fp_t start, fp_t eps, fp_t limit = 0);
    fp_t get_t_max() const;
    fp_t get_t_max_a() const;
    fp_t solve_a(fp_t t) const;
    std::pair<TensorND, fp_t> eval(fp_t a) const;
    TensorND eval_xt(fp_t a) const;

private:
    sanm_pade* m_h = nullptr;
    size_t m_len = 0;
    std::vector<double> m_rows;  // (K+1) x len, kept alive for the handle
};

namespace unary_polynomial {
using coeff_t = std::span<const fp_t>;
fp_t eval(coeff_t f, fp_t x);
fp_t solve_quad(fp_t a, fp_t b, fp_t c);
fp_t solve_eqn(coeff_t f, fp_t xmin, fp_t xmax, fp_t b = 0, fp_t eps = 1e-6);
fp_t stable_x_range(int order);
Maybe<std::vector<std::complex<fp_t>>> roots(coeff_t f, bool only_real,
                                             int max_iter = 300,
                                             fp_t tol = 1e-8);
TensorND eval_tensor(std::span<const TensorND> f, fp_t x);
}  // namespace unary_polynomial

}  // namespace sanm
