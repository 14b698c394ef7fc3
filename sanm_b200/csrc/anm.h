/* sanm_b200/csrc/anm.h -- continuation drivers: ANMSolverVecScale,
 * ANMEqnSolver and ANMImplicitSolver of the reference (libsanm/anm.h:87-305,
 * libsanm/anm.cpp:93-615) with the whole order loop resident on the device.
 *
 * Per ANM iteration (libsanm/anm.cpp:193-312) every [n]-sized or [N,..]-sized
 * operation is a kernel on the library stream and every scalar of the
 * recurrence (t_1, t_k, x_gt . x_1, the per-order sanity residuals) lives in
 * device memory, so the K orders are enqueued without a single host
 * synchronisation.  The K+1 series vectors are copied to the host once at the
 * end, where range estimation, the Pade acceptance test and root finding run
 * in the reference's own scalar arithmetic (pade.cpp, poly.cpp).
 */
#pragma once

#include "graph.h"
#include "pade.h"
#include "remap.h"
#include "normal_eq.h"
#include "sparse_lu.h"
#include "taylor.h"

#include <memory>

namespace sanm_b200 {

struct AnmHyper {
    // defaults of libsanm/anm.h:100-114, 244-248
    bool use_pade = false;
    bool sanity_check = true;
    int order = 8;
    double maxr = 1e-6;
    double solution_check_tol = 1e-4;
    double xcoeff_l2_penalty = 0;
    double converge_rms = 1e-5;
};

//! K = R_out blockdiag(J_e) R_in assembled into a fixed CSR pattern
//! (libsanm/anm.cpp:362-438, 520-608)
class KAssembler {
public:
    /*!
     * \param n number of unknowns (columns >= n of remap_in, i.e. the extra
     *      unknown t of the implicit solver, are routed to grad_t)
     */
    KAssembler(std::shared_ptr<const Remap> rin,
               std::shared_ptr<const Remap> rout, uint64_t n, uint64_t batch,
               int odim, int idim, bool elemwise,
               std::shared_ptr<const Remap> asm_rin = nullptr,
               std::shared_ptr<const Remap> asm_rout = nullptr);
    const std::vector<uint32_t>& rowptr() const { return m_pat->rowptr; }
    const std::vector<uint32_t>& colidx() const { return m_pat->colidx; }
    uint64_t nnz() const { return m_pat->colidx.size(); }
    //! content hash of the pattern (key of the caches that depend on it)
    uint64_t pattern_hash() const { return m_pat->hash; }

    //! structural pattern of K; shared between solver instances (cache.h)
    struct Pattern {
        std::vector<uint32_t> rowptr, colidx;
        rt::DBuf<uint32_t> d_rowptr, d_colidx;
        uint64_t hash = 0;
        //! sizes and strided samples of the two maps the pattern was built
        //! from: a cache hit is accepted only if they agree (a 64-bit hash
        //! alone is not proof)
        std::vector<uint64_t> witness;
    };
    //! vals (device, nnz) and grad_t (device, n; may be null) from the
    //! per-item Jacobian (device, [N][odim][idim] or [N][dim]; with
    //! jac_pitch != 0 the FULL form is stored [odim*idim][jac_pitch])
    void assemble(double* vals, double* grad_t, const double* jac,
                  int64_t jac_pitch = 0) const;

private:
    std::shared_ptr<const Remap> m_rin, m_rout;  //!< define the pattern
    //! the maps the numeric assembly runs over (a rank's shard of the items,
    //! local item numbering; default: the same maps)
    std::shared_ptr<const Remap> m_asm_rin, m_asm_rout;
    uint64_t m_n, m_batch;
    int m_odim, m_idim;
    bool m_elemwise;
    std::shared_ptr<const Pattern> m_pat;
};

//! element-wise K assembly for the mesh remaps of fea (k_assemble.cu)
class FusedKAssembly {
public:
    //! nullptr when the structure does not fit (a tet reading more than 13
    //! distinct unknowns, 32-bit index overflow)
    static std::unique_ptr<FusedKAssembly> build(
            const FusedIO& io, uint64_t n, const std::vector<uint32_t>& k_rowptr,
            const std::vector<uint32_t>& k_colidx);
    //! the host half of build(): pure CPU work, no CUDA or profiler calls, so
    //! that it can run on a helper thread beside the symbolic LU analysis
    //! (both only need the K pattern); finish() uploads on the caller's thread
    struct HostLists {
        bool ok = false;
        uint64_t n_tets = 0, nnz = 0, n = 0;
        int64_t npad = 0;
        std::vector<uint8_t> tab;
        std::vector<uint32_t> kptr, ksrc;
    };
    static HostLists build_host(const FusedIO& io, uint64_t n,
                                const std::vector<uint32_t>& k_rowptr,
                                const std::vector<uint32_t>& k_colidx);
    static std::unique_ptr<FusedKAssembly> finish(HostLists&& h);
    //! jac: [81][jac_pitch] (TaylorProp::ensure_jacobian of a fused graph)
    void assemble(double* vals, double* grad_t, const double* jac,
                  int64_t jac_pitch, const FusedIO& io) const;
    //! the sizes this object was built for (checked on a cache hit)
    bool matches(uint64_t n_tets, uint64_t n, uint64_t nnz) const {
        return m_n_tets == n_tets && m_n == n && m_nnz == nnz;
    }

private:
    FusedKAssembly() = default;
    uint64_t m_n_tets = 0, m_nnz = 0, m_n = 0;
    int64_t m_npad = 0;
    rt::DBuf<uint8_t> m_tab;
    rt::DBuf<uint32_t> m_kptr, m_ksrc;
    rt::DBuf<double> m_cbuf;
};

class AnmSolver {
public:
    enum Kind { VECSCALE = 0, EQN = 1, IMPLICIT = 2 };

    /*!
     * \param v VECSCALE: the vector v; EQN: the vector y; IMPLICIT: unused
     */
    AnmSolver(Kind kind, std::shared_ptr<const Graph> graph, int f,
              std::shared_ptr<const Remap> remap_inp,
              std::shared_ptr<const Remap> remap_out, const double* x0,
              const Shape& x_shape, double t0, const double* v,
              const Shape& v_shape, const AnmHyper& hyper);
    ~AnmSolver();

    Kind kind() const { return m_kind; }
    uint64_t nr_unknown() const { return m_n; }

    // ANMDriverHelper, libsanm/anm.h:116-139
    void update_approx();
    double get_t_upper() const { return m_t_max; }
    double solve_a(double t) const;
    void eval(double a, double* x_out, double* t_out) const;
    size_t nr_iter() const { return m_iter; }
    size_t nr_xt_coeffs() const { return m_nr_coeffs; }
    const double* xt_coeff(size_t i) const;

    // ANMEqnSolver, libsanm/anm.h:253-263
    void next_iter();
    bool converged() const { return m_converged; }
    double residual_rms() const { return m_residual_rms; }
    void get_x(double* out) const;

    // ANMImplicitSolver, libsanm/anm.h:285
    void fx0(double* out) const;

    struct Stats {
        double tet_order_steps = 0;
    };
    const Stats& stats() const { return m_stats; }
    const SparseLUStats& lu_stats() const { return m_lu.stats(); }
    //! continuation steps retried from the plain series range after a Pade
    //! jump failed the consistency check (only with SANM_PADE_RECOVER=1; the
    //! reference aborts, libsanm/anm.cpp:509-518)
    size_t nr_pade_recovered() const { return m_nr_pade_rejected; }
    //! the tangent matrix K of the last expansion as assembled on the device
    //! (what the reference hands to SparseSolver::prepare, libsanm/anm.cpp:
    //! 362-438 / 520-608); rowptr has n + 1 entries.  Pointers may be null to
    //! query the sizes only.
    void k_csr(uint64_t* n, uint64_t* nnz, uint32_t* rowptr, uint32_t* colidx,
               double* vals) const;

private:
    const Kind m_kind;
    const AnmHyper m_hyper;
    std::shared_ptr<const Graph> m_graph;
    std::shared_ptr<const Remap> m_rin, m_rout;
    //! element-phase view: this rank's shard when running on several GPUs
    std::shared_ptr<const Graph> m_graph_e;
    std::shared_ptr<const Remap> m_rin_e, m_rout_e;
    bool m_sharded = false;
    const Shape m_x_shape;
    const uint64_t m_n;   //!< unknowns without t
    const uint64_t m_ld;  //!< pitch of one series vector (>= n + 1)
    const double m_max_a_bound;

    std::unique_ptr<TaylorProp> m_prop;
    std::unique_ptr<KAssembler> m_kasm;
    std::shared_ptr<const FusedKAssembly> m_kfused;
    SparseLU m_lu;

    // host state
    std::vector<double> m_xt0;          //!< [n + 1]
    //! (order+1) x ld series, host copy: filled on demand (xt_coeff(), verbose
    //! trace, recovery), the series itself lives in m_d_xt
    mutable std::vector<double> m_h_xt;
    mutable bool m_h_xt_valid = false;
    bool m_series_on_device = false;
    std::vector<double> m_norm2;        //!< squared norms of the series vectors
    std::vector<double> m_t_coeffs;
    std::vector<double> m_v_host;       //!< VECSCALE: v ; EQN: y
    std::vector<double> m_fx0_host;     //!< IMPLICIT: f(x0, t0)
    size_t m_nr_coeffs = 0;
    size_t m_iter = 0;
    double m_t_max = 0, m_t_max_a = 0;
    std::unique_ptr<PadeApproximation> m_pade;
    bool m_fx_check_failed = false;
    double m_a_bound_taylor = 0;
    size_t m_nr_pade_rejected = 0;
    bool m_converged = false;
    double m_residual_rms = 0;
    Stats m_stats;

    // device state
    rt::DBuf<double> m_d_xt;   //!< (order+1) x ld
    rt::DBuf<double> m_d_v;    //!< grad_t of the vec-scale forms
    rt::DBuf<double> m_d_y;    //!< EQN: y
    rt::DBuf<double> m_d_grad_t, m_d_bi, m_d_xbi, m_d_xgt, m_d_fx, m_d_tmp;
    rt::DBuf<double> m_d_kvals;
    //! xcoeff_l2_penalty != 0: the factor is of K'K + p I (normal_eq.h)
    std::unique_ptr<NormalEquations> m_neq;
    rt::DBuf<double> m_d_k2vals, m_d_atb;
    //! x = K^-1 b, or the Tikhonov solution when the penalty is set
    void lu_solve(double* x, const double* b);
    rt::DBuf<double> m_d_sc;   //!< device scalars
    rt::DBuf<double> m_d_partial;

    void solve_expansion_coeffs();
    bool on_fx0_computed();
    void estimate_valid_range();
    std::vector<double> eval_xt(double a) const;
    double t0() const { return m_t_coeffs[0]; }
    std::vector<const double*> coeff_ptrs() const;
    void ensure_host_series() const;
    void dot_to(int slot, const double* a, const double* b, uint64_t n);
};

}  // namespace sanm_b200
