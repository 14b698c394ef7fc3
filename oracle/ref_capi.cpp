/* oracle/ref_capi.cpp -- TEST INFRASTRUCTURE: the oracle.
 *
 * A C API over the UNMODIFIED reference classes (compiled from
 * /root/reference by oracle/Makefile into oracle/_ref/libref_capi.so).  It has
 * the same shape as the product C-ABI in include/sanm_b200.h (prefix `ref_`
 * instead of `sanm_`), so tests/ and bench.py's cpu_baseline / --impl reference
 * legs can drive the reference and the CUDA path with identical inputs.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py may load this library; the
 * product (sanm_b200/) never does.
 *
 * Reference entry points wrapped here:
 *   graph builders            libsanm/oprs.h:14-105, libsanm/oprs.cpp
 *   TaylorCoeffProp           libsanm/symbolic.h:337-383
 *   ParallelTaylorCoeffProp   libsanm/symbolic.h:391-423
 *   eval_unary_func           libsanm/symbolic.h:426
 *   SparseLinearDesc          libsanm/anm.h:24-85
 *   SparseSolver              libsanm/sparse_solver.h:17-87 (Eigen::SparseLU shim)
 *   ANMSolverVecScale / ANMEqnSolver / ANMImplicitSolver   libsanm/anm.h:98-305
 *   PadeApproximation         libsanm/pade.h:20-58
 *   unary_polynomial::roots / solve_eqn   libsanm/unary_polynomial.h:27-47
 * (this file is compiled with -fno-access-control so that the private
 * denominator coefficients of PadeApproximation can be compared as well)
 */
#include "libsanm/anm.h"
#include "libsanm/oprs.h"
#include "libsanm/sparse_solver.h"
#include "libsanm/unary_polynomial.h"

#include <chrono>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

using namespace sanm;
using namespace sanm::symbolic;

namespace sanm_oracle {
void last_prepared_csr(std::vector<int>& ia, std::vector<int>& ja,
                       std::vector<double>& a);
}

namespace {
thread_local std::string g_err;

template <typename Fn>
int guard(Fn&& fn) {
    try {
        fn();
        return 0;
    } catch (SANMNumericalError& e) {
        g_err = e.what();
        return 3;
    } catch (SANMAssertionError& e) {
        g_err = e.what();
        return 2;
    } catch (std::exception& e) {
        g_err = e.what();
        return 1;
    }
}

TensorShape mkshape(int rank, const uint64_t* dims) {
    TensorShape s;
    s.rank = rank;
    for (int i = 0; i < rank; ++i) s.dim[i] = dims[i];
    return s;
}

TensorND mktensor(const double* data, int rank, const uint64_t* dims) {
    TensorND t{mkshape(rank, dims)};
    memcpy(t.woptr(), data, sizeof(double) * t.shape().total_nr_elems());
    return t;
}

void export_tensor(const TensorND& t, double* out) {
    memcpy(out, t.ptr(), sizeof(double) * t.shape().total_nr_elems());
}

struct Graph {
    ComputingGraph cg;
    std::vector<VarNode*> vars;
    int add(VarNode* v) {
        vars.push_back(v);
        return (int)vars.size() - 1;
    }
    SymbolVar sv(int i) { return i < 0 ? SymbolVar{} : SymbolVar{vars.at(i)}; }
};

struct Prop {
    std::unique_ptr<TaylorCoeffProp> p;
    VarNode* in;
    TensorND last_y;
};

class CsrDesc final : public SparseLinearDescCompressed {
    TensorShape m_out, m_in;

public:
    CsrDesc(TensorShape out, TensorShape in, const uint64_t* rowptr,
            const uint64_t* idx, const double* coeff)
            : m_out{out}, m_in{in} {
        size_t n = out.total_nr_elems();
        m_all_input_elem.resize(rowptr[n]);
        for (size_t i = 0; i < rowptr[n]; ++i) {
            m_all_input_elem[i] = {coeff[i], (size_t)idx[i]};
        }
        m_oidx_input_elem.resize(n);
        for (size_t i = 0; i < n; ++i) {
            m_oidx_input_elem[i] = {(leastsize_t)rowptr[i],
                                    (leastsize_t)rowptr[i + 1]};
        }
    }
    TensorShape out_shape() const override { return m_out; }
    TensorShape inp_shape() const override { return m_in; }
};

struct Remap {
    SparseLinearDescPtr d;
};

struct Sparse {
    std::unique_ptr<SparseSolver> s;
    SparseSolver::SparseMatBuilder* b = nullptr;
    size_t n;
};

struct Hyper {
    int use_pade, sanity_check, order;
    double maxr, solution_check_tol, xcoeff_l2_penalty, converge_rms;
};

ANMDriverHelper::HyperParam cvt(const Hyper* h) {
    ANMDriverHelper::HyperParam r;
    r.use_pade = h->use_pade;
    r.sanity_check = h->sanity_check;
    r.order = h->order;
    r.maxr = h->maxr;
    r.solution_check_tol = h->solution_check_tol;
    r.xcoeff_l2_penalty = h->xcoeff_l2_penalty;
    return r;
}

struct Anm {
    int kind;  // 0 vecscale, 1 eqn, 2 implicit
    std::unique_ptr<ANMSolverVecScale> vs;
    std::unique_ptr<ANMEqnSolver> eq;
    std::unique_ptr<ANMImplicitSolver> im;
    ANMDriverHelper* drv() {
        if (kind == 0) return vs.get();
        if (kind == 2) return im.get();
        // private base of ANMEqnSolver (anm.h:241): reachable for the series
        // dump of tests/test_pade_parity.py through a C-style cast
        if (kind == 1) return (ANMSolverVecScale*)eq.get();
        return nullptr;
    }
};
}  // namespace

extern "C" {

const char* ref_last_error() { return g_err.c_str(); }

int ref_set_num_threads(int n) {
    return guard([&] { sanm::set_num_threads(n); });
}

/* ---------------- graph ---------------- */
int ref_graph_create(Graph** out) {
    return guard([&] { *out = new Graph; });
}
int ref_graph_destroy(Graph* g) {
    delete g;
    return 0;
}
int ref_graph_placeholder(Graph* g, int* out) {
    return guard([&] { *out = g->add(placeholder(g->cg).node()); });
}
int ref_graph_constant(Graph* g, const double* data, int rank,
                       const uint64_t* dims, int* out) {
    return guard([&] {
        *out = g->add(constant(g->cg, mktensor(data, rank, dims)).node());
    });
}
int ref_graph_linear_combine(Graph* g, int n, const double* coeffs,
                             const int* vars, double bias, int* out) {
    return guard([&] {
        std::vector<std::pair<fp_t, SymbolVar>> v;
        for (int i = 0; i < n; ++i) v.emplace_back(coeffs[i], g->sv(vars[i]));
        *out = g->add(linear_combine(v, bias).node());
    });
}
int ref_graph_multiply(Graph* g, int a, int b, int* out) {
    return guard([&] { *out = g->add((g->sv(a) * g->sv(b)).node()); });
}
int ref_graph_pow(Graph* g, int x, double e, int* out) {
    return guard([&] { *out = g->add(g->sv(x).pow(e).node()); });
}
int ref_graph_log(Graph* g, int x, int* out) {
    return guard([&] { *out = g->add(g->sv(x).log().node()); });
}
int ref_graph_reduce_sum(Graph* g, int x, int axis, int keepdim, int* out) {
    return guard(
            [&] { *out = g->add(g->sv(x).reduce_sum(axis, keepdim).node()); });
}
int ref_graph_batched_matmul(Graph* g, int a, int b, int* out) {
    return guard(
            [&] { *out = g->add(g->sv(a).batched_matmul(g->sv(b)).node()); });
}
int ref_graph_batched_transpose(Graph* g, int x, int* out) {
    return guard([&] { *out = g->add(g->sv(x).batched_transpose().node()); });
}
int ref_graph_batched_det(Graph* g, int x, int* out) {
    return guard([&] { *out = g->add(g->sv(x).batched_det().node()); });
}
int ref_graph_batched_mul_eye(Graph* g, int x, uint64_t dim, int* out) {
    return guard([&] { *out = g->add(g->sv(x).batched_mul_eye(dim).node()); });
}
int ref_graph_batched_mat_inv_mul(Graph* g, int x, int a, int is_left,
                                  int* out) {
    return guard([&] {
        *out = g->add(batched_mat_inv_mul(g->sv(x), g->sv(a), is_left).node());
    });
}
int ref_graph_batched_svd_w(Graph* g, int x, int require_rotation, int* out3) {
    return guard([&] {
        auto r = g->sv(x).batched_svd_w(require_rotation);
        for (int i = 0; i < 3; ++i) out3[i] = g->add(r[i].node());
    });
}
int ref_graph_slice(Graph* g, int x, int axis, int has_begin, int begin,
                    int has_end, int end, int stride, int* out) {
    return guard([&] {
        Maybe<int> b, e;
        if (has_begin) b = begin;
        if (has_end) e = end;
        *out = g->add(g->sv(x).slice(axis, b, e, stride).node());
    });
}
int ref_graph_concat(Graph* g, int n, const int* vars, int axis, int* out) {
    return guard([&] {
        std::vector<SymbolVar> xs;
        for (int i = 0; i < n; ++i) xs.push_back(g->sv(vars[i]));
        *out = g->add(concat(xs, axis).node());
    });
}

/* ---------------- element sweep (timed CPU baseline) ----------------
 * The element side of ANMDriverHelper::solve_expansion_coeffs
 * (libsanm/anm.cpp:205-290) run with the reference's own
 * ParallelTaylorCoeffProp (libsanm/symbolic.h:391-423; worker threads =
 * sanm::get_num_threads()) and SparseLinearDesc::apply, with x_0..x_K given.
 * x: [(K+1), n_in]; b: [(K+1), n_out]; seconds: wall time of the loop.
 * The shared zero/one constant pools are pre-sized first so that the
 * reference's multi-thread race on TensorStorage::constant<>
 * (libsanm/tensor.cpp:251-293) cannot trigger. */
int ref_sweep_run(Graph* g, int f, Remap* rin, Remap* rout, int order,
                  const double* x, double* b, double* seconds) {
    return guard([&] {
        const size_t nin = rin->d->inp_shape().total_nr_elems(),
                     nout = rout->d->out_shape().total_nr_elems();
        {
            size_t big = rin->d->out_shape().total_nr_elems() * 9;
            TensorND pre{TensorShape{big}};
            pre.fill_with_inplace(0);
            pre.fill_with_inplace(1);
        }
        auto xk = [&](int k) {
            uint64_t d[1] = {nin};
            TensorND t{rin->d->inp_shape()};
            memcpy(t.woptr(), x + (size_t)k * nin, sizeof(double) * nin);
            (void)d;
            return t;
        };
        auto t0 = std::chrono::steady_clock::now();
        // workers attach to the caller's profiler node, which must exist
        // (libsanm/symbolic.cpp:318-323), as it does inside the driver
        SANM_SCOPED_PROFILER("solve_expansion_coeffs");
        ParallelTaylorCoeffProp prop{g->vars.at(f)};
        TensorND y0 = prop.push_xi(rin->d->apply(xk(0))).gather_yi();
        export_tensor(rout->d->apply(y0), b);
        for (int k = 1; k <= order; ++k) {
            TensorND bk = rout->d->apply(prop.compute_next_order_bias());
            if (bk.is_zero() || bk.shape().total_nr_elems() != nout) {
                memset(b + (size_t)k * nout, 0, sizeof(double) * nout);
            } else {
                export_tensor(bk, b + (size_t)k * nout);
            }
            if (k < order) prop.push_xi(rin->d->apply(xk(k)));
        }
        *seconds = std::chrono::duration<double>(
                           std::chrono::steady_clock::now() - t0)
                           .count();
    });
}

/* ---------------- Taylor propagation ---------------- */
int ref_prop_create(Graph* g, int out_var, Prop** out) {
    return guard([&] {
        auto p = std::make_unique<Prop>();
        p->p.reset(new TaylorCoeffProp{g->vars.at(out_var), true});
        sanm_assert(p->p->inputs().size() == 1);
        p->in = p->p->inputs()[0];
        *out = p.release();
    });
}
int ref_prop_destroy(Prop* p) {
    delete p;
    return 0;
}
int ref_prop_push_xi(Prop* p, const double* x, int rank, const uint64_t* dims) {
    return guard([&] { p->last_y = p->p->push_xi(mktensor(x, rank, dims)); });
}
int ref_prop_out_shape(Prop* p, int* rank, uint64_t* dims) {
    return guard([&] {
        *rank = p->last_y.rank();
        for (int i = 0; i < *rank; ++i) dims[i] = p->last_y.shape(i);
    });
}
int ref_prop_gather_yi(Prop* p, double* out) {
    return guard([&] { export_tensor(p->last_y, out); });
}
int ref_prop_compute_next_order_bias(Prop* p, double* out) {
    return guard([&] {
        const TensorND& b = p->p->compute_next_order_bias();
        if (out) export_tensor(b, out);
    });
}
// type: 1 ELEMWISE [batch, dim], 2 FULL [batch, odim, idim]
int ref_prop_get_jacobian(Prop* p, int* type, uint64_t* batch, uint64_t* odim,
                          uint64_t* idim, double* out) {
    return guard([&] {
        const StSparseLinearTrans& j = p->p->get_jacobian(p->in);
        *type = j.type();
        *batch = j.batch();
        *odim = j.out_dim();
        *idim = j.inp_dim();
        if (out) export_tensor(j.coeff(), out);
    });
}
int ref_eval_unary_func(Graph* g, int out_var, const double* x, int rank,
                        const uint64_t* dims, int* orank, uint64_t* odims,
                        double* out) {
    return guard([&] {
        TensorND y = eval_unary_func(g->vars.at(out_var),
                                     mktensor(x, rank, dims));
        *orank = y.rank();
        for (int i = 0; i < *orank; ++i) odims[i] = y.shape(i);
        if (out) export_tensor(y, out);
    });
}

/* ---------------- remap ---------------- */
int ref_remap_create(int out_rank, const uint64_t* out_dims, int in_rank,
                     const uint64_t* in_dims, const uint64_t* rowptr,
                     const uint64_t* idx, const double* coeff, Remap** out) {
    return guard([&] {
        auto r = new Remap;
        r->d = std::make_shared<CsrDesc>(mkshape(out_rank, out_dims),
                                         mkshape(in_rank, in_dims), rowptr, idx,
                                         coeff);
        *out = r;
    });
}
int ref_remap_identity(int in_rank, const uint64_t* in_dims, int out_rank,
                       const uint64_t* out_dims, Remap** out) {
    return guard([&] {
        auto r = new Remap;
        r->d = SparseLinearDesc::make_identity(mkshape(in_rank, in_dims),
                                               mkshape(out_rank, out_dims));
        *out = r;
    });
}
int ref_remap_destroy(Remap* r) {
    delete r;
    return 0;
}
int ref_remap_apply(Remap* r, const double* x, double* y) {
    return guard([&] {
        auto s = r->d->inp_shape();
        uint64_t d[3] = {s.dim[0], s.dim[1], s.dim[2]};
        export_tensor(r->d->apply(mktensor(x, s.rank, d)), y);
    });
}

/* ---------------- sparse solver ---------------- */
int ref_sparse_create(uint64_t n, Sparse** out) {
    return guard([&] {
        auto s = new Sparse;
        s->n = n;
        s->s.reset(new SparseSolver{n});
        *out = s;
    });
}
int ref_sparse_destroy(Sparse* s) {
    delete s;
    return 0;
}
// rows must be non-decreasing, starting at row 0 (one builder)
int ref_sparse_add_constraints(Sparse* s, uint64_t nnz, const uint64_t* cidx,
                               const uint64_t* xidx, const double* coeff) {
    return guard([&] {
        if (!s->b) s->b = s->s->make_builder(0);
        for (uint64_t i = 0; i < nnz; ++i) {
            s->b->add_constraint(cidx[i], xidx[i], coeff[i]);
        }
    });
}
int ref_sparse_prepare(Sparse* s, double l2) {
    return guard([&] { s->s->prepare(l2); });
}
int ref_sparse_solve(Sparse* s, const double* b, double* x) {
    return guard([&] {
        uint64_t d[1] = {s->n};
        export_tensor(s->s->solve(mktensor(b, 1, d)), x);
    });
}
int ref_sparse_apply(Sparse* s, const double* x, double* y) {
    return guard([&] {
        uint64_t d[1] = {s->n};
        export_tensor(s->s->apply(mktensor(x, 1, d)), y);
    });
}
int ref_sparse_coeff_l2(Sparse* s, double* out) {
    return guard([&] { *out = s->s->coeff_l2(); });
}
// CSR of the matrix given to the most recent prepare() anywhere in the process
// (two-call protocol: first with NULL arrays to get sizes)
int ref_last_prepared_csr(uint64_t* n, uint64_t* nnz, int* ia, int* ja,
                          double* a) {
    return guard([&] {
        std::vector<int> via, vja;
        std::vector<double> va;
        sanm_oracle::last_prepared_csr(via, vja, va);
        *n = via.size() - 1;
        *nnz = vja.size();
        if (ia) memcpy(ia, via.data(), sizeof(int) * via.size());
        if (ja) memcpy(ja, vja.data(), sizeof(int) * vja.size());
        if (a) memcpy(a, va.data(), sizeof(double) * va.size());
    });
}

/* ---------------- ANM drivers ---------------- */
int ref_anm_vecscale_create(Graph* g, int f, Remap* rin, Remap* rout,
                            const double* x0, int x_rank,
                            const uint64_t* x_dims, double t0, const double* v,
                            int v_rank, const uint64_t* v_dims,
                            const Hyper* hyper, Anm** out) {
    return guard([&] {
        auto a = std::make_unique<Anm>();
        a->kind = 0;
        a->vs.reset(new ANMSolverVecScale{
                g->vars.at(f), rin->d, rout->d, mktensor(x0, x_rank, x_dims),
                t0, mktensor(v, v_rank, v_dims), cvt(hyper)});
        *out = a.release();
    });
}
int ref_anm_eqn_create(Graph* g, int f, Remap* rin, Remap* rout,
                       const double* x0, int x_rank, const uint64_t* x_dims,
                       const double* y, int y_rank, const uint64_t* y_dims,
                       const Hyper* hyper, Anm** out) {
    return guard([&] {
        auto a = std::make_unique<Anm>();
        a->kind = 1;
        ANMEqnSolver::HyperParam hp;
        static_cast<ANMDriverHelper::HyperParam&>(hp) = cvt(hyper);
        hp.converge_rms = hyper->converge_rms;
        a->eq.reset(new ANMEqnSolver{g->vars.at(f), rin->d, rout->d,
                                     mktensor(x0, x_rank, x_dims),
                                     mktensor(y, y_rank, y_dims), hp});
        *out = a.release();
    });
}
int ref_anm_implicit_create(Graph* g, int f, Remap* rin, Remap* rout,
                            const double* x0, uint64_t n, double t0,
                            const Hyper* hyper, Anm** out) {
    return guard([&] {
        auto a = std::make_unique<Anm>();
        a->kind = 2;
        uint64_t d[1] = {n};
        a->im.reset(new ANMImplicitSolver{g->vars.at(f), rin->d, rout->d,
                                          mktensor(x0, 1, d), t0, cvt(hyper)});
        *out = a.release();
    });
}
int ref_anm_destroy(Anm* a) {
    delete a;
    return 0;
}
int ref_anm_update_approx(Anm* a) {
    return guard([&] {
        sanm_assert(a->drv());
        a->drv()->update_approx();
    });
}
int ref_anm_get_t_upper(Anm* a, double* out) {
    return guard([&] {
        sanm_assert(a->drv());
        *out = a->drv()->get_t_upper();
    });
}
int ref_anm_solve_a(Anm* a, double t, double* out) {
    return guard([&] {
        sanm_assert(a->drv());
        *out = a->drv()->solve_a(t);
    });
}
int ref_anm_eval(Anm* a, double av, double* x_out, double* t_out) {
    return guard([&] {
        sanm_assert(a->drv());
        auto r = a->drv()->eval(av);
        export_tensor(r.first, x_out);
        *t_out = r.second;
    });
}
int ref_anm_nr_iter(Anm* a, uint64_t* out) {
    return guard([&] {
        *out = a->kind == 1 ? a->eq->get_nr_ieter() : a->drv()->get_nr_ieter();
    });
}
// number of [x;t] coefficient vectors currently held (order+1, or 1 after an
// early return); eqn solver keeps them private in the reference -> kind 0/2 only
int ref_anm_nr_xt_coeffs(Anm* a, uint64_t* out) {
    return guard([&] {
        sanm_assert(a->drv());
        *out = a->drv()->xt_coeffs().size();
    });
}
int ref_anm_xt_coeff(Anm* a, uint64_t i, double* out) {
    return guard([&] {
        sanm_assert(a->drv());
        export_tensor(a->drv()->xt_coeffs()[i], out);
    });
}
int ref_anm_eqn_next_iter(Anm* a) {
    return guard([&] {
        sanm_assert(a->kind == 1);
        a->eq->next_iter();
    });
}
int ref_anm_eqn_converged(Anm* a, int* out) {
    return guard([&] {
        sanm_assert(a->kind == 1);
        *out = a->eq->converged();
    });
}
int ref_anm_eqn_residual_rms(Anm* a, double* out) {
    return guard([&] {
        sanm_assert(a->kind == 1);
        *out = a->eq->residual_rms();
    });
}
int ref_anm_eqn_get_x(Anm* a, double* out) {
    return guard([&] {
        sanm_assert(a->kind == 1);
        export_tensor(a->eq->get_x(), out);
    });
}
int ref_anm_implicit_fx0(Anm* a, double* out) {
    return guard([&] {
        sanm_assert(a->kind == 2);
        export_tensor(a->im->fx0(), out);
    });
}

/* ---------------- scalar polynomial tools ----------------
 * unary_polynomial::roots (libsanm/unary_polynomial.cpp:154-334) and
 * solve_eqn (:101-108); same signatures as sanm_poly_* */
int ref_poly_roots(const double* f, uint64_t n, int only_real, double* re,
                   double* im, uint64_t* nr_roots, int* converged) {
    return guard([&] {
        auto r = unary_polynomial::roots({f, f + n}, only_real != 0);
        *converged = r.valid();
        *nr_roots = 0;
        if (r.valid()) {
            *nr_roots = r.val().size();
            for (size_t i = 0; i < r.val().size(); ++i) {
                re[i] = r.val()[i].real();
                im[i] = r.val()[i].imag();
            }
        }
    });
}
int ref_poly_solve_eqn(const double* f, uint64_t n, double xmin, double xmax,
                       double b, double eps, double* out) {
    return guard([&] {
        *out = unary_polynomial::solve_eqn({f, f + n}, xmin, xmax, b, eps);
    });
}

/* ---------------- PadeApproximation (libsanm/pade.h:20-58) ----------------
 * xs: row-major [nr_coeff][len]; same signatures as sanm_pade_* */
struct RefPade {
    TensorArray xs;
    std::unique_ptr<PadeApproximation> p;
};
int ref_pade_create(const double* xs, uint64_t nr_coeff, uint64_t len,
                    int anm_cond, int sanity_check, RefPade** out) {
    return guard([&] {
        auto h = std::make_unique<RefPade>();
        for (uint64_t i = 0; i < nr_coeff; ++i) {
            TensorND t{TensorShape{len}};
            memcpy(t.woptr(), xs + i * len, sizeof(double) * len);
            h->xs.emplace_back(std::move(t));
        }
        h->p.reset(new PadeApproximation{h->xs, anm_cond != 0,
                                         sanity_check != 0});
        *out = h.release();
    });
}
int ref_pade_destroy(RefPade* p) {
    delete p;
    return 0;
}
int ref_pade_estimate_valid_range(RefPade* p, double start, double eps,
                                  double limit, int* accepted) {
    return guard(
            [&] { *accepted = p->p->estimate_valid_range(start, eps, limit); });
}
int ref_pade_get_t_max(RefPade* p, double* t_max, double* t_max_a) {
    return guard([&] {
        *t_max = p->p->get_t_max();
        *t_max_a = p->p->get_t_max_a();
    });
}
int ref_pade_solve_a(RefPade* p, double t, double* out) {
    return guard([&] { *out = p->p->solve_a(t); });
}
int ref_pade_eval_xt(RefPade* p, double a, double* out) {
    return guard([&] { export_tensor(p->p->eval_xt(a), out); });
}
/* denominator coefficients d_0..d_{n-1} (m_d) and the lower-order set
 * (m_d_lo); *nr = 0 when the constructor rejected the series */
int ref_pade_get_denominator(RefPade* p, int lower_order, double* d,
                             uint64_t* nr) {
    return guard([&] {
        const std::vector<fp_t>& v = lower_order ? p->p->m_d_lo : p->p->m_d;
        *nr = v.size();
        for (size_t i = 0; i < v.size(); ++i) d[i] = v[i];
    });
}

}  // extern "C"
