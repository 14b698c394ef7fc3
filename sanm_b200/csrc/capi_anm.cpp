/* sanm_b200/csrc/capi_anm.cpp -- C ABI: ANM drivers, profiler, polynomial
 * helpers. */
#include "anm.h"
#include "capi_util.h"
#include "pade.h"
#include "poly.h"
#include "prof.h"

#include <cstring>

using namespace sanm_b200;

struct sanm_anm {
    std::unique_ptr<AnmSolver> a;
};

struct sanm_pade {
    std::vector<double> rows;  // [nr_coeff][len], owned copy of the series
    size_t len = 0;
    std::unique_ptr<PadeApproximation> p;
};

namespace {
AnmHyper cvt(const sanm_anm_hyper_t* h) {
    AnmHyper r;
    if (!h) return r;
    r.use_pade = h->use_pade;
    r.sanity_check = h->sanity_check;
    r.order = h->order;
    r.maxr = h->maxr;
    r.solution_check_tol = h->solution_check_tol;
    r.xcoeff_l2_penalty = h->xcoeff_l2_penalty;
    r.converge_rms = h->converge_rms;
    return r;
}
std::shared_ptr<const Graph> snapshot(sanm_graph_t* g) {
    return std::make_shared<Graph>(*g->g);
}
}  // namespace

extern "C" {

void sanm_anm_hyper_default(sanm_anm_hyper_t* h) {
    AnmHyper d;
    h->use_pade = d.use_pade;
    h->sanity_check = d.sanity_check;
    h->order = d.order;
    h->maxr = d.maxr;
    h->solution_check_tol = d.solution_check_tol;
    h->xcoeff_l2_penalty = d.xcoeff_l2_penalty;
    h->converge_rms = d.converge_rms;
}

int sanm_anm_vecscale_create(sanm_graph_t* g, int f, sanm_remap_t* rin,
                             sanm_remap_t* rout, const double* x0, int x_rank,
                             const uint64_t* x_dims, double t0, const double* v,
                             int v_rank, const uint64_t* v_dims,
                             const sanm_anm_hyper_t* hyper, sanm_anm_t** out) {
    return capi::guard([&] {
        *out = new sanm_anm{std::make_unique<AnmSolver>(
                AnmSolver::VECSCALE, snapshot(g), f, rin->r, rout->r, x0,
                Shape{x_rank, x_dims}, t0, v, Shape{v_rank, v_dims},
                cvt(hyper))};
    });
}
int sanm_anm_eqn_create(sanm_graph_t* g, int f, sanm_remap_t* rin,
                        sanm_remap_t* rout, const double* x0, int x_rank,
                        const uint64_t* x_dims, const double* y, int y_rank,
                        const uint64_t* y_dims, const sanm_anm_hyper_t* hyper,
                        sanm_anm_t** out) {
    return capi::guard([&] {
        *out = new sanm_anm{std::make_unique<AnmSolver>(
                AnmSolver::EQN, snapshot(g), f, rin->r, rout->r, x0,
                Shape{x_rank, x_dims}, 0.0, y, Shape{y_rank, y_dims},
                cvt(hyper))};
    });
}
int sanm_anm_implicit_create(sanm_graph_t* g, int f, sanm_remap_t* rin,
                             sanm_remap_t* rout, const double* x0, uint64_t n,
                             double t0, const sanm_anm_hyper_t* hyper,
                             sanm_anm_t** out) {
    return capi::guard([&] {
        *out = new sanm_anm{std::make_unique<AnmSolver>(
                AnmSolver::IMPLICIT, snapshot(g), f, rin->r, rout->r, x0,
                Shape{n}, t0, nullptr, Shape{}, cvt(hyper))};
    });
}
int sanm_anm_destroy(sanm_anm_t* a) {
    delete a;
    return 0;
}
int sanm_anm_update_approx(sanm_anm_t* a) {
    return capi::guard([&] {
        sanm_check(a->a->kind() != AnmSolver::EQN);
        a->a->update_approx();
    });
}
int sanm_anm_get_t_upper(sanm_anm_t* a, double* out) {
    return capi::guard([&] { *out = a->a->get_t_upper(); });
}
int sanm_anm_solve_a(sanm_anm_t* a, double t, double* out) {
    return capi::guard([&] { *out = a->a->solve_a(t); });
}
int sanm_anm_eval(sanm_anm_t* a, double av, double* x_out, double* t_out) {
    return capi::guard([&] { a->a->eval(av, x_out, t_out); });
}
int sanm_anm_nr_iter(sanm_anm_t* a, uint64_t* out) {
    return capi::guard([&] { *out = a->a->nr_iter(); });
}
int sanm_anm_nr_xt_coeffs(sanm_anm_t* a, uint64_t* out) {
    return capi::guard([&] { *out = a->a->nr_xt_coeffs(); });
}
int sanm_anm_xt_coeff(sanm_anm_t* a, uint64_t i, double* out) {
    return capi::guard([&] {
        std::memcpy(out, a->a->xt_coeff(i),
                    (a->a->nr_unknown() + 1) * sizeof(double));
    });
}
int sanm_anm_eqn_next_iter(sanm_anm_t* a) {
    return capi::guard([&] { a->a->next_iter(); });
}
int sanm_anm_eqn_converged(sanm_anm_t* a, int* out) {
    return capi::guard([&] {
        sanm_check(a->a->kind() == AnmSolver::EQN);
        *out = a->a->converged();
    });
}
int sanm_anm_eqn_residual_rms(sanm_anm_t* a, double* out) {
    return capi::guard([&] {
        sanm_check(a->a->kind() == AnmSolver::EQN);
        *out = a->a->residual_rms();
    });
}
int sanm_anm_eqn_get_x(sanm_anm_t* a, double* out) {
    return capi::guard([&] {
        sanm_check(a->a->kind() == AnmSolver::EQN);
        a->a->get_x(out);
    });
}
int sanm_anm_implicit_fx0(sanm_anm_t* a, double* out) {
    return capi::guard([&] { a->a->fx0(out); });
}
int sanm_anm_element_stats(sanm_anm_t* a, double* tet_order_steps,
                           double* element_ms, double* solve_ms,
                           double* factor_ms, double* assemble_ms) {
    return capi::guard([&] {
        *tet_order_steps = a->a->stats().tet_order_steps;
        double ms, tot = 0;
        uint64_t n;
        for (const char* tag : {"taylor_next_order", "push_xi", "remap_out"}) {
            prof::get(tag, &ms, &n);
            tot += ms;
        }
        *element_ms = tot;
        prof::get("sparse_solve", solve_ms, &n);
        prof::get("sparse_prep", factor_ms, &n);
        prof::get("build_sparse_coeff", assemble_ms, &n);
    });
}
int sanm_anm_lu_stats(sanm_anm_t* a, uint64_t* n, uint64_t* nnz_a,
                      uint64_t* factor_entries, double* flops,
                      uint64_t* nr_supernodes, uint64_t* nr_levels,
                      uint64_t* max_front) {
    return capi::guard([&] {
        const auto& s = a->a->lu_stats();
        *n = s.n;
        *nnz_a = s.nnz_a;
        *factor_entries = s.factor_entries;
        *flops = s.flops;
        *nr_supernodes = s.nr_supernodes;
        *nr_levels = s.nr_levels;
        *max_front = s.max_front;
    });
}

int sanm_anm_k_csr(sanm_anm_t* a, uint64_t* n, uint64_t* nnz, uint32_t* rowptr,
                   uint32_t* colidx, double* vals) {
    return capi::guard([&] { a->a->k_csr(n, nnz, rowptr, colidx, vals); });
}
int sanm_anm_robustness_stats(sanm_anm_t* a, uint64_t* pade_recovered,
                              uint64_t* lu_perturbed_pivots,
                              uint64_t* lu_refine_steps) {
    return capi::guard([&] {
        *pade_recovered = a->a->nr_pade_recovered();
        *lu_perturbed_pivots = a->a->lu_stats().nr_perturbed;
        *lu_refine_steps = a->a->lu_stats().nr_refine_steps;
    });
}

/* ---------------- profiler ---------------- */
int sanm_prof_dump(char* buf, uint64_t cap) {
    return capi::guard([&] {
        std::string s = prof::dump();
        if (cap) {
            size_t n = std::min<size_t>(cap - 1, s.size());
            std::memcpy(buf, s.data(), n);
            buf[n] = 0;
        }
    });
}
int sanm_prof_reset(void) {
    return capi::guard([] { prof::reset(); });
}
int sanm_prof_get(const char* tag, double* ms, uint64_t* calls) {
    return capi::guard([&] { prof::get(tag, ms, calls); });
}

/* ---------------- polynomial helpers (host) ---------------- */
int sanm_poly_roots(const double* f, uint64_t n, int only_real, double* re,
                    double* im, uint64_t* nr_roots, int* converged) {
    return capi::guard([&] {
        std::vector<std::complex<double>> out;
        *converged = poly::roots(std::vector<double>(f, f + n), only_real, out);
        *nr_roots = out.size();
        for (size_t i = 0; i < out.size(); ++i) {
            re[i] = out[i].real();
            im[i] = out[i].imag();
        }
    });
}
int sanm_poly_solve_eqn(const double* f, uint64_t n, double xmin, double xmax,
                        double b, double eps, double* out) {
    return capi::guard([&] {
        *out = poly::solve_eqn(std::vector<double>(f, f + n), xmin, xmax, b, eps);
    });
}


/* ---------------- Pade approximant (host decisions) ---------------- */
int sanm_pade_create(const double* xs, uint64_t nr_coeff, uint64_t len,
                     int anm_cond, int sanity_check, sanm_pade_t** out) {
    return capi::guard([&] {
        auto h = std::make_unique<sanm_pade>();
        h->rows.assign(xs, xs + nr_coeff * len);
        h->len = len;
        std::vector<const double*> ptrs(nr_coeff);
        for (uint64_t i = 0; i < nr_coeff; ++i)
            ptrs[i] = h->rows.data() + i * len;
        h->p = std::make_unique<PadeApproximation>(std::move(ptrs), len,
                                                   anm_cond != 0,
                                                   sanity_check != 0);
        *out = h.release();
    });
}
int sanm_pade_destroy(sanm_pade_t* p) {
    delete p;
    return 0;
}
int sanm_pade_estimate_valid_range(sanm_pade_t* p, double start, double eps,
                                   double limit, int* accepted) {
    return capi::guard(
            [&] { *accepted = p->p->estimate_valid_range(start, eps, limit); });
}
int sanm_pade_get_t_max(sanm_pade_t* p, double* t_max, double* t_max_a) {
    return capi::guard([&] {
        *t_max = p->p->get_t_max();
        *t_max_a = p->p->get_t_max_a();
    });
}
int sanm_pade_solve_a(sanm_pade_t* p, double t, double* out) {
    return capi::guard([&] { *out = p->p->solve_a(t); });
}
int sanm_pade_get_denominator(sanm_pade_t* p, int lower_order, double* d,
                              uint64_t* nr) {
    return capi::guard([&] {
        const auto& v = p->p->denominator(lower_order != 0);
        *nr = v.size();
        std::memcpy(d, v.data(), v.size() * sizeof(double));
    });
}
int sanm_pade_eval_xt(sanm_pade_t* p, double a, double* out) {
    return capi::guard([&] {
        auto v = p->p->eval_xt(a);
        std::memcpy(out, v.data(), v.size() * sizeof(double));
    });
}

}  // extern "C"
