/* sanm_b200/csrc/capi_sparse.cpp -- C ABI: SparseSolver entry points. */
#include "capi_util.h"
#include "sparse_solver.h"

using namespace sanm_b200;

struct sanm_sparse {
    std::unique_ptr<SparseSolver> s;
};

extern "C" {

int sanm_sparse_create(uint64_t n, sanm_sparse_t** out) {
    return capi::guard(
            [&] { *out = new sanm_sparse{std::make_unique<SparseSolver>(n)}; });
}
int sanm_sparse_destroy(sanm_sparse_t* s) {
    delete s;
    return 0;
}
int sanm_sparse_add_constraints(sanm_sparse_t* s, uint64_t nnz,
                                const uint64_t* cidx, const uint64_t* xidx,
                                const double* coeff) {
    return capi::guard([&] { s->s->add_constraints(nnz, cidx, xidx, coeff); });
}
int sanm_sparse_prepare(sanm_sparse_t* s, double l2) {
    return capi::guard([&] { s->s->prepare(l2); });
}
int sanm_sparse_solve(sanm_sparse_t* s, const double* b, double* x) {
    return capi::guard([&] { s->s->solve_host(b, x); });
}
int sanm_sparse_apply(sanm_sparse_t* s, const double* x, double* y) {
    return capi::guard([&] { s->s->apply_host(x, y); });
}
int sanm_sparse_coeff_l2(sanm_sparse_t* s, double* out) {
    return capi::guard([&] { *out = s->s->coeff_l2(); });
}
int sanm_sparse_stats(sanm_sparse_t* s, uint64_t* nnz_lu, double* flops,
                      uint64_t* nr_supernodes, uint64_t* nr_levels) {
    return capi::guard([&] {
        const auto& st = s->s->stats();
        *nnz_lu = st.factor_entries;
        *flops = st.flops;
        *nr_supernodes = st.nr_supernodes;
        *nr_levels = st.nr_levels;
    });
}

int sanm_sparse_robustness(sanm_sparse_t* s, uint64_t* nr_perturbed_pivots,
                           uint64_t* nr_refine_steps) {
    return capi::guard([&] {
        const auto& st = s->s->stats();
        *nr_perturbed_pivots = st.nr_perturbed;
        *nr_refine_steps = st.nr_refine_steps;
    });
}

}  // extern "C"
