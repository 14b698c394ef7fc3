/* sanm_b200/csrc/capi_core.cpp -- C ABI (include/sanm_b200.h): library,
 * graph, Taylor propagation and remap entry points. */
#include "capi_util.h"

using namespace sanm_b200;

extern "C" {

const char* sanm_last_error(void) { return capi::last_error().c_str(); }

int sanm_set_device(int device) {
    return capi::guard([&] {
#ifndef SANM_HOSTEMU
        SANM_CUDA_CHECK(cudaSetDevice(device));
#else
        (void)device;
#endif
        rt::check_device();
    });
}

int sanm_set_stream(void* cuda_stream) {
    return capi::guard([&] { rt::set_stream(cuda_stream); });
}

uint64_t sanm_launch_count(void) { return rt::launch_count(); }

int sanm_synchronize(void) {
    return capi::guard([] { rt::sync(); });
}

/* ---------------- graph ---------------- */
int sanm_graph_create(sanm_graph_t** out) {
    return capi::guard([&] { *out = new sanm_graph{std::make_shared<Graph>()}; });
}
int sanm_graph_destroy(sanm_graph_t* g) {
    delete g;
    return 0;
}
int sanm_graph_placeholder(sanm_graph_t* g, int* out) {
    return capi::guard([&] { *out = g->g->placeholder(); });
}
int sanm_graph_constant(sanm_graph_t* g, const double* data, int rank,
                        const uint64_t* dims, int* out) {
    return capi::guard([&] { *out = g->g->constant(data, Shape{rank, dims}); });
}
int sanm_graph_linear_combine(sanm_graph_t* g, int n, const double* coeffs,
                              const int* vars, double bias, int* out) {
    return capi::guard(
            [&] { *out = g->g->linear_combine(n, coeffs, vars, bias); });
}
int sanm_graph_multiply(sanm_graph_t* g, int a, int b, int* out) {
    return capi::guard([&] { *out = g->g->multiply(a, b); });
}
int sanm_graph_pow(sanm_graph_t* g, int x, double e, int* out) {
    return capi::guard([&] { *out = g->g->pow(x, e); });
}
int sanm_graph_log(sanm_graph_t* g, int x, int* out) {
    return capi::guard([&] { *out = g->g->log(x); });
}
int sanm_graph_reduce_sum(sanm_graph_t* g, int x, int axis, int keepdim,
                          int* out) {
    return capi::guard([&] { *out = g->g->reduce_sum(x, axis, keepdim); });
}
int sanm_graph_batched_matmul(sanm_graph_t* g, int a, int b, int* out) {
    return capi::guard([&] { *out = g->g->batched_matmul(a, b); });
}
int sanm_graph_batched_transpose(sanm_graph_t* g, int x, int* out) {
    return capi::guard([&] { *out = g->g->batched_transpose(x); });
}
int sanm_graph_batched_det(sanm_graph_t* g, int x, int* out) {
    return capi::guard([&] { *out = g->g->batched_det(x); });
}
int sanm_graph_batched_mul_eye(sanm_graph_t* g, int x, uint64_t dim, int* out) {
    return capi::guard([&] { *out = g->g->batched_mul_eye(x, dim); });
}
int sanm_graph_batched_mat_inv_mul(sanm_graph_t* g, int x, int a, int is_left,
                                   int* out) {
    return capi::guard(
            [&] { *out = g->g->batched_mat_inv_mul(x, a, is_left); });
}
int sanm_graph_batched_svd_w(sanm_graph_t* g, int x, int require_rotation,
                             int* out3) {
    return capi::guard([&] { g->g->batched_svd_w(x, require_rotation, out3); });
}
int sanm_graph_slice(sanm_graph_t* g, int x, int axis, int has_begin, int begin,
                     int has_end, int end, int stride, int* out) {
    return capi::guard([&] {
        *out = g->g->slice(x, axis, has_begin, begin, has_end, end, stride);
    });
}
int sanm_graph_concat(sanm_graph_t* g, int n, const int* vars, int axis,
                      int* out) {
    return capi::guard([&] { *out = g->g->concat(n, vars, axis); });
}

/* ---------------- Taylor propagation ---------------- */
int sanm_prop_create(sanm_graph_t* g, int out_var, int max_order,
                     sanm_prop_t** out) {
    return capi::guard([&] {
        // the graph is snapshotted: later additions to g do not affect p
        auto snap = std::make_shared<Graph>(*g->g);
        *out = new sanm_prop{
                std::make_unique<TaylorProp>(snap, out_var, max_order)};
    });
}
int sanm_prop_destroy(sanm_prop_t* p) {
    delete p;
    return 0;
}
int sanm_prop_push_xi(sanm_prop_t* p, const double* x, int rank,
                      const uint64_t* dims) {
    return capi::guard([&] { p->p->push_xi_host(x, Shape{rank, dims}); });
}
int sanm_prop_out_shape(sanm_prop_t* p, int* rank, uint64_t* dims) {
    return capi::guard([&] {
        sanm_check(p->p->prepared(), "push_xi has not been called");
        const Shape& s = p->p->out_shape();
        *rank = s.rank;
        for (int i = 0; i < s.rank; ++i) dims[i] = s.dim[i];
    });
}
int sanm_prop_gather_yi(sanm_prop_t* p, double* out) {
    return capi::guard([&] {
        sanm_check(p->p->xi_known(), "gather_yi needs a preceding push_xi");
        p->p->output_to_host(p->p->order(), out);
        p->p->check_device_error();
    });
}
int sanm_prop_compute_next_order_bias(sanm_prop_t* p, double* out) {
    return capi::guard([&] {
        p->p->compute_next_order_bias();
        if (out) p->p->output_to_host(p->p->order(), out);
        p->p->check_device_error();
    });
}
int sanm_prop_get_jacobian(sanm_prop_t* p, int* type, uint64_t* batch,
                           uint64_t* odim, uint64_t* idim, double* out) {
    return capi::guard([&] {
        sanm_check(p->p->prepared(), "push_xi has not been called");
        *type = p->p->jacobian_type();
        *batch = p->p->batch();
        *odim = p->p->out_elems();
        *idim = p->p->in_elems();
        if (out) p->p->jacobian_to_host(out);
    });
}
int sanm_eval_unary_func(sanm_graph_t* g, int out_var, const double* x,
                         int rank, const uint64_t* dims, int* orank,
                         uint64_t* odims, double* out) {
    // libsanm/symbolic.cpp:44-60: an order-0 evaluation
    return capi::guard([&] {
        auto snap = std::make_shared<Graph>(*g->g);
        TaylorProp prop{snap, out_var, 1};
        prop.push_xi_host(x, Shape{rank, dims});
        const Shape& s = prop.out_shape();
        *orank = s.rank;
        for (int i = 0; i < s.rank; ++i) odims[i] = s.dim[i];
        if (out) prop.output_to_host(0, out);
        prop.check_device_error();
    });
}
int sanm_prop_arena_bytes(sanm_prop_t* p, uint64_t* out) {
    return capi::guard([&] { *out = p->p->arena_bytes(); });
}

/* ---------------- remap ---------------- */
int sanm_remap_create(int out_rank, const uint64_t* out_dims, int in_rank,
                      const uint64_t* in_dims, const uint64_t* rowptr,
                      const uint64_t* idx, const double* coeff,
                      sanm_remap_t** out) {
    return capi::guard([&] {
        rt::check_device();
        *out = new sanm_remap{std::make_shared<Remap>(
                Shape{out_rank, out_dims}, Shape{in_rank, in_dims}, rowptr, idx,
                coeff)};
    });
}
int sanm_remap_identity(int in_rank, const uint64_t* in_dims, int out_rank,
                        const uint64_t* out_dims, sanm_remap_t** out) {
    return capi::guard([&] {
        rt::check_device();
        *out = new sanm_remap{Remap::make_identity(Shape{in_rank, in_dims},
                                                   Shape{out_rank, out_dims})};
    });
}
int sanm_remap_destroy(sanm_remap_t* r) {
    delete r;
    return 0;
}
int sanm_remap_apply(sanm_remap_t* r, const double* x, double* y) {
    // libsanm/anm.cpp:55-75
    return capi::guard([&] {
        rt::DBuf<double> dx, dy(r->r->nr_out());
        dx.upload(x, r->r->nr_in());
        r->r->apply(dy.ptr(), VecLayout{}, dx.ptr(), VecLayout{});
        dy.download(y, r->r->nr_out());
    });
}

}  // extern "C"
