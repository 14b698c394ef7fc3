/* cxx/tests_support.cpp -- pieces that only the reference's own Catch2 cases
 * need on top of the host layer (linked into _build/gpu_tests, not into
 * gpu_fea or libsanm_host.a).
 *
 * tests/symbolic.cpp "Symbolic.Analytic" evaluates ln() on the host through
 * UnaryAnalyticTrait::make_log() (declared in the reference's
 * libsanm/analytic_unary.h, which this file compiles against where it lies)
 * to check what the solver -- the device path -- returned.  Only eval() and
 * eval_derivative() are reached from the tests; Taylor propagation of unary
 * functions lives in the device library (csrc/taylor_dev.h). */
#include "libsanm/analytic_unary.h"

#include <cmath>

namespace sanm {
namespace {
class HostLogTrait final : public UnaryAnalyticTrait {
public:
    void eval(TensorND& dst, const TensorND& src) const override {
        map(dst, src, [](fp_t x) { return std::log(x); });
    }
    void eval_derivative(TensorND& dst, const TensorND& src) const override {
        map(dst, src, [](fp_t x) { return 1.0 / x; });
    }

private:
    template <class F>
    static void map(TensorND& dst, const TensorND& src, F f) {
        TensorND out{src.shape()};
        const fp_t* s = src.ptr();
        fp_t* d = out.woptr();
        for (size_t i = 0, n = src.shape().total_nr_elems(); i < n; ++i) d[i] = f(s[i]);
        dst = out;
    }
    void do_prop_taylor_coeff(TensorND&, const TensorArray&, const TensorArray&,
                              TaylorCoeffUserDataPtr*) const override {
        throw SANMError{"host-side Taylor propagation of unary functions is not "
                        "part of the GPU host layer"};
    }
};
}  // namespace

UnaryAnalyticTraitPtr UnaryAnalyticTrait::make_log() {
    static const HostLogTrait inst;
    return UnaryAnalyticTraitPtr{&inst, [](const UnaryAnalyticTrait*) {}};
}
}  // namespace sanm

/* tests/tensor.cpp "Tensor.SVD" also exercises three HOST tensor routines of
 * libsanm/tensor_svd.h.  Their device counterparts are taylor_dev.h
 * svdw3_fwd_p / svdw3_fwd_full / svdwn_fwd_p (checked against the oracle in
 * tests/test_taylor_parity.py); the host layer does not carry a second, CPU
 * implementation, so that one test case is excluded from gpu_tests
 * (tests/test_reference_suite.py) and these definitions only satisfy the
 * linker. */
#include "libsanm/tensor_svd.h"

namespace sanm {
void svd_w_grad_revmode(StSparseLinearTrans&, const TensorND&, const TensorND&,
                        const TensorND&, const StSparseLinearTrans&,
                        const StSparseLinearTrans&, const StSparseLinearTrans&) {
    throw SANMError{"svd_w_grad_revmode: device-side only in the GPU host layer"};
}
void svd_w_taylor_fwd(TensorND&, TensorND&, TensorND&, const TensorND&,
                      const TensorND&, const TensorND&, const TensorND&,
                      const TensorND&, const TensorND*, const TensorND&) {
    throw SANMError{"svd_w_taylor_fwd: device-side only in the GPU host layer"};
}
void svd_w_taylor_fwd_p(TensorND&, TensorND&, const TensorND&, const TensorND&,
                        const TensorND&, const TensorND&, const TensorND&,
                        const TensorND&, const TensorND&) {
    throw SANMError{"svd_w_taylor_fwd_p: device-side only in the GPU host layer"};
}
}  // namespace sanm
