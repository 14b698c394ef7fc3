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

/* tests/tensor.cpp "Tensor.SVD" exercises three HOST tensor routines of
 * libsanm/tensor_svd.h: the reverse-mode gradient of SVD-W and the two order-k
 * forward solves on TensorNDs of any size n (the test goes up to 9 x 9).  On
 * the GPU path these computations live on the device (taylor_dev.h
 * svdw3_fwd_full / svdwn_fwd_p, checked against the oracle in
 * tests/test_taylor_parity.py; the Jacobian comes from forward mode).  The
 * plain host versions below exist so that the reference's test case runs; they
 * follow the same formulas as the device functions, written with loops over a
 * runtime n, row-major n x n matrices.  Not linked into gpu_fea. */
#include "libsanm/tensor_svd.h"

#include <vector>

namespace sanm {
namespace {
using Mat = std::vector<fp_t>;

inline fp_t clip_div(fp_t x, fp_t y) {  // libsanm/tensor_svd.cpp:28-31
    return x * y / (y * y + 1e-12);
}
//! C = op(A) op(B)
void mm(fp_t* C, const fp_t* A, const fp_t* B, size_t n, bool ta, bool tb) {
    for (size_t r = 0; r < n; ++r)
        for (size_t c = 0; c < n; ++c) {
            fp_t s = 0;
            for (size_t q = 0; q < n; ++q)
                s += (ta ? A[q * n + r] : A[r * n + q]) * (tb ? B[c * n + q] : B[q * n + c]);
            C[r * n + c] = s;
        }
}
}  // namespace

/* d(outputs)/dM accumulated over the three outputs of M = U diag(S) V^T,
 * W = U V^T (V = W^T U):
 *   dS_i / dM_kl = U_ki V_li
 *   y (antisymmetric, W = U y V^T):  dy_ij / dM_kl = (U_ki V_lj - U_kj V_li) / (s_i + s_j)
 *   x (antisymmetric, dU = U x):     dx_ij / dM_kl = (U_ki V_lj s_j + U_kj V_li s_i) / (s_j^2 - s_i^2)
 * (libsanm/tensor_svd.cpp:147-272). */
void svd_w_grad_revmode(StSparseLinearTrans& grad, const TensorND& mU,
                        const TensorND& mS, const TensorND& mW,
                        const StSparseLinearTrans& mdU,
                        const StSparseLinearTrans& mdS,
                        const StSparseLinearTrans& mdW) {
    const size_t batch = mU.shape(0), n = mU.shape(1), n2 = n * n;
    sanm_assert(n >= 2);
    size_t out_dim = 0;
    auto take = [&](const StSparseLinearTrans& t, TensorND& full, size_t in_dim) {
        if (!t.valid()) return (const fp_t*)nullptr;
        full = t.check_batched(true).as_full();
        sanm_assert(full.shape() == (TensorShape{batch, t.out_dim(), in_dim}));
        sanm_assert(!out_dim || out_dim == t.out_dim());
        out_dim = t.out_dim();
        return full.ptr();
    };
    TensorND fU, fS, fW, mdM;
    const fp_t* pdU = take(mdU, fU, n2);
    const fp_t* pdS = take(mdS, fS, n);
    const fp_t* pdW = take(mdW, fW, n2);
    sanm_assert(out_dim, "no output grad");
    if (grad.valid()) {
        mdM = grad.check_batched(true).as_full();
        sanm_assert(mdM.shape() == (TensorShape{batch, out_dim, n2}));
        grad = {};
    } else {
        mdM.set_shape({batch, out_dim, n2}).fill_with_inplace(0);
    }
    fp_t* pM = mdM.rwptr();
    Mat V(n2), dy(n2 * n2), dx(n2 * n2), g(n2);
    for (size_t b = 0; b < batch; ++b) {
        const fp_t* U = mU.ptr() + b * n2;
        const fp_t* S = mS.ptr() + b * n;
        const fp_t* W = mW.ptr() + b * n2;
        mm(V.data(), W, U, n, true, false);  // V = W^T U
        // dy[(i,j)][(k,l)], dx[(i,j)][(k,l)]
        for (size_t i = 0; i < n; ++i)
            for (size_t j = 0; j < n; ++j)
                for (size_t k = 0; k < n; ++k)
                    for (size_t l = 0; l < n; ++l) {
                        const fp_t cij = U[k * n + i] * V[l * n + j],
                                   cji = U[k * n + j] * V[l * n + i];
                        const size_t at = (i * n + j) * n2 + k * n + l;
                        dy[at] = i == j ? 0 : clip_div(cij - cji, S[i] + S[j]);
                        dx[at] = i == j ? 0
                                        : clip_div(cij * S[j] + cji * S[i],
                                                   S[j] * S[j] - S[i] * S[i]);
                    }
        for (size_t o = 0; o < out_dim; ++o) {
            fp_t* out = pM + (b * out_dim + o) * n2;
            if (pdS) {
                const fp_t* ds = pdS + (b * out_dim + o) * n;
                for (size_t k = 0; k < n; ++k)
                    for (size_t l = 0; l < n; ++l) {
                        fp_t s = 0;
                        for (size_t i = 0; i < n; ++i) s += ds[i] * U[k * n + i] * V[l * n + i];
                        out[k * n + l] += s;
                    }
            }
            if (pdW) {
                // dOut/dy = U^T (dOut/dW) V, then through dy/dM
                const fp_t* dw = pdW + (b * out_dim + o) * n2;
                Mat t(n2);
                mm(t.data(), U, dw, n, true, false);
                mm(g.data(), t.data(), V.data(), n, false, false);
                for (size_t m = 0; m < n2; ++m) {
                    fp_t s = 0;
                    for (size_t y = 0; y < n2; ++y) s += g[y] * dy[y * n2 + m];
                    out[m] += s;
                }
            }
            if (pdU) {
                // dOut/dx = U^T (dOut/dU), then through dx/dM
                const fp_t* du = pdU + (b * out_dim + o) * n2;
                mm(g.data(), U, du, n, true, false);
                for (size_t m = 0; m < n2; ++m) {
                    fp_t s = 0;
                    for (size_t x = 0; x < n2; ++x) s += g[x] * dx[x * n2 + m];
                    out[m] += s;
                }
            }
        }
    }
    grad.reset(StSparseLinearTrans::FULL, true, mdM);
}

/* Order-k forward solve, all three outputs (libsanm/tensor_svd.cpp:275-387);
 * the formulas of taylor_dev.h svdw3_fwd_full for a runtime n.  mBu == nullptr:
 * only W_k. */
void svd_w_taylor_fwd(TensorND& mUk, TensorND& mSk, TensorND& mWk,
                      const TensorND& mMk, const TensorND& mMbiask,
                      const TensorND& mU0, const TensorND& mS0, const TensorND& mW0,
                      const TensorND* mBu, const TensorND& mBw) {
    const size_t batch = mMk.shape(0), n = mMk.shape(1), n2 = n * n;
    TensorND oU, oS, oW{TensorShape{batch, n, n}};
    if (mBu) {
        oU.set_shape({batch, n, n});
        oS.set_shape({batch, n});
    }
    Mat V0(n2), D(n2), t0(n2), t1(n2), E(n2), X(n2), Q(n2);
    for (size_t b = 0; b < batch; ++b) {
        const fp_t *Mk = mMk.ptr() + b * n2, *Mb = mMbiask.ptr() + b * n2,
                   *U0 = mU0.ptr() + b * n2, *S0 = mS0.ptr() + b * n,
                   *W0 = mW0.ptr() + b * n2, *Bw = mBw.ptr() + b * n2;
        const fp_t* Bu = mBu ? mBu->ptr() + b * n2 : nullptr;
        mm(V0.data(), W0, U0, n, true, false);
        for (size_t q = 0; q < n2; ++q) D[q] = Mk[q] - Mb[q];
        // E = V0^T D^T U0
        mm(t1.data(), D.data(), V0.data(), n, false, false);
        mm(E.data(), t1.data(), U0, n, true, false);
        mm(t0.data(), V0.data(), Bw, n, true, false);
        mm(t1.data(), t0.data(), V0.data(), n, false, false);  // V0^T Bw V0
        for (size_t r = 0; r < n; ++r)
            for (size_t c = 0; c < n; ++c)
                X[r * n + c] = clip_div(E[c * n + r] - E[r * n + c] - t1[r * n + c] * S0[c],
                                        S0[r] + S0[c]);
        mm(t0.data(), U0, X.data(), n, false, false);
        mm(oW.woptr() + b * n2, t0.data(), V0.data(), n, false, true);  // U0 X V0^T
        if (!Bu) continue;
        for (size_t r = 0; r < n; ++r)
            for (size_t c = 0; c < n; ++c)
                E[r * n + c] += (Bu[c * n + r] - X[c * n + r]) * S0[c];
        fp_t* Sk = oS.woptr() + b * n;
        for (size_t j = 0; j < n; ++j) {
            Sk[j] = E[j * n + j];
            for (size_t i = 0; i < j; ++i) {
                const fp_t v = clip_div(E[i * n + j], S0[i] - S0[j]);
                Q[i * n + j] = v;
                Q[j * n + i] = -Bu[i * n + j] - v;
            }
            Q[j * n + j] = -Bu[j * n + j] / 2;
        }
        mm(oU.woptr() + b * n2, U0, Q.data(), n, false, true);  // U0 Q^T
    }
    mWk = oW;
    if (mBu) {
        mUk = oU;
        mSk = oS;
    }
}

/* Order-k forward solve of the polar mode, M = P W with P = U S U^T
 * (libsanm/tensor_svd.cpp:389-475); the formulas of taylor_dev.h svdwn_fwd_p. */
void svd_w_taylor_fwd_p(TensorND& mPk, TensorND& mWk, const TensorND& mMk,
                        const TensorND& mU, const TensorND& mS, const TensorND& mW,
                        const TensorND& mBm, const TensorND& mBp,
                        const TensorND& mBpw) {
    const size_t batch = mMk.shape(0), n = mMk.shape(1), n2 = n * n;
    TensorND oP{TensorShape{batch, n, n}}, oW{TensorShape{batch, n, n}};
    Mat D(n2), t0(n2), t1(n2), G(n2), x(n2);
    for (size_t b = 0; b < batch; ++b) {
        const fp_t *Mk = mMk.ptr() + b * n2, *U = mU.ptr() + b * n2,
                   *S = mS.ptr() + b * n, *W = mW.ptr() + b * n2,
                   *Bm = mBm.ptr() + b * n2, *Bp = mBp.ptr() + b * n2,
                   *Bpw = mBpw.ptr() + b * n2;
        fp_t *Pk = oP.woptr() + b * n2, *Wk = oW.woptr() + b * n2;
        for (size_t q = 0; q < n2; ++q) D[q] = Bm[q] - Bp[q];
        mm(t0.data(), U, D.data(), n, true, false);
        mm(t1.data(), t0.data(), U, n, false, false);  // U^T D U; E = its transpose
        mm(t0.data(), U, W, n, true, false);           // U^T W
        mm(x.data(), t0.data(), Mk, n, false, true);   // (U^T W) Mk^T
        mm(G.data(), x.data(), U, n, false, false);
        for (size_t r = 0; r < n; ++r)
            for (size_t c = 0; c < n; ++c) G[r * n + c] *= S[r];
        for (size_t r = 0; r < n; ++r)
            for (size_t c = 0; c < n; ++c)
                x[r * n + c] = clip_div(t1[c * n + r] + G[r * n + c] + G[c * n + r],
                                        S[r] + S[c]);
        mm(t0.data(), U, x.data(), n, false, false);
        mm(t1.data(), t0.data(), U, n, false, true);  // PkT = U x U^T
        for (size_t r = 0; r < n; ++r)
            for (size_t c = 0; c < n; ++c) Pk[r * n + c] = t1[c * n + r];
        mm(t0.data(), W, t1.data(), n, true, false);  // W^T PkT
        for (size_t r = 0; r < n; ++r)
            for (size_t c = 0; c < n; ++c)
                G[r * n + c] = Mk[c * n + r] - Bpw[c * n + r] - t0[r * n + c];
        mm(t0.data(), G.data(), U, n, false, false);
        for (size_t c = 0; c < n; ++c) {
            const fp_t si = clip_div(1.0, S[c]);
            for (size_t r = 0; r < n; ++r) t0[r * n + c] *= si;
        }
        mm(t1.data(), t0.data(), U, n, false, true);  // WkT
        for (size_t r = 0; r < n; ++r)
            for (size_t c = 0; c < n; ++c) Wk[r * n + c] = t1[c * n + r];
    }
    mPk = oP;
    mWk = oW;
}
}  // namespace sanm
