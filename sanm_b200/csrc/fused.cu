/* sanm_b200/csrc/fused.cu -- see fused.h / fused_dev.h. */
#include "fused.h"
#include "prof.h"
#include "taylor.h"

namespace sanm_b200 {

using namespace fused;

namespace {

struct KArgs {
    double* arena;
    double* aux;
    const double* ds;
    double* out;
    const double* dminv;
    const double* xbias;
    int* err;
    int64_t npad;
    int K1;
    Params pr;
};

template <class M>
SANM_HD inline double* slot_ptr(const KArgs& a, size_t t, int k) {
    const size_t tile = t / TILE, lane = t % TILE;
    return a.arena + ((tile * a.K1 + k) * M::S) * TILE + lane;
}

template <class M>
SANM_HD inline double* aux_ptr(const KArgs& a, size_t t) {
    const size_t tile = t / TILE, lane = t % TILE;
    return a.aux + tile * (M::AUX * TILE) + lane;
}

template <int N>
SANM_HD inline void ld_soa(double* dst, const double* base, int64_t np,
                           size_t t) {
#pragma unroll
    for (int c = 0; c < N; ++c) dst[c] = base[(int64_t)c * np + t];
}
template <int N>
SANM_HD inline void st_soa(double* base, const double* src, int64_t np,
                           size_t t) {
#pragma unroll
    for (int c = 0; c < N; ++c) base[(int64_t)c * np + t] = src[c];
}

template <class M>
struct Eval0Kernel {
    KArgs a;
    SANM_D void operator()(size_t t) const {
        double Ds[9], dm[9], F0[9], s0[M::S], aux[M::AUX ? M::AUX : 1], P0[9];
        ld_soa<9>(Ds, a.ds, a.npad, t);
        if (a.xbias) {
            double xb[9];
            ld_soa<9>(xb, a.xbias, a.npad, t);
#pragma unroll
            for (int q = 0; q < 9; ++q) Ds[q] += xb[q];
        }
        ld_soa<9>(dm, a.dminv, a.npad, t);
        taylor::mm3(F0, Ds, dm);
        M::eval0(F0, s0, aux, P0, a.pr, a.err);
        st_strided<M::S>(slot_ptr<M>(a, t, 0), s0);
        if (M::AUX) st_strided<M::AUX>(aux_ptr<M>(a, t), aux);
        st_soa<9>(a.out, P0, a.npad, t);
    }
};

/*!
 * Coefficient phase of order k_push (x_k known: finish slot k) followed by the
 * bias phase of order kb (usually k_push + 1); either may be skipped (<= 0).
 */
template <class M>
struct StepKernel {
    KArgs a;
    int k_push, kb, want_out;
    SANM_D void operator()(size_t t) const {
        constexpr int S = M::S;
        double s0[S], aux[M::AUX ? M::AUX : 1];
        ld_strided<S>(s0, slot_ptr<M>(a, t, 0));
        if (M::AUX) ld_strided<M::AUX>(aux, aux_ptr<M>(a, t));
        if (k_push > 0) {
            double Ds[9], dm[9], sk[S], Pt[9];
            ld_soa<9>(Ds, a.ds, a.npad, t);
            ld_soa<9>(dm, a.dminv, a.npad, t);
            taylor::mm3(sk, Ds, dm);  // F_k
            M::tangent(sk, s0, aux, a.pr, sk, Pt);
            double* pk = slot_ptr<M>(a, t, k_push);
            if (k_push >= 2) {
                // bias parts were stored by the bias phase of this order
#pragma unroll
                for (int c = 9; c < S; ++c) sk[c] += pk[c * TILE];
            }
            st_strided<S>(pk, sk);
            if (want_out) {
                if (k_push >= 2) {
                    double Pb[9];
                    ld_soa<9>(Pb, a.out, a.npad, t);
#pragma unroll
                    for (int q = 0; q < 9; ++q) Pt[q] += Pb[q];
                }
                st_soa<9>(a.out, Pt, a.npad, t);
            }
        }
        if (kb > 0) {
            typename M::Acc acc;
            M::zero(acc);
            double Si[S], Sj[S];
            for (int i = 1; 2 * i < kb; ++i) {
                const int j = kb - i;
                ld_strided<S>(Si, slot_ptr<M>(a, t, i));
                ld_strided<S>(Sj, slot_ptr<M>(a, t, j));
                M::accum(acc, Si, Sj, i, j);
                M::accum(acc, Sj, Si, j, i);
            }
            if (kb % 2 == 0) {
                const int i = kb / 2;
                ld_strided<S>(Si, slot_ptr<M>(a, t, i));
                M::accum(acc, Si, Si, i, i);
            }
            double bs[S], Pb[9];
            M::bias(acc, s0, aux, kb, a.pr, bs, Pb);
            double* pb = slot_ptr<M>(a, t, kb);
#pragma unroll
            for (int c = 9; c < S; ++c) pb[c * TILE] = bs[c];
            st_soa<9>(a.out, Pb, a.npad, t);
        }
    }
};

//! forward-mode Jacobian: nine unit tangents per tet, [N][9][9] row-major
template <class M>
struct JacKernel {
    KArgs a;
    double* jac;
    SANM_D void operator()(size_t t) const {
        constexpr int S = M::S;
        double s0[S], aux[M::AUX ? M::AUX : 1], dm[9];
        ld_strided<S>(s0, slot_ptr<M>(a, t, 0));
        if (M::AUX) ld_strided<M::AUX>(aux, aux_ptr<M>(a, t));
        ld_soa<9>(dm, a.dminv, a.npad, t);
        double* jt = jac + t * 81;
        for (int j = 0; j < 9; ++j) {
            const int r0 = j / 3, m0 = j % 3;
            double ts[S], Pt[9];
#pragma unroll
            for (int q = 0; q < 9; ++q) ts[q] = 0;
            // d F[r0][c] / d Ds[r0][m0] = DmInv[m0][c]
            for (int c = 0; c < 3; ++c) ts[r0 * 3 + c] = dm[m0 * 3 + c];
            M::tangent(ts, s0, aux, a.pr, ts, Pt);
#pragma unroll
            for (int o = 0; o < 9; ++o) jt[o * 9 + j] = Pt[o];
        }
    }
};

}  // namespace

FusedProp::FusedProp(const Graph& g, const FusedSpec& spec, uint64_t batch,
                     int max_order)
        : m_mat{spec.material},
          m_params{spec.params},
          m_n{batch},
          m_npad{(int64_t)((batch + TILE - 1) / TILE * TILE)},
          m_K{max_order} {
    switch (m_mat) {
        case MAT_NHC:
            m_S = NHC::S;
            break;
        case MAT_NHI:
            m_S = NHI::S;
            break;
        case MAT_ARAP:
            m_S = ARAP::S;
            break;
        default:
            sanm_throw(SANM_ERR_ASSERT, "bad fused material");
    }
    const size_t ntiles = (size_t)m_npad / TILE;
    m_arena.alloc(ntiles * (size_t)(m_K + 1) * m_S * TILE);
    if (m_mat == MAT_ARAP) m_aux.alloc(ntiles * ARAP::AUX * TILE);
    m_ds.alloc(9 * (size_t)m_npad);
    m_out.alloc(9 * (size_t)m_npad);
    m_dminv.alloc(9 * (size_t)m_npad);
    m_err.alloc(1);
    m_err.zero();
    // padding lanes read these: keep them finite
    m_ds.zero();
    m_dminv.zero();
    rt::DBuf<double> stage(batch * 9);
    auto upload = [&](int gop, rt::DBuf<double>& dst) {
        const GraphOp& c = g.ops[gop];
        sanm_check(c.cdata.size() == batch * 9);
        rt::h2d(stage.ptr(), c.cdata.data(), batch * 9 * sizeof(double));
        aos_to_soa(dst.ptr(), stage.ptr(), batch, 9, m_npad);
    };
    upload(spec.dminv_gop, m_dminv);
    if (spec.xbias_gop >= 0) {
        m_xbias.alloc(9 * (size_t)m_npad);
        m_xbias.zero();
        upload(spec.xbias_gop, m_xbias);
    }
    rt::sync();  // `stage` is released on return
}

void FusedProp::reset() {
    m_pending = -1;
    m_out_is_bias = false;
    m_err.zero();
}

template <class M>
void FusedProp::launch_step(int k_push, int kb, bool want_out) {
    KArgs a{m_arena.ptr(), m_aux.ptr(), m_ds.ptr(),  m_out.ptr(),
            m_dminv.ptr(), m_xbias.ptr(), m_err.ptr(), m_npad,
            m_K + 1,       m_params};
    if (k_push == 0) {
        SANM_PROF("k_fused_eval");
        rt::for_each(m_n, Eval0Kernel<M>{a});
    } else {
        SANM_PROF("k_fused_step");
        rt::for_each(m_n, StepKernel<M>{a, k_push, kb, (int)want_out});
    }
}

void FusedProp::step(int k_push, int kb, bool want_out) {
    switch (m_mat) {
        case MAT_NHC:
            launch_step<NHC>(k_push, kb, want_out);
            break;
        case MAT_NHI:
            launch_step<NHI>(k_push, kb, want_out);
            break;
        default:
            launch_step<ARAP>(k_push, kb, want_out);
            break;
    }
}

void FusedProp::push(int k) {
    sanm_check(m_pending < 0 && k >= 0 && k <= m_K);
    if (k == 0) {
        step(0, -1, true);
        m_out_is_bias = false;
    } else {
        m_pending = k;
    }
}

void FusedProp::flush() {
    if (m_pending < 0) return;
    // the output of a bare push is the full coefficient: bias (still in the
    // output buffer from the bias phase of this order) + tangent
    sanm_check(m_pending == 1 || m_out_is_bias);
    step(m_pending, -1, true);
    m_pending = -1;
    m_out_is_bias = false;
}

const double* FusedProp::out_slot() {
    flush();
    return m_out.ptr();
}

void FusedProp::next_bias(int kb) {
    sanm_check(kb >= 1 && kb <= m_K);
    if (kb == 1) {
        // the first-order bias is identically zero (libsanm/symbolic.cpp:278-285)
        sanm_check(m_pending < 0);
        rt::dzero(m_out.ptr(), m_out.size() * sizeof(double));
    } else {
        sanm_check(m_pending == kb - 1 || m_pending < 0);
        step(m_pending, kb, false);
        m_pending = -1;
    }
    m_out_is_bias = true;
}

void FusedProp::jacobian(double* jac) {
    SANM_PROF("k_fused_jacobian");
    KArgs a{m_arena.ptr(), m_aux.ptr(), m_ds.ptr(),  m_out.ptr(),
            m_dminv.ptr(), m_xbias.ptr(), m_err.ptr(), m_npad,
            m_K + 1,       m_params};
    switch (m_mat) {
        case MAT_NHC:
            rt::for_each(m_n, JacKernel<NHC>{a, jac});
            break;
        case MAT_NHI:
            rt::for_each(m_n, JacKernel<NHI>{a, jac});
            break;
        default:
            rt::for_each(m_n, JacKernel<ARAP>{a, jac});
            break;
    }
}

}  // namespace sanm_b200
