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
    uint64_t n;
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

//! device view of FusedIO (fused_io.h) plus the nodal vectors of one launch
struct IOArgs {
    const uint32_t* uidx;
    const uint32_t* gidx;
    const double* gcoef;
    int wu;
    const double* nrm;
    const uint16_t* lrow;
    const uint32_t* blk_ptr;
    const uint32_t* con_ptr;
    const uint16_t* con_idx;
    double* partial;
    const double* x;  //!< nodal coefficient being pushed
};

//! Ds = remap_in(x) for one tet (fea/mesh_template.h:73-110)
template <bool IO>
SANM_HD inline void load_ds(double* Ds, const KArgs& a, const IOArgs& io,
                            size_t t) {
    if (!IO) {
        ld_soa<9>(Ds, a.ds, a.npad, t);
        return;
    }
#pragma unroll
    for (int q = 0; q < 9; ++q) {
        double s = 0;
        for (int w = 0; w < io.wu; ++w) {
            const uint32_t u = io.uidx[((int64_t)w * 9 + q) * a.npad + t];
            if (u != 0xFFFFFFFFu) {
                const double v = io.x[u >> 1];
                s += (u & 1) ? -v : v;
            }
        }
        if (io.gidx) {
            const uint32_t g = io.gidx[(int64_t)q * a.npad + t];
            if (g != 0xFFFFFFFFu) s += io.x[g] * io.gcoef[(int64_t)q * a.npad + t];
        }
        Ds[q] = s;
    }
}

//! the 12 nodal contributions P n_d of one tet into shared memory
//! (fea/mesh_template.h:145-160)
SANM_HD inline void emit_forces(double* vals, int lane, const double* P,
                                const KArgs& a, const IOArgs& io, size_t t) {
#pragma unroll
    for (int d = 0; d < 4; ++d) {
        const double n0 = io.nrm[((int64_t)d * 3 + 0) * a.npad + t];
        const double n1 = io.nrm[((int64_t)d * 3 + 1) * a.npad + t];
        const double n2 = io.nrm[((int64_t)d * 3 + 2) * a.npad + t];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            vals[(d * 3 + c) * TILE + lane] =
                    n0 * P[c * 3] + n1 * P[c * 3 + 1] + n2 * P[c * 3 + 2];
        }
    }
}

//! first level of the assembly: one partial sum per distinct row of the block
SANM_HD inline void block_reduce(const rt::BlockCtx& c, const double* vals,
                                 const IOArgs& io) {
    const uint32_t base = io.blk_ptr[c.bid], end = io.blk_ptr[c.bid + 1];
    for (uint32_t g = base + c.tid; g < end; g += c.nthr) {
        double s = 0;
        for (uint32_t p = io.con_ptr[g], pe = io.con_ptr[g + 1]; p < pe; ++p) {
            s += vals[io.con_idx[p]];
        }
        io.partial[g] = s;
    }
}

template <class M, bool IO>
struct Eval0Kernel {
    KArgs a;
    IOArgs io;
    SANM_D void operator()(const rt::BlockCtx& c) const {
        double* vals = reinterpret_cast<double*>(c.smem);
        for (int lane = c.tid; lane < TILE; lane += c.nthr) {
            const size_t t = c.bid * TILE + lane;
            if (t >= a.n) continue;
            double Ds[9], dm[9], F0[9], s0[M::S], aux[M::AUX ? M::AUX : 1],
                    P0[9];
            load_ds<IO>(Ds, a, io, t);
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
            if (IO) emit_forces(vals, lane, P0, a, io, t);
        }
        if (IO) {
            c.sync();
            block_reduce(c, vals, io);
        }
    }
};

/*!
 * Coefficient phase of order k_push (x_k known: finish slot k) followed by the
 * bias phase of order kb (usually k_push + 1); either may be skipped (<= 0).
 * One block owns one tile of TILE tets.
 */
template <class M, bool IO>
struct StepKernel {
    KArgs a;
    IOArgs io;
    int k_push, kb, want_out;
    SANM_D void operator()(const rt::BlockCtx& c) const {
        double* vals = reinterpret_cast<double*>(c.smem);
        for (int lane = c.tid; lane < TILE; lane += c.nthr) {
            const size_t t = c.bid * TILE + lane;
            if (t < a.n) item(t, lane, vals);
        }
        if (IO && kb > 0) {
            c.sync();
            block_reduce(c, vals, io);
        }
    }

    //! slot 0 (and the ARAP factors).  Loaded once per phase with a
    //! default-cached load instead of being kept live across the convolution
    //! loop: its 2 S registers then hold loads in flight, and ptxas no longer
    //! spills (NH-C 88, NH-I 208, ARAP 88 bytes of stack before).  The second
    //! read of a launch hits L2 -- the streamed slots are evict-first.
    //! Measured (B200, K = 20): NH-C 2.51 -> 2.56 G tet*order/s at 2 M tets
    //! and 2.65 -> 2.69 G at 10 M (step kernel 0.887 -> 0.904 of the copy
    //! bandwidth), NH-I 1.90 -> 1.99 G, ARAP 1.59 -> 1.68 G.
    SANM_D void load_base(double* s0, double* aux, size_t t) const {
        const double* p = slot_ptr<M>(a, t, 0);
#pragma unroll
        for (int q = 0; q < M::S; ++q) s0[q] = p[q * TILE];
        if (M::AUX) ld_strided<M::AUX>(aux, aux_ptr<M>(a, t));
    }

    SANM_D void push_phase(size_t t, const double* s0, const double* aux) const {
        constexpr int S = M::S;
        double Ds[9], dm[9], sk[S], Pt[9];
        load_ds<IO>(Ds, a, io, t);
        ld_soa<9>(dm, a.dminv, a.npad, t);
        taylor::mm3(sk, Ds, dm);  // F_k
        M::tangent(sk, s0, aux, a.pr, sk, Pt);
        double* pk = slot_ptr<M>(a, t, k_push);
        if (k_push >= 2) {
            // bias parts were stored by the bias phase of this order
#pragma unroll
            for (int q = 9; q < S; ++q) sk[q] += pk[q * TILE];
        }
        st_strided<S>(pk, sk);
        if (!IO && want_out) {
            if (k_push >= 2) {
                double Pb[9];
                ld_soa<9>(Pb, a.out, a.npad, t);
#pragma unroll
                for (int q = 0; q < 9; ++q) Pt[q] += Pb[q];
            }
            st_soa<9>(a.out, Pt, a.npad, t);
        }
    }

    SANM_D void convolve(typename M::Acc& acc, size_t t) const {
        constexpr int S = M::S;
        M::zero(acc);
        double Si[S], Sj[S];
        for (int i = 1; 2 * i < kb; ++i) {
            const int j = kb - i;
            ld_strided<S>(Si, slot_ptr<M>(a, t, i));
            ld_strided<S>(Sj, slot_ptr<M>(a, t, j));
            M::accum(acc, Si, Sj, i, j, a.pr);
            M::accum(acc, Sj, Si, j, i, a.pr);
        }
        if (kb % 2 == 0) {
            const int i = kb / 2;
            ld_strided<S>(Si, slot_ptr<M>(a, t, i));
            M::accum(acc, Si, Si, i, i, a.pr);
        }
    }

    SANM_D void bias_phase(const typename M::Acc& acc, size_t t, int lane,
                           double* vals, const double* s0,
                           const double* aux) const {
        constexpr int S = M::S;
        double bs[S], Pb[9];
        M::bias(acc, s0, aux, kb, a.pr, bs, Pb);
        double* pb = slot_ptr<M>(a, t, kb);
#pragma unroll
        for (int q = 9; q < S; ++q) pb[q * TILE] = bs[q];
        if (IO) {
            emit_forces(vals, lane, Pb, a, io, t);
        } else {
            st_soa<9>(a.out, Pb, a.npad, t);
        }
    }

    SANM_D void item(size_t t, int lane, double* vals) const {
        constexpr int S = M::S;
        if (k_push > 0) {
            double s0[S], aux[M::AUX ? M::AUX : 1];
            load_base(s0, aux, t);
            push_phase(t, s0, aux);
        }
        if (kb > 0) {
            typename M::Acc acc;
            convolve(acc, t);
            double s0[S], aux[M::AUX ? M::AUX : 1];
            load_base(s0, aux, t);
            bias_phase(acc, t, lane, vals, s0, aux);
        }
    }
};

//! forward-mode Jacobian: nine unit tangents per tet, stored
//! structure-of-arrays [81][npad] (entry o * 9 + j) so that stores coalesce
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
        double* jt = jac + t;
        // unrolled: with a runtime j the unit tangent is indexed dynamically and
        // lives in local memory (224-byte stack frame before)
#pragma unroll
        for (int j = 0; j < 9; ++j) {
            const int r0 = j / 3, m0 = j % 3;
            double ts[S], Pt[9];
#pragma unroll
            for (int q = 0; q < 9; ++q) ts[q] = 0;
            // d F[r0][c] / d Ds[r0][m0] = DmInv[m0][c]
#pragma unroll
            for (int c = 0; c < 3; ++c) ts[r0 * 3 + c] = dm[m0 * 3 + c];
            M::tangent(ts, s0, aux, a.pr, ts, Pt);
#pragma unroll
            for (int o = 0; o < 9; ++o) jt[(int64_t)(o * 9 + j) * a.npad] = Pt[o];
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
    m_pending_x = nullptr;
    m_out_is_bias = false;
    m_err.zero();
}

void FusedProp::attach_io(std::shared_ptr<const FusedIO> io) {
    sanm_check(!io || (io->n_tets == m_n && io->npad == m_npad));
    m_io = std::move(io);
}

template <class M>
void FusedProp::launch_step(int k_push, int kb, bool want_out, const double* x) {
    KArgs a{m_arena.ptr(), m_aux.ptr(),   m_ds.ptr(),  m_out.ptr(),
            m_dminv.ptr(), m_xbias.ptr(), m_err.ptr(), m_npad,
            m_n,           m_K + 1,       m_params};
    const size_t ntiles = (size_t)m_npad / TILE;
    const bool use_io = x != nullptr;
    IOArgs io{};
    if (use_io) {
        sanm_check(m_io && !want_out);
        io = IOArgs{m_io->uidx.ptr(),    m_io->gidx.ptr(),    m_io->gcoef.ptr(),
                    m_io->wu,            m_io->nrm.ptr(),     m_io->lrow.ptr(),
                    m_io->blk_ptr.ptr(), m_io->con_ptr.ptr(), m_io->con_idx.ptr(),
                    m_io->partial.ptr(), x};
    }
    const size_t smem = use_io ? 12 * TILE * sizeof(double) : 0;
    if (k_push == 0) {
        SANM_PROF("k_fused_eval");
        if (use_io) {
            rt::for_each_block(ntiles, TILE, smem, Eval0Kernel<M, true>{a, io});
        } else {
            rt::for_each_block(ntiles, TILE, smem, Eval0Kernel<M, false>{a, io});
        }
    } else {
        SANM_PROF("k_fused_step");
        // occupancy target per energy model (fused_dev.h M::MINB, measured):
        // the kernel is HBM-bound and wants as many loads in flight as the
        // register file allows; __launch_bounds__(128, 1) (255 registers) gives
        // 1.76 G tet*order/s where (128, 5) gives 2.49 G
        if (use_io) {
            // SANM_STEP_MINB = 4 | 5 | 6 overrides the occupancy target (tuning
            // knob, tools/step_variants.sh)
            static const int minb = getenv("SANM_STEP_MINB") ? atoi(getenv("SANM_STEP_MINB")) : M::MINB;
            StepKernel<M, true> k{a, io, k_push, kb, (int)want_out};
            if (minb == M::MINB) {
                rt::for_each_block<TILE, M::MINB>(ntiles, TILE, smem, k);
            } else if (minb == 4) {
                rt::for_each_block<TILE, 4>(ntiles, TILE, smem, k);
            } else if (minb == 5) {
                rt::for_each_block<TILE, 5>(ntiles, TILE, smem, k);
            } else {
                rt::for_each_block<TILE, 6>(ntiles, TILE, smem, k);
            }
        } else {
            rt::for_each_block(
                    ntiles, TILE, smem,
                    StepKernel<M, false>{a, io, k_push, kb, (int)want_out});
        }
    }
}

void FusedProp::step(int k_push, int kb, bool want_out, const double* x) {
    switch (m_mat) {
        case MAT_NHC:
            launch_step<NHC>(k_push, kb, want_out, x);
            break;
        case MAT_NHI:
            launch_step<NHI>(k_push, kb, want_out, x);
            break;
        default:
            launch_step<ARAP>(k_push, kb, want_out, x);
            break;
    }
}

void FusedProp::push(int k) {
    sanm_check(m_pending < 0 && k >= 0 && k <= m_K);
    if (k == 0) {
        step(0, -1, true, nullptr);
        m_out_is_bias = false;
    } else {
        m_pending = k;
        m_pending_x = nullptr;
    }
}

void FusedProp::push_from(int k, const double* x, double* fx) {
    sanm_check(m_io && m_pending < 0 && k >= 0 && k <= m_K);
    if (k == 0) {
        step(0, -1, false, x);
        m_out_is_bias = false;
        if (fx) {
            SANM_PROF("k_assemble_final");
            m_io->finalize(fx, m_local_rows);
        }
    } else {
        m_pending = k;
        m_pending_x = x;
    }
}

void FusedProp::flush() {
    if (m_pending < 0) return;
    // the output of a bare push is the full coefficient: bias (still in the
    // output buffer from the bias phase of this order) + tangent
    sanm_check(!m_pending_x, "bare push is not available with fused I/O");
    sanm_check(m_pending == 1 || m_out_is_bias);
    step(m_pending, -1, true, nullptr);
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
        sanm_check((m_pending == kb - 1 || m_pending < 0) && !m_pending_x);
        step(m_pending, kb, false, nullptr);
        m_pending = -1;
    }
    m_out_is_bias = true;
}

void FusedProp::next_bias_to(int kb, double* b) {
    sanm_check(m_io && kb >= 1 && kb <= m_K);
    if (kb == 1) {
        sanm_check(m_pending < 0);
        if (m_local_rows) {
            rt::dzero(b + m_io->row_lo,
                      (m_io->row_hi - m_io->row_lo) * sizeof(double));
        } else {
            rt::dzero(b, m_io->nout * sizeof(double));
        }
        return;
    }
    sanm_check(m_pending == kb - 1 && m_pending_x,
               "fused I/O needs push_from before next_bias_to");
    step(m_pending, kb, false, m_pending_x);
    m_pending = -1;
    m_pending_x = nullptr;
    SANM_PROF("k_assemble_final");
    m_io->finalize(b, m_local_rows);
}

void FusedProp::jacobian(double* jac) {
    SANM_PROF("k_fused_jacobian");
    KArgs a{m_arena.ptr(), m_aux.ptr(),   m_ds.ptr(),  m_out.ptr(),
            m_dminv.ptr(), m_xbias.ptr(), m_err.ptr(), m_npad,
            m_n,           m_K + 1,       m_params};
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
