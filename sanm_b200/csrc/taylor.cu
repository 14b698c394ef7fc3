/* sanm_b200/csrc/taylor.cu -- see taylor.h.  Kernels: one thread per batch item
 * (tetrahedron), walking the lowered operator list (taylor_dev.h). */
#include "taylor.h"
#include "cache.h"
#include "prof.h"
#include "remap.h"

#include <cstdlib>

namespace sanm_b200 {

using namespace taylor;

namespace {

struct OrderKernel {
    DProgram prog;
    int k, phase, sb_zero;
    SANM_D void operator()(size_t t) const {
        run_item(prog, t, k, phase, sb_zero);
    }
};

struct Aos2Soa {
    double* soa;
    const double* aos;
    uint64_t n;
    int e;
    int64_t npad;
    SANM_D void operator()(size_t i) const {
        // i enumerates the SoA side so that writes are coalesced
        size_t c = i / n, t = i % n;
        soa[c * npad + t] = aos[t * e + c];
    }
};

struct Soa2Aos {
    double* aos;
    const double* soa;
    uint64_t n;
    int e;
    int64_t npad;
    SANM_D void operator()(size_t i) const {
        size_t c = i / n, t = i % n;
        aos[t * e + c] = soa[c * npad + t];
    }
};

struct FillUnit {
    double* slot;  // [e][npad]
    uint64_t n;
    int64_t npad;
    int j;  // -1: all ones
    SANM_D void operator()(size_t i) const {
        size_t c = i / n, t = i % n;
        slot[c * npad + t] = (j < 0 || (int)c == j) ? 1.0 : 0.0;
    }
};

struct ExtractJacCol {
    double* jac;        // [n][odim][idim]
    const double* out;  // [odim][npad]
    uint64_t n;
    int odim, idim, j;
    int64_t npad;
    SANM_D void operator()(size_t i) const {
        size_t o = i / n, t = i % n;
        jac[(t * odim + o) * idim + j] = out[o * npad + t];
    }
};

}  // namespace

void aos_to_soa(double* soa, const double* aos, uint64_t n, int e,
                int64_t npad) {
    rt::for_each(n * e, Aos2Soa{soa, aos, n, e, npad});
}
void soa_to_aos(double* aos, const double* soa, uint64_t n, int e,
                int64_t npad) {
    rt::for_each(n * e, Soa2Aos{aos, soa, n, e, npad});
}

TaylorProp::TaylorProp(std::shared_ptr<const Graph> g, int out_var,
                       int max_order)
        : m_graph{std::move(g)}, m_out_var{out_var}, m_max_order{max_order} {
    sanm_check(max_order >= 1 && max_order <= MAX_ORDER, "max_order=%d",
               max_order);
    m_graph->check_var(out_var);
}

TaylorProp::~TaylorProp() = default;

void TaylorProp::reset() {
    m_order = -1;
    m_xi_known = false;
    m_jac_done = false;
    if (m_flags.size()) m_flags.zero();
    if (m_fused) m_fused->reset();
}

int TaylorProp::state_doubles() const {
    if (m_fused) return m_fused->state_doubles();
    int s = 0;
    for (auto& v : m_low.vars) {
        if (v.needs_stack && !v.is_const) s += v.e;
    }
    for (auto& o : m_low.ops) s += o.eauxk;
    return s;
}

void TaylorProp::prepare(const Shape& x_shape) {
    rt::check_device();
    if (m_prepared && in_shape() == x_shape) {
        reset();
        return;
    }
    m_low = lower_graph(*m_graph, m_out_var, x_shape);
    const uint64_t n = m_low.batch;
    m_fused.reset();
    {
        FusedSpec spec = match_fea_graph(*m_graph, m_out_var, x_shape);
        if (spec.material != fused::MAT_NONE) {
            m_fused.reset(new FusedProp{*m_graph, spec, n, m_max_order});
            m_fused->set_local_rows(m_local_rows);
            m_npad = m_fused->npad();
            m_arena.release();
            m_prepared = true;
            reset();
            return;
        }
    }
    m_npad = (int64_t)((n + 31) / 32 * 32);
    const int K = m_max_order;

    // arena layout (offsets in doubles)
    size_t off = 0;
    auto take = [&](size_t doubles) {
        size_t r = off;
        off += doubles;
        return r;
    };
    struct VOff {
        size_t p0, pk;
    };
    std::vector<VOff> voff(m_low.vars.size());
    for (size_t i = 0; i < m_low.vars.size(); ++i) {
        auto& v = m_low.vars[i];
        voff[i].p0 = take((size_t)v.e * m_npad);
        voff[i].pk = 0;
        if (!v.is_const) {
            voff[i].pk = take((size_t)v.e * m_npad * (v.needs_stack ? K : 1));
        }
    }
    struct OOff {
        size_t sb, aux0, auxk;
    };
    std::vector<OOff> ooff(m_low.ops.size());
    int nflags = 1;  // flag 0 = program error word
    for (size_t i = 0; i < m_low.ops.size(); ++i) {
        auto& o = m_low.ops[i];
        ooff[i].sb = take((size_t)o.esb * m_npad);
        ooff[i].aux0 = take((size_t)o.eaux0 * m_npad);
        ooff[i].auxk = take((size_t)o.eauxk * m_npad * (K + 1));
        if (o.need_flag) ++nflags;
    }
    m_arena.alloc(off);
    m_arena.zero();
    m_flags.alloc(nflags);
    m_flags.zero();

    double* base = m_arena.ptr();
    m_hvars.resize(m_low.vars.size());
    for (size_t i = 0; i < m_low.vars.size(); ++i) {
        auto& v = m_low.vars[i];
        DVar d{};
        d.p0 = base + voff[i].p0;
        d.pk = v.is_const ? nullptr : base + voff[i].pk;
        d.kstride = v.needs_stack ? (int64_t)v.e * m_npad : 0;
        d.e = v.e;
        d.d1 = v.d1;
        d.d2 = v.d2;
        d.is_const = v.is_const;
        m_hvars[i] = d;
    }
    m_hops.resize(m_low.ops.size());
    int flag_i = 1;
    for (size_t i = 0; i < m_low.ops.size(); ++i) {
        auto& o = m_low.ops[i];
        const GraphOp& go = m_graph->ops[o.gop];
        DOp d{};
        d.type = o.type;
        d.nin = (int)o.in.size();
        d.nout = (int)o.out.size();
        for (int j = 0; j < d.nin; ++j) d.in[j] = o.in[j];
        for (int j = 0; j < d.nout; ++j) d.out[j] = o.out[j];
        for (size_t j = 0; j < go.f.size(); ++j) d.f[j] = go.f[j];
        d.const_op = o.const_op;
        d.sb = o.esb ? base + ooff[i].sb : nullptr;
        d.aux0 = o.eaux0 ? base + ooff[i].aux0 : nullptr;
        d.auxk = o.eauxk ? base + ooff[i].auxk : nullptr;
        d.auxk_stride = (int64_t)o.eauxk * m_npad;
        d.flag = o.need_flag ? m_flags.ptr() + flag_i++ : nullptr;
        switch (o.type) {
            case OP_REDUCE_SUM:
                d.i[0] = go.ip[0];
                break;
            case OP_MATMUL: {
                auto &a = m_low.vars[o.in[0]], &b = m_low.vars[o.in[1]];
                d.i[0] = a.d1 == 3 && a.d2 == 3 && b.d2 == 3;
                break;
            }
            case OP_MULEYE:
                d.i[0] = go.ip[0];
                break;
            case OP_MATINVMUL:
                d.i[0] = go.ip[0];
                d.i[1] = go.ip[1];
                d.i[2] = m_low.vars[o.in[0]].d1 == 3;
                break;
            case OP_SVDW:
                d.i[0] = go.ip[0];
                d.i[1] = o.svdw_full;
                break;
            case OP_SLICE: {
                // resolve begin against the per-item axis size
                auto& xv = m_low.vars[o.in[0]];
                int axis = go.ip[0], stride = go.ip[5];
                int size = (int)xv.shape.dim[axis], begin;
                if (go.ip[1]) {
                    begin = go.ip[2];
                    if (begin < 0) begin += size;
                } else {
                    begin = stride > 0 ? 0 : size - 1;
                }
                // map tensor axis to the (d1, d2) view of the item
                d.i[0] = axis;
                if (xv.shape.rank == 2) d.i[0] = 1;
                d.i[1] = begin;
                d.i[2] = stride;
                break;
            }
            case OP_CONCAT:
                d.i[0] = go.ip[0];
                if (m_low.vars[o.in[0]].shape.rank == 2) d.i[0] = 1;
                break;
            default:
                break;
        }
        m_hops[i] = d;

        if (o.type == OP_CONSTANT) {
            // upload the constant, transposed to SoA
            auto& v = m_low.vars[o.out[0]];
            m_stage.ensure(go.cdata.size());
            rt::h2d(m_stage.ptr(), go.cdata.data(),
                    go.cdata.size() * sizeof(double));
            aos_to_soa(m_hvars[o.out[0]].p0, m_stage.ptr(), n, v.e, m_npad);
        }
    }
    m_dvars.upload(m_hvars);
    m_dops.upload(m_hops);
    m_prog.vars = m_dvars.ptr();
    m_prog.ops = m_dops.ptr();
    m_prog.nops = (int)m_hops.size();
    m_prog.npad = m_npad;
    m_prog.err = m_flags.ptr();
    m_prepared = true;
    reset();
}

double* TaylorProp::input_slot(int k) const {
    sanm_check(m_prepared && k >= 0 && k <= m_max_order);
    if (m_fused) return m_fused->ds_slot();
    return vbase(m_hvars[m_low.input], k);
}

const double* TaylorProp::output_slot(int k) const {
    sanm_check(m_prepared && k >= 0 && k <= m_max_order);
    if (m_fused) return m_fused->out_slot();
    return vbase(m_hvars[m_low.output], k);
}

void TaylorProp::launch(int k, int phase, int sb_zero) {
    static const char* const TAGS[] = {"k_order_eval", "k_order_bias",
                                       "k_order_coef"};
    SANM_PROF(m_in_jacobian ? "k_jacobian" : TAGS[phase]);
    rt::for_each(m_low.batch, OrderKernel{m_prog, k, phase, sb_zero});
}

void TaylorProp::push_xi_filled() {
    // libsanm/symbolic.cpp:162-204
    sanm_check(m_prepared && !m_xi_known, "push_xi called twice in a row");
    if (m_fused) {
        if (m_order < 0) m_order = 0;
        m_fused->push(m_order);
        m_xi_known = true;
        return;
    }
    if (m_order < 0) {
        m_order = 0;
        launch(0, PH_EVAL, 0);
    } else {
        sanm_check(m_order >= 1);
        launch(m_order, PH_COEF, m_order == 1);
    }
    m_xi_known = true;
}

void TaylorProp::compute_next_order_bias() {
    // libsanm/symbolic.cpp:249-289
    sanm_check(m_prepared && m_xi_known,
               "compute_next_order_bias needs a preceding push_xi");
    sanm_check(m_order + 1 <= m_max_order, "order %d exceeds max_order %d",
               m_order + 1, m_max_order);
    if (m_fused) {
        // the Jacobian depends on the order-0 slot only: computed on demand
        ++m_order;
        m_xi_known = false;
        m_fused->next_bias(m_order);
        return;
    }
    ensure_jacobian();
    ++m_order;
    m_xi_known = false;
    if (m_order == 1) {
        // bias of the first-order term is identically zero
        // (libsanm/symbolic.cpp:278-285)
        auto& ov = m_low.vars[m_low.output];
        rt::dzero(vbase(m_hvars[m_low.output], 1),
                  (size_t)ov.e * m_npad * sizeof(double));
    } else {
        launch(m_order, PH_BIAS, 0);
    }
}

bool TaylorProp::attach_io(const Remap& rin, const Remap& rout) {
    if (!m_fused) return false;
    if (m_fused->has_io()) return true;
    const char* off = getenv("SANM_NO_FUSED_IO");
    if (off && off[0] == '1') return false;
    // derived tables depend on the two maps only: shared between solver
    // instances over the same mesh (cache.h)
    static LruCache<const FusedIO> cache{6};
    uint64_t key = rin.hash_full() * 0x9E3779B97F4A7C15ull ^ rout.hash_full();
    key = hash_bytes(&key, 8, m_low.batch);
    std::shared_ptr<const FusedIO> io = cache.get(key);
    if (!io || io->n_tets != m_low.batch) {
        std::unique_ptr<FusedIO> fresh = FusedIO::analyse(rin, rout, m_low.batch);
        if (!fresh) return false;
        fresh->hash = key;
        io = std::move(fresh);
        cache.put(key, io);
    }
    m_fused->attach_io(std::move(io));
    return true;
}

void TaylorProp::push_xi_from(const double* x, double* fx) {
    sanm_check(m_prepared && m_fused && m_fused->has_io() && !m_xi_known);
    if (m_order < 0) m_order = 0;
    m_fused->push_from(m_order, x, fx);
    m_xi_known = true;
}

void TaylorProp::compute_next_order_bias_to(double* b) {
    sanm_check(m_prepared && m_fused && m_fused->has_io() && m_xi_known);
    sanm_check(m_order + 1 <= m_max_order, "order %d exceeds max_order %d",
               m_order + 1, m_max_order);
    ++m_order;
    m_xi_known = false;
    m_fused->next_bias_to(m_order, b);
}

const double* TaylorProp::ensure_jacobian() {
    if (m_jac_done) return m_jac.ptr();
    if (m_fused) {
        sanm_check(m_order >= 0, "Jacobian needs the order-0 point");
        m_jac.ensure((size_t)81 * m_npad);
        m_fused->jacobian(m_jac.ptr());
        m_jac_done = true;
        return m_jac.ptr();
    }
    sanm_check(m_order == 0 && m_xi_known,
               "Jacobian must be requested right after the first push_xi");
    const uint64_t n = m_low.batch;
    const int idim = in_elems(), odim = out_elems();
    double* in1 = vbase(m_hvars[m_low.input], 1);
    const double* out1 = vbase(m_hvars[m_low.output], 1);
    m_in_jacobian = true;
    if (m_low.elemwise) {
        m_jac.ensure(n * odim);
        rt::for_each(n * idim, FillUnit{in1, n, m_npad, -1});
        launch(1, PH_COEF, 1);
        soa_to_aos(m_jac.ptr(), out1, n, odim, m_npad);
    } else {
        m_jac.ensure(n * odim * idim);
        for (int j = 0; j < idim; ++j) {
            rt::for_each(n * idim, FillUnit{in1, n, m_npad, j});
            launch(1, PH_COEF, 1);
            rt::for_each(n * odim,
                         ExtractJacCol{m_jac.ptr(), out1, n, odim, idim, j, m_npad});
        }
    }
    m_in_jacobian = false;
    m_jac_done = true;
    return m_jac.ptr();
}

void TaylorProp::jacobian_to_host(double* out) {
    const double* j = ensure_jacobian();
    const uint64_t n = m_low.batch;
    const size_t per = jacobian_type() == JAC_FULL
                               ? (size_t)out_elems() * in_elems()
                               : (size_t)out_elems();
    if (jacobian_pitch()) {
        m_stage.ensure(n * per);
        soa_to_aos(m_stage.ptr(), j, n, (int)per, jacobian_pitch());
        j = m_stage.ptr();
    }
    rt::d2h(out, j, n * per * sizeof(double));
}

void TaylorProp::push_xi_host(const double* x, const Shape& shape) {
    if (m_order < 0) {
        prepare(shape);
    } else {
        sanm_check(shape == in_shape(), "push_xi shape %s != %s",
                   shape.str().c_str(), in_shape().str().c_str());
    }
    const uint64_t n = m_low.batch;
    const int e = in_elems();
    m_stage.ensure(n * e);
    rt::h2d(m_stage.ptr(), x, n * e * sizeof(double));
    aos_to_soa(input_slot(m_order < 0 ? 0 : m_order), m_stage.ptr(), n, e, m_npad);
    push_xi_filled();
}

void TaylorProp::output_to_host(int k, double* out) const {
    const uint64_t n = m_low.batch;
    const int e = out_elems();
    auto& st = const_cast<rt::DBuf<double>&>(m_stage);
    st.ensure(n * e);
    soa_to_aos(st.ptr(), output_slot(k), n, e, m_npad);
    rt::d2h(out, st.ptr(), n * e * sizeof(double));
}

void TaylorProp::check_device_error() {
    if (!m_prepared) return;
    int err = 0;
    rt::d2h(&err, m_fused ? m_fused->err_flag() : m_flags.ptr(), sizeof(int));
    if (err & 1) {
        sanm_throw(SANM_ERR_NUMERICAL, "0^p when p is not integer");
    }
}

}  // namespace sanm_b200
