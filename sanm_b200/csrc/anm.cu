/* sanm_b200/csrc/anm.cu -- see anm.h. */
#include "anm.h"

#include <future>
#include "pade_dev.h"
#include "cache.h"
#include "comm.h"
#include "shard.h"
#include "poly.h"
#include "prof.h"

#include <algorithm>
#include <cmath>
#include <cstring>

namespace sanm_b200 {

namespace {

/* device scalar slots */
enum Slot : int {
    SC_T1 = 0,
    SC_XGT_X1 = 1,
    SC_TI = 2,
    SC_DOT = 3,
    SC_RMS2 = 4,
    SC_CHECK0 = 8,  //!< + 2*i: residual check of order i ; + 2*i + 1: x1.xi + t1 ti
};

constexpr int DOT_CHUNK = 8192;
constexpr int DOT_THREADS = 256;

//! partial[b] = sum over chunk b of a[i]*b[i]; fixed tree => deterministic
struct DotPartial {
    const double* a;
    const double* b;
    uint64_t n;
    double* partial;
    SANM_D void operator()(const rt::BlockCtx& c) const {
        double* red = reinterpret_cast<double*>(c.smem);
        const uint64_t i0 = c.bid * DOT_CHUNK;
        const uint64_t i1 = i0 + DOT_CHUNK < n ? i0 + DOT_CHUNK : n;
        double s = 0;
        for (uint64_t i = i0 + c.tid; i < i1; i += c.nthr) s += a[i] * b[i];
        red[c.tid] = s;
        c.sync();
        for (int w = c.nthr / 2; w > 0; w >>= 1) {
            if (c.tid < w) red[c.tid] += red[c.tid + w];
            c.sync();
        }
        if (c.tid == 0) partial[c.bid] = red[0];
    }
};

struct DotFinal {
    const double* partial;
    uint64_t nblk;
    double* out;
    SANM_D void operator()(const rt::BlockCtx& c) const {
        double* red = reinterpret_cast<double*>(c.smem);
        double s = 0;
        for (uint64_t i = c.tid; i < nblk; i += c.nthr) s += partial[i];
        red[c.tid] = s;
        c.sync();
        for (int w = c.nthr / 2; w > 0; w >>= 1) {
            if (c.tid < w) red[c.tid] += red[c.tid + w];
            c.sync();
        }
        if (c.tid == 0) out[0] = red[0];
    }
};

//! per-chunk max of |v0 - v1| / max(min(|v0|, |v1|), 1) with v1 = -(gt*ti + bi)
//! (the metric of TensorND::assert_allclose, libsanm/tensor.cpp:670-684)
struct ResidualPartial {
    const double* ax;
    const double* gt;
    const double* bi;  // may be null (order 1: bias is zero)
    const double* sc;
    uint64_t n;
    double* partial;
    SANM_D void operator()(const rt::BlockCtx& c) const {
        double* red = reinterpret_cast<double*>(c.smem);
        const double ti = sc[SC_TI];
        const uint64_t i0 = c.bid * DOT_CHUNK;
        const uint64_t i1 = i0 + DOT_CHUNK < n ? i0 + DOT_CHUNK : n;
        double m = 0;
        for (uint64_t i = i0 + c.tid; i < i1; i += c.nthr) {
            double v0 = ax[i], v1 = -(gt[i] * ti + (bi ? bi[i] : 0.0));
            double a0 = fabs(v0), a1 = fabs(v1);
            double den = a0 < a1 ? a0 : a1;
            den = den > 1 ? den : 1;
            double e = fabs(v0 - v1) / den;
            // NaN must not be lost by the max
            m = (e > m || e != e) ? e : m;
        }
        red[c.tid] = m;
        c.sync();
        for (int w = c.nthr / 2; w > 0; w >>= 1) {
            if (c.tid < w) {
                double o = red[c.tid + w];
                if (o > red[c.tid] || o != o) red[c.tid] = o;
            }
            c.sync();
        }
        if (c.tid == 0) partial[c.bid] = red[0];
    }
};

struct MaxFinal {
    const double* partial;
    uint64_t nblk;
    double* out;
    SANM_D void operator()(const rt::BlockCtx& c) const {
        if (c.tid == 0) {
            double m = 0;
            for (uint64_t i = 0; i < nblk; ++i) {
                double o = partial[i];
                if (o > m || o != o) m = o;
            }
            out[0] = m;
        }
    }
};

struct ScalarT1 {
    double* sc;
    SANM_D void operator()(size_t) const {
        // t1 = 1 / sqrt(|xgt|^2 + 1), libsanm/anm.cpp:245
        double t1 = 1.0 / sqrt(sc[SC_DOT] + 1.0);
        sc[SC_T1] = t1;
        sc[SC_TI] = t1;
    }
};

struct ScalarTi {
    double* sc;
    SANM_D void operator()(size_t) const {
        // ti = xbi . x1 / (t1 - xgt . x1), libsanm/anm.cpp:254
        sc[SC_TI] = sc[SC_DOT] / (sc[SC_T1] - sc[SC_XGT_X1]);
    }
};

struct ScalarXdot {
    double* sc;
    int slot;
    SANM_D void operator()(size_t) const {
        sc[slot] = sc[SC_DOT] + sc[SC_T1] * sc[SC_TI];
    }
};

struct ScalarCopy {
    double* sc;
    int dst, src;
    SANM_D void operator()(size_t) const { sc[dst] = sc[src]; }
};

//! x_i = -ti * xgt - xbi ; t_i appended  (libsanm/anm.cpp:261-264)
struct CombineX {
    double* xt;  // [n + 1]
    const double* xgt;
    const double* xbi;  // may be null (order 1)
    const double* sc;
    uint64_t n;
    SANM_D void operator()(size_t i) const {
        const double ti = sc[SC_TI];
        if (i == n) {
            xt[i] = ti;
        } else {
            xt[i] = xgt[i] * (-ti) - (xbi ? xbi[i] : 0.0);
        }
    }
};

struct AddVec {
    double* out;
    const double* a;
    const double* b;
    SANM_D void operator()(size_t i) const { out[i] = a[i] + b[i]; }
};

/* ---- K assembly */
struct AssembleK {
    const uint32_t *k_rowptr, *k_colidx;
    const uint32_t *o_rowptr, *o_idx;
    const double* o_coeff;
    const uint32_t *i_rowptr, *i_idx;
    const double* i_coeff;
    const double* jac;
    int64_t jac_pitch;  //!< 0: [N][odim][idim]; else [odim*idim][pitch]
    double* vals;
    double* grad_t;
    uint32_t n;
    int odim, idim, elemwise;

    SANM_D double jget(uint32_t e, uint32_t o, int ii) const {
        return jac_pitch ? jac[(int64_t)(o * idim + ii) * jac_pitch + e]
                         : jac[((uint64_t)e * odim + o) * idim + ii];
    }

    SANM_D void add(uint32_t row, uint32_t col, double v, double& gt) const {
        if (col >= n) {
            gt += v;
            return;
        }
        uint32_t lo = k_rowptr[row], hi = k_rowptr[row + 1];
        while (lo + 1 < hi) {
            uint32_t mid = (lo + hi) >> 1;
            if (k_colidx[mid] <= col) {
                lo = mid;
            } else {
                hi = mid;
            }
        }
        vals[lo] += v;
    }

    SANM_D void operator()(size_t row) const {
        for (uint32_t p = k_rowptr[row]; p < k_rowptr[row + 1]; ++p) vals[p] = 0;
        double gt = 0;
        for (uint32_t p = o_rowptr[row]; p < o_rowptr[row + 1]; ++p) {
            const uint32_t fi = o_idx[p];
            const uint32_t e = fi / odim, o = fi % odim;
            const double w = o_coeff[p];
            if (elemwise) {
                const double g = jac[(uint64_t)e * odim + o] * w;
                const uint32_t r2 = e * idim + o;
                for (uint32_t q = i_rowptr[r2]; q < i_rowptr[r2 + 1]; ++q) {
                    add((uint32_t)row, i_idx[q], g * i_coeff[q], gt);
                }
            } else {
                for (int ii = 0; ii < idim; ++ii) {
                    const double g = jget(e, o, ii) * w;
                    const uint32_t r2 = e * idim + ii;
                    for (uint32_t q = i_rowptr[r2]; q < i_rowptr[r2 + 1]; ++q) {
                        add((uint32_t)row, i_idx[q], g * i_coeff[q], gt);
                    }
                }
            }
        }
        if (grad_t) grad_t[row] = gt;
    }
};

}  // namespace

/* ======================= KAssembler ======================= */

KAssembler::KAssembler(std::shared_ptr<const Remap> rin,
                       std::shared_ptr<const Remap> rout, uint64_t n,
                       uint64_t batch, int odim, int idim, bool elemwise,
                       std::shared_ptr<const Remap> asm_rin,
                       std::shared_ptr<const Remap> asm_rout)
        : m_rin{std::move(rin)},
          m_rout{std::move(rout)},
          m_asm_rin{asm_rin ? std::move(asm_rin) : m_rin},
          m_asm_rout{asm_rout ? std::move(asm_rout) : m_rout},
          m_n{n},
          m_batch{batch},
          m_odim{odim},
          m_idim{idim},
          m_elemwise{elemwise} {
    sanm_check(m_rout->nr_out() == n, "remap_out has %zu outputs, expected %zu",
               (size_t)m_rout->nr_out(), (size_t)n);
    sanm_check(m_rout->nr_in() == batch * odim);
    sanm_check(m_rin->nr_out() == batch * idim);
    // the pattern depends on the structure of the two maps only
    static LruCache<const Pattern> cache{6};
    uint64_t key = m_rin->hash_struct() * 0x9E3779B97F4A7C15ull ^
                   m_rout->hash_struct();
    key = hash_bytes(&key, 8, n * 131 + odim * 17 + idim * 3 + elemwise);
    const auto& orp = m_rout->h_rowptr();
    const auto& oix = m_rout->h_idx();
    const auto& irp = m_rin->h_rowptr();
    const auto& iix = m_rin->h_idx();
    // independent witness of the inputs: sizes + 128 strided samples per array
    std::vector<uint64_t> witness{n, (uint64_t)odim, (uint64_t)idim,
                                  (uint64_t)elemwise, orp.size(), oix.size(),
                                  irp.size(), iix.size()};
    for (const auto* v : {&orp, &oix, &irp, &iix}) {
        const size_t step = std::max<size_t>(1, v->size() / 128);
        for (size_t q = 0; q < v->size(); q += step) witness.push_back((*v)[q]);
    }
    m_pat = cache.get(key);
    if (m_pat && m_pat->rowptr.size() == n + 1 && m_pat->witness == witness) return;
    // structural pattern (independent of values): for every output row the
    // union, over the items it touches, of the unknowns those items read
    std::vector<std::vector<uint32_t>> rows(n);
#pragma omp parallel num_threads(::sanm_b200::rt::host_threads())
    {
        std::vector<uint32_t> items, cand;
#pragma omp for schedule(dynamic, 256)
        for (int64_t r = 0; r < (int64_t)n; ++r) {
            cand.clear();
            if (m_elemwise) {
                for (uint32_t p = orp[r]; p < orp[r + 1]; ++p) {
                    uint32_t e = oix[p] / odim, o = oix[p] % odim;
                    uint32_t r2 = e * idim + o;
                    for (uint32_t q = irp[r2]; q < irp[r2 + 1]; ++q) {
                        if (iix[q] < n) cand.push_back(iix[q]);
                    }
                }
            } else {
                items.clear();
                for (uint32_t p = orp[r]; p < orp[r + 1]; ++p) {
                    items.push_back(oix[p] / odim);
                }
                std::sort(items.begin(), items.end());
                items.erase(std::unique(items.begin(), items.end()), items.end());
                for (uint32_t e : items) {
                    for (uint32_t q = irp[(uint64_t)e * idim];
                         q < irp[(uint64_t)(e + 1) * idim]; ++q) {
                        if (iix[q] < n) cand.push_back(iix[q]);
                    }
                }
            }
            std::sort(cand.begin(), cand.end());
            cand.erase(std::unique(cand.begin(), cand.end()), cand.end());
            rows[r] = cand;
        }
    }
    auto pat = std::make_shared<Pattern>();
    auto& m_rowptr = pat->rowptr;
    auto& m_colidx = pat->colidx;
    m_rowptr.assign(n + 1, 0);
    for (uint64_t r = 0; r < n; ++r) {
        sanm_check(!rows[r].empty(), "empty row %zu", (size_t)r);
        sanm_check(m_rowptr[r] + rows[r].size() < (1ull << 32));
        m_rowptr[r + 1] = m_rowptr[r] + (uint32_t)rows[r].size();
    }
    m_colidx.resize(m_rowptr[n]);
    for (uint64_t r = 0; r < n; ++r) {
        std::copy(rows[r].begin(), rows[r].end(), m_colidx.begin() + m_rowptr[r]);
    }
    pat->d_rowptr.upload(m_rowptr);
    pat->d_colidx.upload(m_colidx);
    pat->hash = hash_vec(m_colidx, hash_vec(m_rowptr, n));
    m_pat = pat;
    pat->witness = std::move(witness);
    cache.put(key, m_pat);
}

void KAssembler::assemble(double* vals, double* grad_t, const double* jac,
                          int64_t jac_pitch) const {
    sanm_check(!jac_pitch || !m_elemwise);
    rt::for_each(m_n, AssembleK{m_pat->d_rowptr.ptr(), m_pat->d_colidx.ptr(),
                                m_asm_rout->d_rowptr(), m_asm_rout->d_idx(),
                                m_asm_rout->d_coeff(), m_asm_rin->d_rowptr(),
                                m_asm_rin->d_idx(), m_asm_rin->d_coeff(), jac,
                                jac_pitch, vals,
                                grad_t, (uint32_t)m_n, m_odim, m_idim,
                                (int)m_elemwise});
}

/* ======================= AnmSolver ======================= */

AnmSolver::AnmSolver(Kind kind, std::shared_ptr<const Graph> graph, int f,
                     std::shared_ptr<const Remap> remap_inp,
                     std::shared_ptr<const Remap> remap_out, const double* x0,
                     const Shape& x_shape, double t0, const double* v,
                     const Shape& v_shape, const AnmHyper& hyper)
        : m_kind{kind},
          m_hyper{hyper},
          m_graph{std::move(graph)},
          m_rin{std::move(remap_inp)},
          m_rout{std::move(remap_out)},
          m_x_shape{x_shape},
          m_n{x_shape.total()},
          m_ld{(x_shape.total() + 1 + 3) / 4 * 4},
          m_max_a_bound{poly::stable_x_range(hyper.order)} {
    rt::check_device();
    sanm_check(hyper.order >= 2 && m_rin && m_rout, "order=%d", hyper.order);
    sanm_check(hyper.order <= taylor::MAX_ORDER, "order %d above %d",
               hyper.order, taylor::MAX_ORDER);
    sanm_check(hyper.xcoeff_l2_penalty >= 0, "xcoeff_l2_penalty=%g",
               hyper.xcoeff_l2_penalty);
    if (kind == IMPLICIT) {
        // libsanm/anm.cpp:501-507
        sanm_check(m_rin->in_shape().rank == 1 && m_rout->out_shape().rank == 1 &&
                   m_rin->in_shape().dim[0] == m_rout->out_shape().dim[0] + 1);
        sanm_check(x_shape == m_rout->out_shape());
    } else {
        // libsanm/anm.cpp:330-340, 451-458
        sanm_check(m_rin->in_shape() == x_shape,
                   "linear map expects %s, got x0 %s",
                   m_rin->in_shape().str().c_str(), x_shape.str().c_str());
        sanm_check(x_shape.total() == v_shape.total(),
                   "currently we assume the system is a full-rank mapping");
        sanm_check(m_rout->out_shape() == v_shape,
                   "output shapes mismatch: %s vs %s",
                   m_rout->out_shape().str().c_str(), v_shape.str().c_str());
        m_v_host.assign(v, v + m_n);
    }
    // multi-GPU: this rank propagates a contiguous range of the batch
    // (libsanm/symbolic.cpp:525-529); nodal vectors stay global
    m_graph_e = m_graph;
    m_rin_e = m_rin;
    m_rout_e = m_rout;
    {
        const uint64_t N = m_rin->out_shape().dim[0];
        if (comm::world() > 1 && N >= (uint64_t)comm::world() &&
            m_rout->in_shape().dim[0] == N) {
            uint64_t lo, hi;
            comm::shard_range(N, comm::rank(), comm::world(), &lo, &hi);
            m_graph_e = shard_graph(*m_graph, N, lo, hi);
            m_rin_e = shard_remap_rows(*m_rin, N, lo, hi);
            m_rout_e = shard_remap_cols(*m_rout, N, lo, hi);
            m_sharded = true;
        }
    }
    m_prop.reset(new TaylorProp{m_graph_e, f, hyper.order});

    m_xt0.assign(x0, x0 + m_n);
    m_xt0.push_back(kind == EQN ? 0.0 : t0);

    const int K = hyper.order;
    m_d_xt.alloc((size_t)(K + 1) * m_ld);
    m_h_xt.assign((size_t)(K + 1) * m_ld, 0.0);
    m_d_v.alloc(m_n);
    m_d_grad_t.alloc(m_n);
    m_d_bi.alloc(m_n);
    m_d_xbi.alloc(m_n);
    m_d_xgt.alloc(m_n);
    m_d_fx.alloc(m_n);
    m_d_tmp.alloc(m_n);
    m_d_sc.alloc(SC_CHECK0 + 2 * (K + 2));
    m_d_partial.alloc((m_ld + DOT_CHUNK - 1) / DOT_CHUNK + 1);
    if (kind == VECSCALE) m_d_v.upload(m_v_host.data(), m_n);
    if (kind == EQN) m_d_y.upload(m_v_host.data(), m_n);
    solve_expansion_coeffs();
}

AnmSolver::~AnmSolver() = default;

void AnmSolver::lu_solve(double* x, const double* b) {
    if (m_neq) {
        // libsanm/sparse_solver.cpp:163-175: right-hand side K'b
        m_neq->apply_transposed(m_d_kvals.ptr(), b, m_d_atb.ptr());
        m_lu.solve(x, m_d_atb.ptr());
    } else {
        m_lu.solve(x, b);
    }
}

void AnmSolver::dot_to(int slot, const double* a, const double* b, uint64_t n) {
    const uint64_t nblk = (n + DOT_CHUNK - 1) / DOT_CHUNK;
    rt::for_each_block<DOT_THREADS>(nblk, DOT_THREADS, DOT_THREADS * sizeof(double),
                                    DotPartial{a, b, n, m_d_partial.ptr()});
    rt::for_each_block<DOT_THREADS>(1, DOT_THREADS, DOT_THREADS * sizeof(double),
                       DotFinal{m_d_partial.ptr(), nblk, m_d_sc.ptr() + slot});
}

std::vector<const double*> AnmSolver::coeff_ptrs() const {
    std::vector<const double*> p(m_nr_coeffs);
    for (size_t i = 0; i < m_nr_coeffs; ++i) p[i] = m_h_xt.data() + i * m_ld;
    return p;
}

void AnmSolver::ensure_host_series() const {
    if (m_h_xt_valid) return;
    rt::d2h(m_h_xt.data(), m_d_xt.ptr(), m_nr_coeffs * m_ld * sizeof(double));
    m_h_xt_valid = true;
}

const double* AnmSolver::xt_coeff(size_t i) const {
    sanm_check(i < m_nr_coeffs);
    ensure_host_series();
    return m_h_xt.data() + i * m_ld;
}

bool AnmSolver::on_fx0_computed() {
    // m_d_fx holds f(x0) mapped to the unknowns
    if (m_kind == EQN) {
        // libsanm/anm.cpp:480-491
        if (m_converged) return false;
        rt::for_each(m_n, AddVec{m_d_v.ptr(), m_d_fx.ptr(), m_d_y.ptr()});
        dot_to(SC_RMS2, m_d_v.ptr(), m_d_v.ptr(), m_n);
        double s2 = 0;
        rt::d2h(&s2, m_d_sc.ptr() + SC_RMS2, sizeof(double));
        m_residual_rms = std::sqrt(s2 / (double)m_n);
        if (m_residual_rms < m_hyper.converge_rms) {
            m_converged = true;
            return false;
        }
        return true;
    }
    std::vector<double> fx(m_n);
    m_d_fx.download(fx.data(), m_n);
    if (m_kind == VECSCALE) {
        // check_t0v_match, libsanm/anm.cpp:343-360
        const double t0v = t0();
        for (uint64_t i = 0; i < m_n; ++i) {
            double a = fx[i], b = m_v_host[i] * t0v;
            double maxerr = std::max(std::min(std::fabs(a), std::fabs(b)), 1.0) *
                            m_hyper.solution_check_tol;
            if (!(std::fabs(a + b) <= maxerr)) {
                m_fx_check_failed = true;
                sanm_throw(SANM_ERR_NUMERICAL,
                           "f(x0)+t0*v is not zero: lhs=%g rhs=%g idx=%zu iter=%zu",
                           a, b, (size_t)i, m_iter);
            }
        }
        return true;
    }
    // IMPLICIT, libsanm/anm.cpp:509-518
    if (m_fx0_host.empty()) {
        m_fx0_host = fx;
    } else {
        for (uint64_t i = 0; i < m_n; ++i) {
            double v0 = m_fx0_host[i], v1 = fx[i];
            double maxerr = m_hyper.solution_check_tol *
                            std::max(std::min(std::fabs(v0), std::fabs(v1)), 1.0);
            if (!(std::fabs(v0 - v1) < maxerr)) m_fx_check_failed = true;
            sanm_check(std::fabs(v0 - v1) < maxerr,
                       "check f(x0, t0)=f(x, t): value mismatch at index %zu: "
                       "v0=%g v1=%g (eps=%g)",
                       (size_t)i, v0, v1, m_hyper.solution_check_tol);
        }
    }
    return true;
}

void AnmSolver::solve_expansion_coeffs() {
    SANM_PROF("solve_expansion_coeffs");
    const int K = m_hyper.order;
    const uint64_t n = m_n;
    m_pade.reset();
    m_t_coeffs.assign(1, m_xt0[n]);

    // x0 -> device series slot 0 and host series
    std::memcpy(m_h_xt.data(), m_xt0.data(), (n + 1) * sizeof(double));
    rt::h2d(m_d_xt.ptr(), m_xt0.data(), (n + 1) * sizeof(double));
    m_nr_coeffs = 1;
    m_h_xt_valid = true;

    const Shape elem_shape = m_rin_e->out_shape();
    m_prop->prepare(elem_shape);
    sanm_check(m_rout_e->nr_in() == m_prop->batch() * m_prop->out_elems(),
               "remap_out input size does not match the function output");
    // fused I/O: gather and assembly happen inside the element kernels
    const bool fio = m_prop->attach_io(*m_rin_e, *m_rout_e);
    // sum over the ranks of a nodal vector assembled from the local tets
    auto reduce = [&](double* v, size_t len) {
        if (m_sharded) comm::allreduce_sum(v, len);
    };
    const VecLayout soa_in{m_prop->in_elems(), m_prop->npad()};
    const VecLayout soa_out{m_prop->out_elems(), m_prop->npad()};
    if (fio) {
        SANM_PROF("push_xi");
        m_prop->push_xi_from(m_d_xt.ptr(), m_d_fx.ptr());
    } else {
        {
            SANM_PROF("push_xi");
            m_rin_e->apply(m_prop->input_slot(0), soa_in, m_d_xt.ptr(),
                           VecLayout{});
            m_prop->push_xi_filled();
        }
        SANM_PROF("remap_out");
        m_rout_e->apply(m_d_fx.ptr(), VecLayout{}, m_prop->output_slot(0),
                        soa_out);
    }
    reduce(m_d_fx.ptr(), n);
    m_prop->check_device_error();
    if (!on_fx0_computed()) return;

    double* sc = m_d_sc.ptr();
    const double* grad_t = m_kind == IMPLICIT ? m_d_grad_t.ptr() : m_d_v.ptr();
    double* x1 = m_d_xt.ptr() + m_ld;
    const uint64_t nblk = (n + DOT_CHUNK - 1) / DOT_CHUNK;

    for (int i = 1; i <= K; ++i) {
        double* xti = m_d_xt.ptr() + (size_t)i * m_ld;
        {
            SANM_PROF("taylor_next_order");
            if (fio) {
                m_prop->compute_next_order_bias_to(m_d_bi.ptr());
            } else {
                m_prop->compute_next_order_bias();
            }
        }
        if (i == 1) {
            // the first-order bias is zero (libsanm/anm.cpp:236); build and
            // factorise K = d(remap_out . f . remap_in)/dx
            {
                SANM_PROF("build_sparse_coeff");
                if (!m_kasm) {
                    SANM_PROF("K_symbolic_host");
                    m_kasm.reset(new KAssembler{
                            m_rin, m_rout, n, m_rin->out_shape().dim[0],
                            m_prop->out_elems(), m_prop->in_elems(),
                            m_prop->jacobian_type() == TaylorProp::JAC_ELEMWISE,
                            m_rin_e, m_rout_e});
                    // The element-wise assembly lists and the symbolic LU
                    // analysis both depend on the K pattern only: the host
                    // half of the lists runs on a helper thread beside the
                    // analysis (on Bob 26 ms beside 39 ms).
                    static LruCache<const FusedKAssembly> kcache{6};
                    uint64_t kkey = 0;
                    std::future<FusedKAssembly::HostLists> lists;
                    if (fio && m_prop->jacobian_pitch()) {
                        const FusedIO& io = *m_prop->fused_io();
                        kkey = io.hash * 0x9E3779B97F4A7C15ull ^ m_kasm->pattern_hash();
                        kkey = hash_bytes(&kkey, 8, n);
                        m_kfused = io.hash ? kcache.get(kkey) : nullptr;
                        if (m_kfused && !m_kfused->matches(io.n_tets, n, m_kasm->nnz())) {
                            m_kfused.reset();  // hash collision
                        }
                        if (!m_kfused) {
                            const KAssembler* ka = m_kasm.get();
                            lists = std::async(std::launch::async, [&io, n, ka] {
                                return FusedKAssembly::build_host(io, n, ka->rowptr(),
                                                                  ka->colidx());
                            });
                        }
                    }
                    {
                        SANM_PROF("lu_analyze_host");
                        if (m_hyper.xcoeff_l2_penalty) {
                            // libsanm/sparse_solver.cpp:366-395: factor K'K + p I
                            m_neq.reset(new NormalEquations);
                            m_neq->analyze(n, m_kasm->rowptr(), m_kasm->colidx());
                            m_lu.analyze(n, m_neq->rowptr2().data(),
                                         m_neq->colidx2().data());
                            m_d_k2vals.alloc(m_neq->nnz2());
                            m_d_atb.alloc(n);
                        } else {
                            // ordering hint: the current point (in `fea`
                            // the vertex coordinates of the unknowns)
                            m_lu.analyze(n, m_kasm->rowptr().data(),
                                         m_kasm->colidx().data(), m_xt0.data());
                        }
                    }
                    m_d_kvals.alloc(m_kasm->nnz());
                    if (lists.valid()) {
                        SANM_PROF("K_elem_lists_host");  // what is left of it
                        m_kfused = FusedKAssembly::finish(lists.get());
                        if (m_kfused && m_prop->fused_io()->hash) {
                            kcache.put(kkey, m_kfused);
                        }
                    }
                }
                SANM_PROF("k_assemble_K");
                double* gt = m_kind == IMPLICIT ? m_d_grad_t.ptr() : nullptr;
                if (m_kfused) {
                    m_kfused->assemble(m_d_kvals.ptr(), gt,
                                       m_prop->ensure_jacobian(),
                                       m_prop->jacobian_pitch(),
                                       *m_prop->fused_io());
                } else {
                    m_kasm->assemble(m_d_kvals.ptr(), gt,
                                     m_prop->ensure_jacobian(),
                                     m_prop->jacobian_pitch());
                }
                reduce(m_d_kvals.ptr(), m_kasm->nnz());
                if (gt) reduce(gt, n);
            }
            {
                SANM_PROF("sparse_prep");
                if (m_neq) {
                    m_neq->form(m_d_kvals.ptr(), m_hyper.xcoeff_l2_penalty,
                                m_d_k2vals.ptr());
                    m_lu.factor(m_d_k2vals.ptr());
                } else {
                    m_lu.factor(m_d_kvals.ptr());
                }
            }
            {
                SANM_PROF("sparse_solve");
                lu_solve(m_d_xgt.ptr(), grad_t);
            }
            dot_to(SC_DOT, m_d_xgt.ptr(), m_d_xgt.ptr(), n);
            rt::for_each(1, ScalarT1{sc});
            rt::for_each(n + 1, CombineX{xti, m_d_xgt.ptr(), nullptr, sc, n});
            dot_to(SC_XGT_X1, x1, m_d_xgt.ptr(), n);
        } else {
            if (!fio) {
                SANM_PROF("remap_out");
                m_rout_e->apply(m_d_bi.ptr(), VecLayout{},
                                m_prop->output_slot(i), soa_out);
            }
            reduce(m_d_bi.ptr(), n);
            {
                SANM_PROF("sparse_solve");
                lu_solve(m_d_xbi.ptr(), m_d_bi.ptr());
            }
            dot_to(SC_DOT, m_d_xbi.ptr(), x1, n);
            rt::for_each(1, ScalarTi{sc});
            rt::for_each(n + 1, CombineX{xti, m_d_xgt.ptr(), m_d_xbi.ptr(), sc, n});
        }
        if (m_hyper.sanity_check && !m_hyper.xcoeff_l2_penalty) {
            // libsanm/anm.cpp:266-285, evaluated after the loop
            SANM_PROF("anm_sanity_check");
            m_lu.apply(m_d_tmp.ptr(), xti);
            rt::for_each_block<DOT_THREADS>(
                    nblk, DOT_THREADS, DOT_THREADS * sizeof(double),
                    ResidualPartial{m_d_tmp.ptr(), grad_t,
                                    i == 1 ? nullptr : m_d_bi.ptr(), sc, n,
                                    m_d_partial.ptr()});
            rt::for_each_block<32>(1, 32, 0,
                               MaxFinal{m_d_partial.ptr(), nblk,
                                        sc + SC_CHECK0 + 2 * i});
            dot_to(SC_DOT, x1, xti, n);
            rt::for_each(1, ScalarXdot{sc, SC_CHECK0 + 2 * i + 1});
        }
        if (i < K) {
            SANM_PROF("push_xi");
            if (fio) {
                m_prop->push_xi_from(xti);
            } else {
                m_rin_e->apply(m_prop->input_slot(i), soa_in, xti, VecLayout{});
                m_prop->push_xi_filled();
            }
        }
    }
    m_stats.tet_order_steps += (double)m_prop->batch() * K;

    // The series stays on the device: its norms, t coefficients
    // and a finiteness flag are reduced there, and the Pade passes read it in
    // place (pade_dev.cu).  SANM_HOST_PADE=1: one copy of the whole series to
    // the host and the host passes of pade.cpp instead.
    m_nr_coeffs = K + 1;
    m_h_xt_valid = false;
    m_prop->check_device_error();
    m_series_on_device = getenv("SANM_HOST_PADE") == nullptr &&
                         device_pade_supported(K + 1);
    if (m_series_on_device) {
        SANM_PROF("series_summary");
        int bad = -1;
        series_dev::summary(m_d_xt.ptr(), K + 1, m_ld, n + 1, m_norm2, m_t_coeffs,
                            &bad);
        sanm_check(bad < 0, "non-finite value in series coefficient %d", bad);
    } else {
        ensure_host_series();
        m_norm2.assign(K + 1, 0.0);
        m_t_coeffs.resize(K + 1);
        for (int i = 0; i <= K; ++i) {
            const double* p = m_h_xt.data() + (size_t)i * m_ld;
            double s2 = 0;
            for (uint64_t j = 0; j <= n; ++j) {
                sanm_check(std::isfinite(p[j]),
                           "non-finite value in series coefficient %d", i);
                s2 += p[j] * p[j];
            }
            m_norm2[i] = s2;
            m_t_coeffs[i] = p[n];
        }
    }
    if (m_hyper.sanity_check && !m_hyper.xcoeff_l2_penalty) {
        std::vector<double> chk(2 * (K + 2));
        rt::d2h(chk.data(), sc + SC_CHECK0, chk.size() * sizeof(double));
        for (int i = 1; i <= K; ++i) {
            sanm_check(chk[2 * i] < 1e-4,
                       "ANM check coeff eqn: order %d: relative mismatch %g", i,
                       chk[2 * i]);
            double xdot = chk[2 * i + 1];
            if (i == 1) {
                sanm_check(std::fabs(xdot - 1) < 1e-4, "xdot=%g", xdot);
            } else {
                sanm_check(std::fabs(xdot) < 1e-4, "i=%d: xdot=%g", i, xdot);
            }
        }
    }
    estimate_valid_range();
    static const bool verbose = getenv("SANM_VERBOSE") != nullptr;
    if (verbose) {
        // same trace as libsanm/anm.cpp:295-309
        printf("\nbound=%g t=%g\n", m_t_max_a, m_t_max);
        printf("x(a):");
        for (int i = 0; i <= K; ++i) printf(" %.3g", std::sqrt(m_norm2[i]));
        printf("\nt(a):");
        for (double t : m_t_coeffs) printf(" %.3g,", t);
        printf("\n");
        fflush(stdout);
    }
    ++m_iter;
}

void AnmSolver::estimate_valid_range() {
    // libsanm/anm.cpp:117-154
    const int K = m_hyper.order;
    const uint64_t len = m_n + 1;
    auto norm = [&](size_t i) { return std::sqrt(m_norm2[i]); };
    double x1 = norm(1), xback = std::max(norm(K), 1e-15);
    double a_bound = std::pow(m_hyper.maxr / xback * x1, 1.0 / double(K - 1));
    a_bound = std::min(a_bound, m_max_a_bound);
    sanm_check((int)m_t_coeffs.size() == K + 1 && m_t_coeffs[1] > 0);
    m_t_max_a = a_bound;
    m_a_bound_taylor = a_bound;
    m_t_max = poly::eval(m_t_coeffs, a_bound);
    sanm_check(m_t_max > m_t_coeffs[0],
               "t does not incr at iter %zu: t0=%g tmax=%g bound=%g", m_iter,
               m_t_coeffs[0], m_t_max, a_bound);
    static const bool env_pade = getenv("SANM_PADE") != nullptr;
    if ((m_hyper.use_pade || env_pade) && a_bound < m_max_a_bound) {
        SANM_PROF("pade");
        if (m_series_on_device) {
            m_pade.reset(new PadeApproximation{
                    make_device_pade_ops(m_d_xt.ptr(), K + 1, m_ld, len),
                    !m_hyper.xcoeff_l2_penalty, false});
        } else {
            m_pade.reset(new PadeApproximation{coeff_ptrs(), len,
                                               !m_hyper.xcoeff_l2_penalty, false});
        }
        if (m_pade->estimate_valid_range(a_bound, m_hyper.maxr, m_max_a_bound)) {
            m_t_max_a = m_pade->get_t_max_a();
            m_t_max = m_pade->get_t_max();
        } else {
            m_pade.reset();
        }
    }
}

std::vector<double> AnmSolver::eval_xt(double a) const {
    // libsanm/anm.cpp:167-173, unary_polynomial.cpp:118-129
    if (m_pade) return m_pade->eval_xt(a);
    const uint64_t len = m_n + 1;
    sanm_check(m_nr_coeffs >= 1);
    if (m_series_on_device && m_nr_coeffs > 1 && !m_h_xt_valid) {
        std::vector<double> ret(len);
        series_dev::horner(m_d_xt.ptr(), m_nr_coeffs, m_ld, len, a, ret.data());
        return ret;
    }
    ensure_host_series();
    std::vector<double> ret(m_h_xt.begin() + (m_nr_coeffs - 1) * m_ld,
                            m_h_xt.begin() + (m_nr_coeffs - 1) * m_ld + len);
    for (size_t i = m_nr_coeffs - 1; i-- > 0;) {
        const double* f = m_h_xt.data() + i * m_ld;
        for (uint64_t j = 0; j < len; ++j) ret[j] = ret[j] * a + f[j];
    }
    return ret;
}

void AnmSolver::update_approx() {
    // libsanm/anm.cpp:156-159
    const bool had_pade = (bool)m_pade;
    const double a_taylor = m_a_bound_taylor;
    const std::vector<double> xt_old(m_h_xt.begin(), m_h_xt.begin() + m_n + 1);
    const size_t nr_old = m_nr_coeffs;
    m_xt0 = eval_xt(m_t_max_a);
    m_fx_check_failed = false;
    // Reference behaviour (libsanm/anm.cpp:343-360, 509-518): a failed
    // consistency check of the new expansion is an error.  Opt-in recovery
    // (hyper-parameter-like switch SANM_PADE_RECOVER=1, not in the reference):
    // the Pade range estimate (pade.cpp:107-173) is a discontinuous function of
    // the series and can accept a range beyond the radius of convergence; with
    // recovery on, the step is retried from the range of the plain series
    // (libsanm/anm.cpp:118-122).  Every recovery is counted
    // (sanm_anm_robustness_stats) so that such a run cannot pass for a clean one.
    const bool recover = getenv("SANM_PADE_RECOVER") != nullptr;  // once per ANM iteration
    // (the retry re-evaluates the OLD series, which the failed expansion
    // overwrites on the device: keep its host copy)
    if (recover) ensure_host_series();
#ifdef SANM_HOSTEMU
    // test-only (host-emulation build): take the recovery path once without a
    // failing check
    const char* inject = getenv("SANM_INJECT_PADE_REJECT");
#else
    const char* inject = nullptr;
#endif
    try {
        if (inject && inject[0] == '1' && had_pade && !m_nr_pade_rejected) {
            m_fx_check_failed = true;
            sanm_throw(SANM_ERR_ASSERT, "injected Pade rejection");
        }
        solve_expansion_coeffs();
        return;
    } catch (Error&) {
        if (!m_fx_check_failed || !had_pade || !recover) throw;
    }
    fprintf(stderr,
            "sanm_b200: Pade range a=%g rejected by the f(x,t) check at "
            "iteration %zu; retrying with the series range a=%g\n",
            m_t_max_a, m_iter, a_taylor);
    std::copy(xt_old.begin(), xt_old.end(), m_h_xt.begin());
    m_nr_coeffs = nr_old;
    m_pade.reset();
    m_xt0 = eval_xt(a_taylor);
    ++m_nr_pade_rejected;
    solve_expansion_coeffs();
}

void AnmSolver::k_csr(uint64_t* n, uint64_t* nnz, uint32_t* rowptr,
                      uint32_t* colidx, double* vals) const {
    sanm_check(m_kasm, "k_csr() before the first expansion");
    const auto& rp = m_kasm->rowptr();
    const auto& ci = m_kasm->colidx();
    *n = rp.size() - 1;
    *nnz = ci.size();
    if (rowptr) std::memcpy(rowptr, rp.data(), rp.size() * sizeof(uint32_t));
    if (colidx) std::memcpy(colidx, ci.data(), ci.size() * sizeof(uint32_t));
    if (vals) rt::d2h(vals, m_d_kvals.ptr(), ci.size() * sizeof(double));
}

void AnmSolver::eval(double a, double* x_out, double* t_out) const {
    std::vector<double> xt = eval_xt(a);
    std::memcpy(x_out, xt.data(), m_n * sizeof(double));
    *t_out = xt[m_n];
}

double AnmSolver::solve_a(double t) const {
    // libsanm/anm.cpp:175-191
    if (t == m_t_max) return m_t_max_a;
    if (m_pade) return m_pade->solve_a(t);
    sanm_check(t >= m_t_coeffs[0] && t < m_t_max);
    double l, r;
    if (m_t_max_a > 0) {
        l = 0;
        r = m_t_max_a;
    } else {
        l = -m_t_max_a;
        r = 0;
    }
    return poly::solve_eqn(m_t_coeffs, l, r, t);
}

void AnmSolver::next_iter() {
    // libsanm/anm.cpp:464-478
    sanm_check(m_kind == EQN);
    if (m_converged) return;
    double a = get_t_upper() >= 1 ? solve_a(1) : m_t_max_a;
    m_xt0 = eval_xt(a);
    m_xt0[m_n] = 0;
    solve_expansion_coeffs();
}

void AnmSolver::get_x(double* out) const {
    std::memcpy(out, m_xt0.data(), m_n * sizeof(double));
}

void AnmSolver::fx0(double* out) const {
    sanm_check(m_kind == IMPLICIT && !m_fx0_host.empty());
    std::memcpy(out, m_fx0_host.data(), m_n * sizeof(double));
}

}  // namespace sanm_b200
