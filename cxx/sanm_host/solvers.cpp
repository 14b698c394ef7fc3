// cxx/sanm_host/solvers.cpp -- drivers, sparse maps and the sparse solver of
// the libsanm-compatible host layer, as calls into include/sanm_b200.h.
#include "sanm_host/solvers.hpp"

#include "sanm_b200.h"

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>

namespace sanm {
using capi::check;
using capi::Dims;

// ------------------------------------------------------------ ScopedProfiler
namespace {
struct HostTag {
    double tot = 0, mn = 1e300, mx = 0;
    size_t nr = 0;
};
struct HostProfile {
    std::mutex mtx;
    std::map<std::string, HostTag> tags;
    ~HostProfile() { ScopedProfiler::report(); }
};
HostProfile& host_profile() {
    static HostProfile p;
    return p;
}
bool g_profile_reported = false;
}  // namespace

struct ScopedProfiler::Stat {};

ScopedProfiler::ScopedProfiler(const char* tag)
        : m_tag(tag), m_t0(std::chrono::steady_clock::now()) {
    host_profile();  // construct before the first report at exit
}

ScopedProfiler::~ScopedProfiler() {
    double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() -
                                              m_t0)
                        .count();
    auto& p = host_profile();
    std::lock_guard<std::mutex> lk{p.mtx};
    auto& t = p.tags[m_tag];
    t.tot += dt;
    t.mn = std::min(t.mn, dt);
    t.mx = std::max(t.mx, dt);
    ++t.nr;
}

ScopedProfiler::Stat* ScopedProfiler::get_node() {
    host_profile();  // arms the report printed at exit
    return nullptr;
}
void ScopedProfiler::set_thread_root_node(Stat*) {}
void ScopedProfiler::clear_thread_root_node() {}

void ScopedProfiler::report() {
    if (g_profile_reported)
        return;
    g_profile_reported = true;
    auto& p = host_profile();
    std::lock_guard<std::mutex> lk{p.mtx};
    std::string dev(1 << 16, '\0');
    bool have_dev = sanm_prof_dump(dev.data(), dev.size()) == SANM_OK &&
                    dev[0] != '\0';
    if (p.tags.empty() && !have_dev)
        return;
    printf("=========== begin profiling results ===========\n");
    for (auto& [name, t] : p.tags)
        printf("%-20s: tot=%.3f avg=%.3f nr=%zu min=%.3f max=%.3f\n",
               name.c_str(), t.tot, t.tot / t.nr, t.nr, t.mn, t.mx);
    if (have_dev)
        printf("--- device time per tag (libsanm_b200) ---\n%s", dev.c_str());
    printf("=========== end profiling results ===========\n");
}

Xorshift128pRng::Xorshift128pRng(uint64_t seed) {
    // two non-zero words drawn from mt19937_64 (libsanm/utils.cpp:71-80);
    // written out so that no <random> state layout is assumed
    constexpr int N = 312, M = 156;
    uint64_t mt[N];
    mt[0] = seed;
    for (int i = 1; i < N; ++i)
        mt[i] = 6364136223846793005ull * (mt[i - 1] ^ (mt[i - 1] >> 62)) + i;
    int idx = N;
    auto next = [&]() {
        if (idx >= N) {
            for (int i = 0; i < N; ++i) {
                uint64_t x = (mt[i] & 0xFFFFFFFF80000000ull) |
                             (mt[(i + 1) % N] & 0x7FFFFFFFull);
                uint64_t xa = x >> 1;
                if (x & 1)
                    xa ^= 0xB5026F5AA96619E9ull;
                mt[i] = mt[(i + M) % N] ^ xa;
            }
            idx = 0;
        }
        uint64_t y = mt[idx++];
        y ^= (y >> 29) & 0x5555555555555555ull;
        y ^= (y << 17) & 0x71D67FFFEDA60000ull;
        y ^= (y << 37) & 0xFFF7EEE000000000ull;
        y ^= y >> 43;
        return y;
    };
    do {
        m_s0 = next();
        m_s1 = next();
    } while (!m_s0 || !m_s1);
}

// ---------------------------------------------------------- SparseLinearDesc
capi::RemapHandle::~RemapHandle() {
    if (h)
        sanm_remap_destroy(h);
}

void SparseLinearDesc::init_multi_thread(size_t) {}

void SparseLinearDesc::set_device_remap(sanm_remap* h) const {
    m_dev = std::make_shared<capi::RemapHandle>(h);
}

sanm_remap* SparseLinearDesc::device_remap() const {
    if (m_dev)
        return m_dev->h;
    const TensorShape os = out_shape(), is = inp_shape();
    const size_t nout = os.total_nr_elems();
    std::vector<uint64_t> rowptr(nout + 1), idx;
    std::vector<double> coeff;
    const_cast<SparseLinearDesc*>(this)->init_multi_thread(1);
    // first pass sizes the arrays, second fills them (get() is cheap: the
    // mesh descriptors return spans into their own storage)
    size_t nnz = 0;
    for (size_t i = 0; i < nout; ++i)
        nnz += get(i, 0).size();
    idx.reserve(nnz);
    coeff.reserve(nnz);
    rowptr[0] = 0;
    for (size_t i = 0; i < nout; ++i) {
        for (const InputElem& e : get(i, 0)) {
            idx.push_back(e.idx);
            coeff.push_back(e.coeff);
        }
        rowptr[i + 1] = idx.size();
    }
    Dims od{os}, id{is};
    sanm_remap* h = nullptr;
    check(sanm_remap_create(od.rank, od.d, id.rank, id.d, rowptr.data(),
                            idx.data(), coeff.data(), &h));
    set_device_remap(h);
    return h;
}

TensorND SparseLinearDesc::apply(const TensorND& x) const {
    sanm_assert(x.shape().total_nr_elems() == inp_shape().total_nr_elems(),
                "shape mismatch: expect %s, got %s",
                inp_shape().str().c_str(), x.shape().str().c_str());
    TensorND y{out_shape()};
    check(sanm_remap_apply(device_remap(), x.ptr(), y.woptr()));
    return y;
}

namespace {
class IdentityDesc final : public SparseLinearDesc {
public:
    IdentityDesc(const TensorShape& in, const TensorShape& out)
            : m_in(in), m_out(out) {
        sanm_assert(in.total_nr_elems() == out.total_nr_elems(),
                    "identity map between %s and %s", in.str().c_str(),
                    out.str().c_str());
        Dims id{in}, od{out};
        sanm_remap* h = nullptr;
        check(sanm_remap_identity(id.rank, id.d, od.rank, od.d, &h));
        set_device_remap(h);
    }
    TensorShape out_shape() const override { return m_out; }
    TensorShape inp_shape() const override { return m_in; }
    Linear1d get(size_t dst_index, size_t thread_id) const override {
        // one slot per thread; valid until the next call from that thread
        if (m_slots.size() <= thread_id)
            m_slots.resize(thread_id + 1);
        m_slots[thread_id] = InputElem{1._fp, dst_index};
        return {&m_slots[thread_id], 1};
    }
    void init_multi_thread(size_t n) override { m_slots.resize(n); }

private:
    TensorShape m_in, m_out;
    mutable std::vector<InputElem> m_slots{1};
};
}  // namespace

SparseLinearDescPtr SparseLinearDesc::make_identity(
        const TensorShape& inp_shape, const Maybe<TensorShape>& out_shape) {
    return std::make_shared<IdentityDesc>(
            inp_shape, out_shape.valid() ? out_shape.val() : inp_shape);
}

SparseLinearDesc::Linear1d SparseLinearDescCompressed::get(size_t dst_index,
                                                           size_t) const {
    sanm_assert(dst_index < m_oidx_input_elem.size());
    auto [b, e] = m_oidx_input_elem[dst_index];
    return {m_all_input_elem.data() + b, m_all_input_elem.data() + e};
}

// ------------------------------------------------------------------ drivers
namespace {
sanm_anm_hyper_t to_c(const ANMDriverHelper::HyperParam& h,
                      fp_t converge_rms = 1e-5) {
    sanm_anm_hyper_t c;
    sanm_anm_hyper_default(&c);
    c.use_pade = h.use_pade;
    c.sanity_check = h.sanity_check;
    c.order = h.order;
    c.maxr = h.maxr;
    c.solution_check_tol = h.solution_check_tol;
    c.xcoeff_l2_penalty = h.xcoeff_l2_penalty;
    c.converge_rms = converge_rms;
    return c;
}
sanm_graph* graph_of(symbolic::VarNode* f) {
    sanm_assert(f);
    return f->owner_graph()->handle();
}
}  // namespace

ANMDriverHelper::ANMDriverHelper(symbolic::VarNode* f,
                                 SparseLinearDescPtr remap_inp,
                                 SparseLinearDescPtr remap_out,
                                 const TensorShape& x_shape,
                                 const HyperParam& hyper_param)
        : m_hyper_param(hyper_param),
          m_func_def(f),
          m_remap_inp(std::move(remap_inp)),
          m_remap_out(std::move(remap_out)),
          m_x_shape(x_shape),
          m_nr_unknown(x_shape.total_nr_elems()) {
    sanm_assert(m_remap_inp && m_remap_out && f);
}

ANMDriverHelper::~ANMDriverHelper() {
    if (m_h)
        sanm_anm_destroy(m_h);
}

bool ANMDriverHelper::verbose_mode() {
    static const bool v = std::getenv("SANM_VERBOSE") != nullptr;
    return v;
}

void ANMDriverHelper::update_approx() {
    check(sanm_anm_update_approx(m_h));
}

fp_t ANMDriverHelper::get_t_upper() const {
    double v;
    check(sanm_anm_get_t_upper(m_h, &v));
    return v;
}

fp_t ANMDriverHelper::solve_a(fp_t t) const {
    double a;
    check(sanm_anm_solve_a(m_h, t, &a));
    return a;
}

std::pair<TensorND, fp_t> ANMDriverHelper::eval(fp_t a) const {
    TensorND x{m_x_shape};
    double t;
    check(sanm_anm_eval(m_h, a, x.woptr(), &t));
    return {x, t};
}

size_t ANMDriverHelper::get_nr_ieter() const {
    uint64_t n;
    check(sanm_anm_nr_iter(m_h, &n));
    return n;
}

std::span<const TensorND> ANMDriverHelper::xt_coeffs() const {
    size_t it = get_nr_ieter();
    if (m_xt_cache_iter != it) {
        uint64_t nr;
        check(sanm_anm_nr_xt_coeffs(m_h, &nr));
        m_xt_cache.assign(nr, TensorND{});
        for (uint64_t i = 0; i < nr; ++i) {
            m_xt_cache[i].set_shape({m_nr_unknown + 1});
            check(sanm_anm_xt_coeff(m_h, i, m_xt_cache[i].woptr()));
        }
        m_xt_cache_iter = it;
    }
    return m_xt_cache;
}

ANMSolverVecScale::ANMSolverVecScale(DeferCreate, symbolic::VarNode* f,
                                     SparseLinearDescPtr remap_inp,
                                     SparseLinearDescPtr remap_out,
                                     const TensorShape& x_shape,
                                     const HyperParam& hyper_param)
        : ANMDriverHelper(f, std::move(remap_inp), std::move(remap_out),
                          x_shape, hyper_param) {}

ANMSolverVecScale::ANMSolverVecScale(symbolic::VarNode* f,
                                     SparseLinearDescPtr remap_inp,
                                     SparseLinearDescPtr remap_out, TensorND x0,
                                     fp_t t0, TensorND v,
                                     const HyperParam& hyper_param)
        : ANMDriverHelper(f, std::move(remap_inp), std::move(remap_out),
                          x0.shape(), hyper_param) {
    sanm_anm_hyper_t h = to_c(hyper_param);
    Dims xd{x0.shape()}, vd{v.shape()};
    check(sanm_anm_vecscale_create(graph_of(f), f->device_id(),
                                   m_remap_inp->device_remap(),
                                   m_remap_out->device_remap(), x0.ptr(),
                                   xd.rank, xd.d, t0, v.ptr(), vd.rank, vd.d,
                                   &h, &m_h));
}

ANMEqnSolver::ANMEqnSolver(symbolic::VarNode* f, SparseLinearDescPtr remap_inp,
                           SparseLinearDescPtr remap_out, TensorND x0,
                           TensorND y, const HyperParam& hyper_param)
        : ANMSolverVecScale(DeferCreate{}, f, std::move(remap_inp),
                            std::move(remap_out), x0.shape(), hyper_param) {
    sanm_anm_hyper_t h = to_c(hyper_param, hyper_param.converge_rms);
    Dims xd{x0.shape()}, yd{y.shape()};
    check(sanm_anm_eqn_create(graph_of(f), f->device_id(),
                              m_remap_inp->device_remap(),
                              m_remap_out->device_remap(), x0.ptr(), xd.rank,
                              xd.d, y.ptr(), yd.rank, yd.d, &h, &m_h));
}

fp_t ANMEqnSolver::residual_rms() const {
    double r;
    check(sanm_anm_eqn_residual_rms(m_h, &r));
    return r;
}

bool ANMEqnSolver::converged() const {
    int c;
    check(sanm_anm_eqn_converged(m_h, &c));
    return c != 0;
}

ANMEqnSolver& ANMEqnSolver::next_iter() {
    check(sanm_anm_eqn_next_iter(m_h));
    return *this;
}

TensorND ANMEqnSolver::get_x() const {
    TensorND x{m_x_shape};
    check(sanm_anm_eqn_get_x(m_h, x.woptr()));
    return x;
}

ANMImplicitSolver::ANMImplicitSolver(symbolic::VarNode* f,
                                     SparseLinearDescPtr remap_inp,
                                     SparseLinearDescPtr remap_out,
                                     const TensorND& x0, fp_t t0,
                                     const HyperParam& hyper_param)
        : ANMDriverHelper(f, std::move(remap_inp), std::move(remap_out),
                          x0.shape(), hyper_param) {
    sanm_anm_hyper_t h = to_c(hyper_param);
    check(sanm_anm_implicit_create(graph_of(f), f->device_id(),
                                   m_remap_inp->device_remap(),
                                   m_remap_out->device_remap(), x0.ptr(),
                                   x0.shape().total_nr_elems(), t0, &h, &m_h));
}

const TensorND& ANMImplicitSolver::fx0() const {
    size_t it = get_nr_ieter();
    if (m_fx0_iter != it) {
        m_fx0.set_shape(m_remap_out->out_shape());
        check(sanm_anm_implicit_fx0(m_h, m_fx0.woptr()));
        m_fx0_iter = it;
    }
    return m_fx0;
}

// -------------------------------------------------------------- SparseSolver
namespace {
int g_solver_threads = 0, g_solver_verbosity = 0;
}

class SparseSolver::Impl {
public:
    struct Builder final : SparseMatBuilder {
        size_t offset;
        std::vector<uint64_t> rows, cols;
        std::vector<double> vals;
        explicit Builder(size_t off) : offset(off) {}
        void add_constraint(size_t cidx, size_t xidx, fp_t coeff) override {
            sanm_assert(cidx >= offset &&
                                (rows.empty() || cidx >= rows.back()),
                        "rows must be added in increasing order starting at "
                        "the builder's offset");
            rows.push_back(cidx);
            cols.push_back(xidx);
            vals.push_back(coeff);
        }
    };
    size_t n;
    sanm_sparse* h = nullptr;
    bool prepared = false;
    std::vector<std::unique_ptr<Builder>> builders;

    explicit Impl(size_t n_) : n(n_) { check(sanm_sparse_create(n_, &h)); }
    ~Impl() {
        if (h)
            sanm_sparse_destroy(h);
    }
};

SparseSolver::SparseSolver(size_t nr_xs)
        : m_pimpl(std::make_unique<Impl>(nr_xs)) {}
SparseSolver::~SparseSolver() = default;

void SparseSolver::set_num_threads(int nr) {
    g_solver_threads = nr;
}
int SparseSolver::get_num_threads() {
    return g_solver_threads > 0 ? g_solver_threads : sanm::get_num_threads();
}
void SparseSolver::set_verbosity(int verbosity) {
    g_solver_verbosity = verbosity;
}

SparseSolver::SparseMatBuilder* SparseSolver::make_builder(size_t cidx_offset) {
    sanm_assert(!m_pimpl->prepared, "make_builder() after prepare()");
    m_pimpl->builders.push_back(std::make_unique<Impl::Builder>(cidx_offset));
    return m_pimpl->builders.back().get();
}

SparseSolver& SparseSolver::prepare(fp_t l2_penalty) {
    auto& im = *m_pimpl;
    sanm_assert(!im.prepared, "prepare() called twice");
    // builders cover increasing, disjoint row ranges
    // (libsanm/sparse_solver.h:39-51): feed them in offset order
    std::vector<Impl::Builder*> order;
    for (auto& b : im.builders)
        order.push_back(b.get());
    std::sort(order.begin(), order.end(),
              [](auto* a, auto* b) { return a->offset < b->offset; });
    for (auto* b : order)
        if (!b->rows.empty())
            check(sanm_sparse_add_constraints(im.h, b->rows.size(),
                                              b->rows.data(), b->cols.data(),
                                              b->vals.data()));
    im.builders.clear();
    check(sanm_sparse_prepare(im.h, l2_penalty));
    im.prepared = true;
    if (g_solver_verbosity) {
        uint64_t nnz_lu, nsn, nlv;
        double flops;
        if (sanm_sparse_stats(im.h, &nnz_lu, &flops, &nsn, &nlv) == SANM_OK)
            printf("sparse LU: n=%zu nnz(L+U)=%llu flops=%.3g supernodes=%llu "
                   "levels=%llu\n",
                   im.n, (unsigned long long)nnz_lu, flops,
                   (unsigned long long)nsn, (unsigned long long)nlv);
    }
    return *this;
}

TensorND SparseSolver::solve(const TensorND& b) const {
    sanm_assert(m_pimpl->prepared, "solve() before prepare()");
    sanm_assert(b.shape().total_nr_elems() == m_pimpl->n);
    TensorND x{b.shape()};
    check(sanm_sparse_solve(m_pimpl->h, b.ptr(), x.woptr()));
    return x;
}

TensorND SparseSolver::apply(const TensorND& x) const {
    sanm_assert(m_pimpl->prepared, "apply() before prepare()");
    sanm_assert(x.shape().total_nr_elems() == m_pimpl->n);
    TensorND y{x.shape()};
    check(sanm_sparse_apply(m_pimpl->h, x.ptr(), y.woptr()));
    return y;
}

fp_t SparseSolver::coeff_l2() const {
    double v;
    check(sanm_sparse_coeff_l2(m_pimpl->h, &v));
    return v;
}

void SparseSolver::dump(const TensorND& b, FILE* fout) const {
    // the matrix lives on the device in factored form; dump what can be shown
    fprintf(fout, "SparseSolver(n=%zu, prepared=%d) on device; |A|_F=%g\n",
            m_pimpl->n, int(m_pimpl->prepared),
            m_pimpl->prepared ? coeff_l2() : 0.);
    if (!b.empty()) {
        fprintf(fout, "b:");
        for (size_t i = 0, n = b.shape().total_nr_elems(); i < n; ++i)
            fprintf(fout, " %g", b.ptr()[i]);
        fprintf(fout, "\n");
    }
}

// --------------------------------------------------------- PadeApproximation
PadeApproximation::PadeApproximation(std::span<const TensorND> xs,
                                     bool anm_cond, bool sanity_check) {
    sanm_assert(!xs.empty());
    m_len = xs[0].shape().total_nr_elems();
    m_rows.resize(xs.size() * m_len);
    for (size_t i = 0; i < xs.size(); ++i) {
        sanm_assert(xs[i].shape().total_nr_elems() == m_len);
        std::memcpy(m_rows.data() + i * m_len, xs[i].ptr(),
                    m_len * sizeof(double));
    }
    check(sanm_pade_create(m_rows.data(), xs.size(), m_len, anm_cond,
                           sanity_check, &m_h));
    m_rows.clear();
    m_rows.shrink_to_fit();  // the handle keeps its own copy
}

PadeApproximation::~PadeApproximation() {
    if (m_h)
        sanm_pade_destroy(m_h);
}

bool PadeApproximation::estimate_valid_range(fp_t start, fp_t eps, fp_t limit) {
    int ok;
    check(sanm_pade_estimate_valid_range(m_h, start, eps, limit, &ok));
    return ok != 0;
}

fp_t PadeApproximation::get_t_max() const {
    double t, a;
    check(sanm_pade_get_t_max(m_h, &t, &a));
    return t;
}

fp_t PadeApproximation::get_t_max_a() const {
    double t, a;
    check(sanm_pade_get_t_max(m_h, &t, &a));
    return a;
}

fp_t PadeApproximation::solve_a(fp_t t) const {
    double a;
    check(sanm_pade_solve_a(m_h, t, &a));
    return a;
}

TensorND PadeApproximation::eval_xt(fp_t a) const {
    TensorND xt{TensorShape{m_len}};
    check(sanm_pade_eval_xt(m_h, a, xt.woptr()));
    return xt;
}

std::pair<TensorND, fp_t> PadeApproximation::eval(fp_t a) const {
    TensorND xt = eval_xt(a);
    fp_t t = xt.ptr()[m_len - 1];
    return {xt.take_sub({m_len - 1}), t};
}

// ---------------------------------------------------------- unary_polynomial
namespace unary_polynomial {

fp_t eval(coeff_t f, fp_t x) {
    fp_t acc = 0;
    for (size_t i = f.size(); i-- > 0;)
        acc = acc * x + f[i];
    return acc;
}

fp_t solve_quad(fp_t a, fp_t b, fp_t c) {
    // larger root of a x^2 + b x + c, or the minimiser when there is none
    // (libsanm/unary_polynomial.cpp:76-83)
    sanm_assert(a > 0, "bad a: %g", a);
    fp_t disc = b * b - 4 * a * c;
    if (disc < 0)
        return -b / (2 * a);
    return (std::sqrt(disc) - b) / (2 * a);
}

fp_t solve_eqn(coeff_t f, fp_t xmin, fp_t xmax, fp_t b, fp_t eps) {
    double out;
    check(sanm_poly_solve_eqn(f.data(), f.size(), xmin, xmax, b, eps, &out));
    return out;
}

fp_t stable_x_range(int order) {
    // x^order stays inside the ~15.9 decimal digits of a double
    // (libsanm/unary_polynomial.cpp:94-100)
    return std::pow(1e15, 1. / static_cast<double>(order));
}

Maybe<std::vector<std::complex<fp_t>>> roots(coeff_t f, bool only_real,
                                             int /*max_iter*/, fp_t /*tol*/) {
    std::vector<double> re(f.size()), im(f.size());
    uint64_t nr = 0;
    int converged = 0;
    check(sanm_poly_roots(f.data(), f.size(), only_real, re.data(), im.data(),
                          &nr, &converged));
    if (!converged)
        return None;
    std::vector<std::complex<fp_t>> out(nr);
    for (uint64_t i = 0; i < nr; ++i)
        out[i] = {re[i], im[i]};
    return out;
}

TensorND eval_tensor(std::span<const TensorND> f, fp_t x) {
    sanm_assert(!f.empty());
    TensorND acc = f.back();
    for (size_t i = f.size() - 1; i-- > 0;) {
        acc *= x;
        acc += f[i];
    }
    return acc;
}

}  // namespace unary_polynomial
}  // namespace sanm
