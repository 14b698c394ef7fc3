// cxx/sanm_host/graph.cpp -- the symbolic layer recorded into the C ABI.
// Every builder below forwards to one `sanm_graph_*` entry of
// include/sanm_b200.h (which cites the reference operator it replaces); the
// propagation classes forward to `sanm_prop_*`.
#include "sanm_host/graph.hpp"

#include "sanm_b200.h"

#include <cstdlib>
#include <unordered_set>

namespace sanm {

void capi::check(int status) {
    if (status == SANM_OK)
        return;
    std::string msg = sanm_last_error();
    switch (status) {
        case SANM_ERR_ASSERT:
            throw SANMAssertionError{std::move(msg)};
        case SANM_ERR_NUMERICAL:
            throw SANMNumericalError{std::move(msg)};
        default:  // incl. SANM_ERR_CUDA: no device means no result
            throw SANMError{std::move(msg)};
    }
}

namespace symbolic {
using capi::check;
using capi::Dims;

// ------------------------------------------------------------ ComputingGraph
ComputingGraph::~ComputingGraph() {
    if (m_handle)
        sanm_graph_destroy(m_handle);
}

sanm_graph* ComputingGraph::handle() {
    if (!m_handle)
        check(sanm_graph_create(&m_handle));
    return m_handle;
}

VarNode* ComputingGraph::add_var(int id, VarNodeArray deps, const char* opr,
                                 bool is_placeholder) {
    m_vars.emplace_back(this, id, std::move(deps), opr, is_placeholder);
    return &m_vars.back();
}

VarNodeArray find_placeholders(VarNode* dest) {
    VarNodeArray found, stack{dest};
    std::unordered_set<VarNode*> seen{dest};
    // iterative DFS; operands are pushed right to left so that the leftmost
    // operand is expanded first
    while (!stack.empty()) {
        VarNode* v = stack.back();
        stack.pop_back();
        if (v->is_placeholder())
            found.push_back(v);
        auto& d = v->deps();
        for (size_t i = d.size(); i-- > 0;)
            if (seen.insert(d[i]).second)
                stack.push_back(d[i]);
    }
    return found;
}

// ------------------------------------------------------------ TensorValueMap
TensorValueMap& TensorValueMap::insert(VarNode* var, TensorND value) {
    auto r = m_map.emplace(var, std::move(value));
    sanm_assert(r.second, "duplicated var in TensorValueMap");
    return *this;
}

TensorArray TensorValueMap::pack(std::span<VarNode* const> vars) const {
    sanm_assert(vars.size() == m_map.size(), "expect %zu values, got %zu",
                vars.size(), m_map.size());
    TensorArray out;
    out.reserve(vars.size());
    for (VarNode* v : vars)
        out.push_back(m_map.at(v));
    return out;
}

// ----------------------------------------------------------------- builders
namespace {
ComputingGraph* graph_of(VarNode* v) {
    sanm_assert(v, "null SymbolVar used as an operand");
    return v->owner_graph();
}

ComputingGraph* common_graph(std::initializer_list<VarNode*> vs) {
    ComputingGraph* g = nullptr;
    for (VarNode* v : vs) {
        if (!v)
            continue;
        sanm_assert(!g || g == v->owner_graph(),
                    "operands belong to different computing graphs");
        g = v->owner_graph();
    }
    sanm_assert(g);
    return g;
}
}  // namespace

SymbolVar SymbolVar::operator+(SymbolVar rhs) const {
    return linear_combine({{1._fp, *this}, {1._fp, rhs}});
}

SymbolVar SymbolVar::operator-(SymbolVar rhs) const {
    return linear_combine({{1._fp, *this}, {-1._fp, rhs}});
}

SymbolVar SymbolVar::operator*(SymbolVar rhs) const {
    auto g = common_graph({m_var, rhs.m_var});
    int out;
    check(sanm_graph_multiply(g->handle(), m_var->device_id(),
                              rhs.m_var->device_id(), &out));
    return g->add_var(out, {m_var, rhs.m_var}, "multiply");
}

SymbolVar SymbolVar::pow(fp_t exp) const {
    if (exp == 1.)
        return *this;  // libsanm/oprs.cpp:28-30
    auto g = graph_of(m_var);
    int out;
    check(sanm_graph_pow(g->handle(), m_var->device_id(), exp, &out));
    return g->add_var(out, {m_var}, "pow");
}

SymbolVar SymbolVar::log() const {
    auto g = graph_of(m_var);
    int out;
    check(sanm_graph_log(g->handle(), m_var->device_id(), &out));
    return g->add_var(out, {m_var}, "log");
}

SymbolVar SymbolVar::reduce_sum(int axis, bool keepdim) const {
    auto g = graph_of(m_var);
    int out;
    check(sanm_graph_reduce_sum(g->handle(), m_var->device_id(), axis, keepdim,
                                &out));
    return g->add_var(out, {m_var}, "reduce_sum");
}

SymbolVar SymbolVar::batched_transpose() const {
    auto g = graph_of(m_var);
    int out;
    check(sanm_graph_batched_transpose(g->handle(), m_var->device_id(), &out));
    return g->add_var(out, {m_var}, "batched_transpose");
}

SymbolVar SymbolVar::batched_matmul(SymbolVar rhs) const {
    auto g = common_graph({m_var, rhs.m_var});
    int out;
    check(sanm_graph_batched_matmul(g->handle(), m_var->device_id(),
                                    rhs.m_var->device_id(), &out));
    return g->add_var(out, {m_var, rhs.m_var}, "batched_matmul");
}

SymbolVar SymbolVar::batched_det() const {
    auto g = graph_of(m_var);
    int out;
    check(sanm_graph_batched_det(g->handle(), m_var->device_id(), &out));
    return g->add_var(out, {m_var}, "batched_det");
}

SymbolVar SymbolVar::batched_mul_eye(size_t dim) const {
    auto g = graph_of(m_var);
    int out;
    check(sanm_graph_batched_mul_eye(g->handle(), m_var->device_id(), dim,
                                     &out));
    return g->add_var(out, {m_var}, "batched_mul_eye");
}

std::array<SymbolVar, 3> SymbolVar::batched_svd_w(bool require_rotation) const {
    auto g = graph_of(m_var);
    int out[3];
    check(sanm_graph_batched_svd_w(g->handle(), m_var->device_id(),
                                   require_rotation, out));
    return {g->add_var(out[0], {m_var}, "svd_w.U"),
            g->add_var(out[1], {m_var}, "svd_w.S"),
            g->add_var(out[2], {m_var}, "svd_w.W")};
}

SymbolVar SymbolVar::slice(int axis, Maybe<int> begin, Maybe<int> end,
                           int stride) {
    auto g = graph_of(m_var);
    int out;
    check(sanm_graph_slice(g->handle(), m_var->device_id(), axis, begin.valid(),
                           begin.valid() ? begin.val() : 0, end.valid(),
                           end.valid() ? end.val() : 0, stride, &out));
    return g->add_var(out, {m_var}, "slice");
}

SymbolVar batched_mat_inv_mul(SymbolVar x, SymbolVar a, bool is_left) {
    auto g = common_graph({x.node(), a.node()});
    int out;
    check(sanm_graph_batched_mat_inv_mul(
            g->handle(), x.node()->device_id(),
            a.node() ? a.node()->device_id() : -1, is_left, &out));
    VarNodeArray deps{x.node()};
    if (a.node())
        deps.push_back(a.node());
    return g->add_var(out, std::move(deps), "batched_mat_inv_mul");
}

SymbolVar linear_combine(const std::vector<std::pair<fp_t, SymbolVar>>& vars,
                         fp_t bias) {
    sanm_assert(!vars.empty());
    ComputingGraph* g = graph_of(vars[0].second.node());
    std::vector<fp_t> coeffs;
    std::vector<int> ids;
    VarNodeArray deps;
    for (auto& [c, v] : vars) {
        sanm_assert(graph_of(v.node()) == g,
                    "operands belong to different computing graphs");
        coeffs.push_back(c);
        ids.push_back(v.node()->device_id());
        deps.push_back(v.node());
    }
    int out;
    check(sanm_graph_linear_combine(g->handle(), static_cast<int>(ids.size()),
                                    coeffs.data(), ids.data(), bias, &out));
    return g->add_var(out, std::move(deps), "linear_combine");
}

SymbolVar placeholder(ComputingGraph& cg) {
    int out;
    check(sanm_graph_placeholder(cg.handle(), &out));
    return cg.add_var(out, {}, "placeholder", true);
}

SymbolVar constant(ComputingGraph& cg, TensorND val) {
    sanm_assert(!val.empty(), "constant needs a value");
    Dims d{val.shape()};
    int out;
    check(sanm_graph_constant(cg.handle(), val.ptr(), d.rank, d.d, &out));
    return cg.add_var(out, {}, "constant");
}

SymbolVar concat(std::span<const SymbolVar> xs, int axis) {
    sanm_assert(!xs.empty());
    ComputingGraph* g = graph_of(xs[0].node());
    std::vector<int> ids;
    VarNodeArray deps;
    for (auto& v : xs) {
        sanm_assert(graph_of(v.node()) == g);
        ids.push_back(v.node()->device_id());
        deps.push_back(v.node());
    }
    int out;
    check(sanm_graph_concat(g->handle(), static_cast<int>(ids.size()),
                            ids.data(), axis, &out));
    return g->add_var(out, std::move(deps), "concat");
}

// --------------------------------------------------------------- DeviceProp
int default_prop_max_order() {
    if (const char* e = std::getenv("SANM_PROP_MAX_ORDER")) {
        int v = std::atoi(e);
        if (v > 0)
            return v;
    }
    return 32;
}

DeviceProp::DeviceProp(VarNode* output, int max_order) {
    sanm_assert(output);
    check(sanm_prop_create(output->owner_graph()->handle(),
                           output->device_id(), max_order, &m_h));
}

DeviceProp::~DeviceProp() {
    if (m_h)
        sanm_prop_destroy(m_h);
}

TensorShape DeviceProp::out_shape() const {
    int rank;
    uint64_t dims[3];
    check(sanm_prop_out_shape(m_h, &rank, dims));
    return capi::shape_from(rank, dims);
}

void DeviceProp::push(const TensorND& xi) {
    // the calls must alternate (libsanm/symbolic.cpp:163,253-255)
    sanm_assert(m_expect_push,
                "push_xi() must be followed by compute_next_order_bias()");
    Dims d{xi.shape()};
    check(sanm_prop_push_xi(m_h, xi.ptr(), d.rank, d.d));
    m_expect_push = false;
}

const TensorND& DeviceProp::next_bias() {
    sanm_assert(!m_expect_push,
                "compute_next_order_bias() must be preceded by push_xi()");
    m_bias.set_shape(out_shape());
    check(sanm_prop_compute_next_order_bias(m_h, m_bias.woptr()));
    m_expect_push = true;
    m_have_bias = true;
    return m_bias;
}

const TensorND& DeviceProp::prev_bias() const {
    sanm_assert(m_have_bias);
    return m_bias;
}

TensorND DeviceProp::yi() const {
    TensorND y{out_shape()};
    check(sanm_prop_gather_yi(m_h, y.woptr()));
    return y;
}

StSparseLinearTrans DeviceProp::jacobian() {
    int type;
    uint64_t batch, odim, idim;
    check(sanm_prop_get_jacobian(m_h, &type, &batch, &odim, &idim, nullptr));
    const bool full = type == 2;
    TensorND c{full ? TensorShape{batch, odim, idim} : TensorShape{batch, odim}};
    check(sanm_prop_get_jacobian(m_h, &type, &batch, &odim, &idim, c.woptr()));
    return {full ? StSparseLinearTrans::FULL : StSparseLinearTrans::ELEMWISE,
            true, c};
}

// ---------------------------------------------------------- TaylorCoeffProp
TaylorCoeffProp::TaylorCoeffProp(VarNode* output, bool output_is_batched)
        : m_inputs(find_placeholders(output)) {
    if (!output_is_batched || m_inputs.size() != 1)
        throw SANMError{
                "TaylorCoeffProp on the device path needs one placeholder "
                "input and a batched output"};
    m_prop = std::make_unique<DeviceProp>(output, default_prop_max_order());
}

TaylorCoeffProp::~TaylorCoeffProp() = default;

const TensorND& TaylorCoeffProp::push_xi(std::span<const TensorND> xi) {
    sanm_assert(xi.size() == 1, "expect one input value, got %zu", xi.size());
    m_prop->push(xi[0]);
    m_yi = m_prop->yi();
    return m_yi;
}

const TensorND& TaylorCoeffProp::compute_next_order_bias() {
    return m_prop->next_bias();
}

const TensorND& TaylorCoeffProp::get_prev_next_order_bias() const {
    return m_prop->prev_bias();
}

const StSparseLinearTrans& TaylorCoeffProp::get_jacobian(VarNode* x) {
    sanm_assert(x == m_inputs[0], "jacobian requested for a non-input var");
    if (!m_jac.valid())
        m_jac = m_prop->jacobian();
    return m_jac;
}

// -------------------------------------------------- ParallelTaylorCoeffProp
class ParallelTaylorCoeffProp::Impl {
public:
    DeviceProp prop;
    explicit Impl(VarNode* y) : prop(y, default_prop_max_order()) {}
};

ParallelTaylorCoeffProp::ParallelTaylorCoeffProp(VarNode* var) {
    auto inputs = find_placeholders(var);
    sanm_assert(inputs.size() == 1,
                "ParallelTaylorCoeffProp needs exactly one input, got %zu",
                inputs.size());
    m_pimpl = std::make_unique<Impl>(var);
}

ParallelTaylorCoeffProp::~ParallelTaylorCoeffProp() = default;

ParallelTaylorCoeffProp& ParallelTaylorCoeffProp::push_xi(const TensorND& val) {
    m_pimpl->prop.push(val);
    return *this;
}

const TensorND& ParallelTaylorCoeffProp::compute_next_order_bias() {
    return m_pimpl->prop.next_bias();
}

const TensorND& ParallelTaylorCoeffProp::get_prev_next_order_bias() const {
    return m_pimpl->prop.prev_bias();
}

StSparseLinearTrans ParallelTaylorCoeffProp::get_jacobian() {
    return m_pimpl->prop.jacobian();
}

TensorND ParallelTaylorCoeffProp::gather_yi() const {
    return m_pimpl->prop.yi();
}

size_t ParallelTaylorCoeffProp::nr_worker() const {
    return 1;
}

void ParallelTaylorCoeffProp::run_on_workers(const WorkerTaskFn& fn) {
    fn(0);
}

TensorND eval_unary_func(VarNode* y, const TensorND& xval) {
    sanm_assert(y);
    Dims d{xval.shape()};
    int orank;
    uint64_t odims[3];
    sanm_graph* g = y->owner_graph()->handle();
    check(sanm_eval_unary_func(g, y->device_id(), xval.ptr(), d.rank, d.d,
                               &orank, odims, nullptr));
    TensorND out{capi::shape_from(orank, odims)};
    check(sanm_eval_unary_func(g, y->device_id(), xval.ptr(), d.rank, d.d,
                               &orank, odims, out.woptr()));
    return out;
}

}  // namespace symbolic
}  // namespace sanm
