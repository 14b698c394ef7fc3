// cxx/sanm_host/graph.hpp -- symbolic graph + Taylor propagation classes of
// the libsanm-compatible host layer.  Interfaces mirrored:
//   libsanm/symbolic.h:108-130 (VarNode), :283-293 (ComputingGraph),
//   :298-317 (TensorValueMap), :337-383 (TaylorCoeffProp),
//   :391-423 (ParallelTaylorCoeffProp), :426 (eval_unary_func),
//   libsanm/oprs.h:14-105 (SymbolVar and the free builder functions).
//
// A ComputingGraph owns a `sanm_graph_t` of the C ABI; every builder call
// records its operator there and returns a VarNode that carries the integer
// variable id.  Propagation objects are thin handles on `sanm_prop_t`: the
// Taylor stacks live in HBM, host TensorNDs are filled on request only.
// The operator set is the closed set of include/sanm_b200.h (the 13 operators
// of libsanm/oprs/*): there is no host-side OperatorMeta plug-in vtable on the
// GPU path -- a user operator would have to be a device function.
#pragma once

#include "sanm_host/tensor.hpp"

#include <array>
#include <deque>
#include <unordered_map>

struct sanm_graph;
struct sanm_prop;

namespace sanm {
namespace symbolic {

class ComputingGraph;
class VarNode;
using VarNodeArray = std::vector<VarNode*>;

//! a variable; its address is its identity (libsanm/symbolic.h:108-112)
class VarNode : public NonCopyable {
public:
    VarNode(ComputingGraph* g, int id, VarNodeArray deps, const char* opr,
            bool is_placeholder)
            : m_graph(g),
              m_id(id),
              m_deps(std::move(deps)),
              m_opr(opr),
              m_placeholder(is_placeholder) {}

    ComputingGraph* owner_graph() const { return m_graph; }
    //! variable id inside the owning sanm_graph_t
    int device_id() const { return m_id; }
    //! variables this one is computed from
    const VarNodeArray& deps() const { return m_deps; }
    const char* opr_name() const { return m_opr; }
    bool is_placeholder() const { return m_placeholder; }

private:
    ComputingGraph* m_graph;
    int m_id;
    VarNodeArray m_deps;
    const char* m_opr;
    bool m_placeholder;
};

class ComputingGraph : public NonCopyable {
public:
    ComputingGraph() = default;
    ~ComputingGraph();

    //! the C-ABI graph (created on first use)
    sanm_graph* handle();
    //! register the output variable of an operator just recorded
    VarNode* add_var(int id, VarNodeArray deps, const char* opr,
                     bool is_placeholder = false);
    size_t next_obj_id() { return m_next_id++; }

private:
    sanm_graph* m_handle = nullptr;
    std::deque<VarNode> m_vars;  // stable addresses
    size_t m_next_id = 0;
};

//! placeholders reachable from \p dest, in discovery order of a DFS that
//! follows operands left to right
VarNodeArray find_placeholders(VarNode* dest);

class TensorValueMap {
public:
    TensorValueMap& insert(VarNode* var, TensorND value);
    const TensorND& get(VarNode* var) const { return m_map.at(var); }
    TensorND& get(VarNode* var) { return m_map.at(var); }
    TensorArray pack(std::span<VarNode* const> vars) const;
    auto begin() { return m_map.begin(); }
    auto end() { return m_map.end(); }
    auto begin() const { return m_map.begin(); }
    auto end() const { return m_map.end(); }

private:
    std::unordered_map<VarNode*, TensorND> m_map;
};

//! sugar over VarNode* (libsanm/oprs.h:14-67)
class SymbolVar {
public:
    SymbolVar() = default;
    SymbolVar(VarNode* var) : m_var(var) {}
    VarNode* node() const { return m_var; }

    SymbolVar operator+(SymbolVar rhs) const;
    inline SymbolVar operator+(fp_t rhs) const;
    SymbolVar operator-(SymbolVar rhs) const;
    SymbolVar operator-(fp_t rhs) const { return *this + (-rhs); }
    SymbolVar operator*(SymbolVar rhs) const;
    inline SymbolVar operator*(fp_t rhs) const;

    SymbolVar reduce_sum(int axis, bool keepdim = true) const;
    SymbolVar batched_transpose() const;
    inline SymbolVar batched_matinv() const;
    SymbolVar batched_matmul(SymbolVar rhs) const;
    SymbolVar batched_det() const;
    SymbolVar batched_mul_eye(size_t dim) const;
    SymbolVar pow(fp_t exp) const;
    SymbolVar log() const;
    std::array<SymbolVar, 3> batched_svd_w(bool require_rotation = false) const;
    SymbolVar slice(int axis, Maybe<int> begin, Maybe<int> end, int stride = 1);

private:
    VarNode* m_var = nullptr;
};

SymbolVar batched_mat_inv_mul(SymbolVar x, SymbolVar a, bool is_left);
SymbolVar linear_combine(const std::vector<std::pair<fp_t, SymbolVar>>& vars,
                         fp_t bias = 0);
SymbolVar placeholder(ComputingGraph& cg);
SymbolVar constant(ComputingGraph& cg, TensorND val);
SymbolVar concat(std::span<const SymbolVar> xs, int axis);

SymbolVar SymbolVar::batched_matinv() const {
    return batched_mat_inv_mul(*this, {}, true);
}
SymbolVar SymbolVar::operator+(fp_t rhs) const {
    return linear_combine({{1._fp, *this}}, rhs);
}
SymbolVar SymbolVar::operator*(fp_t rhs) const {
    return linear_combine({{rhs, *this}}, 0);
}
static inline SymbolVar operator-(fp_t x, SymbolVar y) {
    return linear_combine({{-1._fp, y}}, x);
}

//! propagation state shared by TaylorCoeffProp and ParallelTaylorCoeffProp
class DeviceProp : public NonCopyable {
public:
    DeviceProp(VarNode* output, int max_order);
    ~DeviceProp();
    void push(const TensorND& xi);
    const TensorND& next_bias();
    const TensorND& prev_bias() const;
    TensorND yi() const;
    StSparseLinearTrans jacobian();
    TensorShape out_shape() const;

private:
    sanm_prop* m_h = nullptr;
    TensorND m_bias;
    bool m_expect_push = true, m_have_bias = false;
};

//! single-shard propagator (libsanm/symbolic.h:337-383).  On the device path
//! it takes one placeholder input and a batched output (what the ANM drivers
//! and the reference's Taylor tests use).
class TaylorCoeffProp : public NonCopyable {
public:
    TaylorCoeffProp(VarNode* output, bool output_is_batched);
    ~TaylorCoeffProp();

    const VarNodeArray& inputs() const { return m_inputs; }
    const TensorND& push_xi(std::span<const TensorND> xi);
    const TensorND& push_xi(const TensorValueMap& vmap) {
        return push_xi(vmap.pack(inputs()));
    }
    const TensorND& push_xi(const TensorND& xi) {
        return push_xi(std::span<const TensorND>{&xi, 1});
    }
    const TensorND& compute_next_order_bias();
    const TensorND& get_prev_next_order_bias() const;
    const StSparseLinearTrans& get_jacobian(VarNode* x);

private:
    VarNodeArray m_inputs;
    std::unique_ptr<DeviceProp> m_prop;
    TensorND m_yi;
    StSparseLinearTrans m_jac;
};

//! batch-parallel propagator (libsanm/symbolic.h:391-423).  The reference
//! splits the batch over worker threads; here the batch is spread over the
//! thread blocks of one GPU (and over ranks when sanm_comm_init_* was called),
//! so there is exactly one host "worker".
class ParallelTaylorCoeffProp : public NonCopyable {
    class Impl;
    std::unique_ptr<Impl> m_pimpl;

public:
    explicit ParallelTaylorCoeffProp(VarNode* var);
    ~ParallelTaylorCoeffProp();
    ParallelTaylorCoeffProp& push_xi(const TensorND& val);
    const TensorND& compute_next_order_bias();
    const TensorND& get_prev_next_order_bias() const;
    StSparseLinearTrans get_jacobian();
    TensorND gather_yi() const;
    size_t nr_worker() const;
    using WorkerTaskFn = std::function<void(size_t worker_id)>;
    void run_on_workers(const WorkerTaskFn& fn);
};

TensorND eval_unary_func(VarNode* y, const TensorND& xval);

//! upper bound on the orders a directly constructed propagator will be asked
//! for (device stacks are preallocated); env SANM_PROP_MAX_ORDER, default 32
int default_prop_max_order();

}  // namespace symbolic

using symbolic::TensorValueMap;
using SymbolVar = symbolic::SymbolVar;
}  // namespace sanm
