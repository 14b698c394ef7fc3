/* sanm_b200/csrc/graph.cpp -- graph builders and lowering (see graph.h). */
#include "graph.h"

#include <cmath>
#include <cstdarg>
#include <functional>

namespace sanm_b200 {

std::string ssprintf(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    char buf[1024];
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    return buf;
}

std::string Shape::str() const {
    std::string s = "{";
    for (int i = 0; i < rank; ++i) {
        if (i) s += ",";
        s += std::to_string(dim[i]);
    }
    return s + "}";
}

using namespace taylor;

int Graph::add_op(GraphOp op, int nout) {
    for (int v : op.in) check_var(v);
    int opi = (int)ops.size();
    int first = (int)var_owner.size();
    for (int i = 0; i < nout; ++i) {
        op.out.push_back(first + i);
        var_owner.push_back(opi);
    }
    ops.emplace_back(std::move(op));
    return first;
}

int Graph::placeholder() {
    GraphOp op;
    op.type = OP_PLACEHOLDER;
    return add_op(std::move(op), 1);
}

int Graph::constant(const double* data, const Shape& shape) {
    sanm_check(shape.rank >= 1 && shape.total() > 0, "empty constant");
    GraphOp op;
    op.type = OP_CONSTANT;
    op.cshape = shape;
    op.cdata.assign(data, data + shape.total());
    return add_op(std::move(op), 1);
}

int Graph::linear_combine(int n, const double* coeffs, const int* vars,
                          double bias) {
    // libsanm/oprs/elem_arith.cpp:108-119
    sanm_check(n > 0, "linear_combine needs inputs");
    sanm_check(n <= MAX_IN, "linear_combine supports at most %d inputs", MAX_IN);
    GraphOp op;
    op.type = OP_LINCOMB;
    op.in.assign(vars, vars + n);
    op.f.assign(coeffs, coeffs + n);
    op.f.push_back(bias);
    return add_op(std::move(op), 1);
}

int Graph::multiply(int a, int b) {
    GraphOp op;
    op.type = OP_MULTIPLY;
    op.in = {a, b};
    return add_op(std::move(op), 1);
}

int Graph::pow(int x, double e) {
    // libsanm/oprs.cpp:28-34: pow(1) is the identity;
    // libsanm/analytic_unary.cpp:165-168: zero exponent rejected
    check_var(x);
    if (e == 1.) return x;
    sanm_check(std::fabs(e) > 1e-9, "zero power not handled");
    GraphOp op;
    op.type = OP_POW;
    op.in = {x};
    op.f = {e};
    return add_op(std::move(op), 1);
}

int Graph::log(int x) {
    GraphOp op;
    op.type = OP_LOG;
    op.in = {x};
    return add_op(std::move(op), 1);
}

int Graph::reduce_sum(int x, int axis, bool keepdim) {
    GraphOp op;
    op.type = OP_REDUCE_SUM;
    op.in = {x};
    op.ip[0] = axis;
    op.ip[1] = keepdim;
    return add_op(std::move(op), 1);
}

int Graph::batched_matmul(int a, int b) {
    GraphOp op;
    op.type = OP_MATMUL;
    op.in = {a, b};
    return add_op(std::move(op), 1);
}

int Graph::batched_transpose(int x) {
    GraphOp op;
    op.type = OP_TRANSPOSE;
    op.in = {x};
    return add_op(std::move(op), 1);
}

int Graph::batched_det(int x) {
    GraphOp op;
    op.type = OP_DET;
    op.in = {x};
    return add_op(std::move(op), 1);
}

int Graph::batched_mul_eye(int x, uint64_t dim) {
    sanm_check(dim > 0 && dim <= (uint64_t)MAXD, "mul_eye dim=%zu", (size_t)dim);
    GraphOp op;
    op.type = OP_MULEYE;
    op.in = {x};
    op.ip[0] = (int)dim;
    return add_op(std::move(op), 1);
}

int Graph::batched_mat_inv_mul(int x, int a, bool is_left) {
    GraphOp op;
    op.type = OP_MATINVMUL;
    op.in = {x};
    if (a >= 0) op.in.push_back(a);
    op.ip[0] = is_left;
    op.ip[1] = a < 0;
    return add_op(std::move(op), 1);
}

void Graph::batched_svd_w(int x, bool require_rotation, int out[3]) {
    GraphOp op;
    op.type = OP_SVDW;
    op.in = {x};
    op.ip[0] = require_rotation;
    int first = add_op(std::move(op), 3);
    for (int i = 0; i < 3; ++i) out[i] = first + i;
}

int Graph::slice(int x, int axis, bool has_begin, int begin, bool has_end,
                 int end, int stride) {
    sanm_check(stride != 0, "slice stride must be non-zero");
    GraphOp op;
    op.type = OP_SLICE;
    op.in = {x};
    op.ip[0] = axis;
    op.ip[1] = has_begin;
    op.ip[2] = begin;
    op.ip[3] = has_end;
    op.ip[4] = end;
    op.ip[5] = stride;
    return add_op(std::move(op), 1);
}

int Graph::concat(int n, const int* vars, int axis) {
    sanm_check(n > 0 && n <= MAX_IN, "concat supports 1..%d inputs", MAX_IN);
    GraphOp op;
    op.type = OP_CONCAT;
    op.in.assign(vars, vars + n);
    op.ip[0] = axis;
    return add_op(std::move(op), 1);
}

/* ------------------------------------------------------------ lowering */

namespace {

Shape infer_elemwise(const std::vector<Shape>& ins) {
    // libsanm/oprs/elem_arith.cpp:12-41
    Shape o;
    for (const Shape& s : ins) {
        if (!o.rank) {
            o = s;
        } else if (o != s) {
            sanm_check(o.is_batched_scalar() || s.is_batched_scalar(),
                       "invalid shape in elem arith: %s vs %s", o.str().c_str(),
                       s.str().c_str());
            if (o.is_batched_scalar() && !s.is_batched_scalar()) o = s;
            sanm_check(o.dim[0] == s.dim[0],
                       "batch mismatch in elem arith: %s vs %s",
                       o.str().c_str(), s.str().c_str());
        }
    }
    return o;
}

void set_dims(LoweredVar& v) {
    v.e = (int)v.shape.per_item();
    v.d1 = v.shape.rank >= 2 ? (int)v.shape.dim[1] : 1;
    v.d2 = v.shape.rank >= 3 ? (int)v.shape.dim[2] : 1;
}

}  // namespace

Lowered lower_graph(const Graph& g, int out_var, const Shape& x_shape) {
    g.check_var(out_var);
    sanm_check(x_shape.rank >= 1 && x_shape.total() > 0, "empty input");
    Lowered L;
    L.batch = x_shape.dim[0];
    L.gvar2lvar.assign(g.var_owner.size(), -1);

    // reachable ops; builder order is already a topological order
    std::vector<char> need(g.ops.size(), 0);
    {
        std::vector<int> stack{g.var_owner[out_var]};
        while (!stack.empty()) {
            int o = stack.back();
            stack.pop_back();
            if (need[o]) continue;
            need[o] = 1;
            for (int v : g.ops[o].in) stack.push_back(g.var_owner[v]);
        }
    }

    int nr_placeholder = 0;
    for (size_t oi = 0; oi < g.ops.size(); ++oi) {
        if (!need[oi]) continue;
        const GraphOp& go = g.ops[oi];
        LoweredOp lo;
        lo.type = go.type;
        lo.gop = (int)oi;
        for (int v : go.in) {
            sanm_check(L.gvar2lvar[v] >= 0);
            lo.in.push_back(L.gvar2lvar[v]);
        }
        for (int v : go.out) {
            L.gvar2lvar[v] = (int)L.vars.size();
            lo.out.push_back((int)L.vars.size());
            L.vars.emplace_back();
        }
        auto in_shape = [&](int i) -> const Shape& {
            return L.vars[lo.in[i]].shape;
        };
        auto& o0 = L.vars[lo.out[0]];
        bool all_const = !lo.in.empty();
        for (int v : lo.in) all_const = all_const && L.vars[v].is_const;

        switch (go.type) {
            case OP_PLACEHOLDER:
                sanm_check(++nr_placeholder == 1,
                           "multiple input variables are not supported");
                o0.shape = x_shape;
                L.input = lo.out[0];
                all_const = false;
                break;
            case OP_CONSTANT:
                sanm_check(go.cshape.dim[0] == L.batch,
                           "constant batch %s does not match input batch %zu",
                           go.cshape.str().c_str(), (size_t)L.batch);
                o0.shape = go.cshape;
                all_const = true;
                break;
            case OP_LINCOMB:
            case OP_MULTIPLY: {
                std::vector<Shape> ins;
                for (size_t i = 0; i < lo.in.size(); ++i) ins.push_back(in_shape(i));
                o0.shape = infer_elemwise(ins);
                break;
            }
            case OP_POW:
            case OP_LOG:
                o0.shape = in_shape(0);
                break;
            case OP_REDUCE_SUM: {
                const Shape& s = in_shape(0);
                int axis = go.ip[0];
                bool keepdim = go.ip[1];
                if (axis == -1) {
                    o0.shape = Shape{s.dim[0], 1};
                } else {
                    sanm_check(axis >= 1 && axis < s.rank,
                               "reduce axis %d unsupported for shape %s (only "
                               "per-item axes)",
                               axis, s.str().c_str());
                    o0.shape = s;
                    if (keepdim) {
                        o0.shape.dim[axis] = 1;
                    } else {
                        o0.shape.rank -= 1;
                        for (int i = axis; i < o0.shape.rank; ++i) {
                            o0.shape.dim[i] = s.dim[i + 1];
                        }
                    }
                }
                break;
            }
            case OP_MATMUL: {
                const Shape &a = in_shape(0), &b = in_shape(1);
                sanm_check(a.rank == 3 && b.rank == 3 && a.dim[0] == b.dim[0] &&
                                   a.dim[2] == b.dim[1],
                           "invalid operand shapes for matmul: %s and %s",
                           a.str().c_str(), b.str().c_str());
                sanm_check(a.dim[1] <= (uint64_t)MAXD && a.dim[2] <= (uint64_t)MAXD &&
                                   b.dim[2] <= (uint64_t)MAXD,
                           "matmul dims above %d unsupported", MAXD);
                o0.shape = Shape{a.dim[0], a.dim[1], b.dim[2]};
                break;
            }
            case OP_TRANSPOSE: {
                const Shape& s = in_shape(0);
                sanm_check(s.rank == 3, "transpose needs rank 3");
                o0.shape = Shape{s.dim[0], s.dim[2], s.dim[1]};
                break;
            }
            case OP_DET: {
                const Shape& s = in_shape(0);
                sanm_check(s.rank == 3 && s.dim[1] == s.dim[2],
                           "invalid shape for det: %s", s.str().c_str());
                sanm_check(s.dim[1] > 1, "scalar determinant is unsupported");
                sanm_check(s.dim[1] <= (uint64_t)MAXD,
                           "determinant of %zux%zu matrices: dims above %d are "
                           "not supported on the device path",
                           (size_t)s.dim[1], (size_t)s.dim[1], MAXD);
                o0.shape = Shape{s.dim[0], 1};
                break;
            }
            case OP_MULEYE: {
                const Shape& s = in_shape(0);
                sanm_check(s.is_batched_scalar(),
                           "the input shape must be a scalar: got %s",
                           s.str().c_str());
                o0.shape = Shape{s.dim[0], (uint64_t)go.ip[0], (uint64_t)go.ip[0]};
                break;
            }
            case OP_MATINVMUL: {
                const Shape& s = in_shape(0);
                sanm_check(s.rank == 3 && s.dim[1] == s.dim[2] &&
                                   s.dim[1] <= (uint64_t)MAXD,
                           "invalid shape for matinv: %s", s.str().c_str());
                if (lo.in.size() > 1) {
                    sanm_check(s == in_shape(1), "matinvmul shape mismatch");
                }
                o0.shape = s;
                break;
            }
            case OP_SVDW: {
                const Shape& s = in_shape(0);
                sanm_check(s.rank == 3 && s.dim[1] == s.dim[2],
                           "invalid shape for SVD-W: %s", s.str().c_str());
                sanm_check(s.dim[1] >= 2 && s.dim[1] <= (uint64_t)MAXD,
                           "SVD-W is implemented for 2x2 .. %dx%d matrices", MAXD,
                           MAXD);
                L.vars[lo.out[0]].shape = s;
                L.vars[lo.out[1]].shape = Shape{s.dim[0], s.dim[1]};
                L.vars[lo.out[2]].shape = s;
                break;
            }
            case OP_SLICE: {
                Shape s = in_shape(0);
                int axis = go.ip[0], stride = go.ip[5];
                sanm_check(axis >= 1 && axis < s.rank,
                           "slice axis %d unsupported (only per-item axes)", axis);
                int size = (int)s.dim[axis], begin, end;
                if (go.ip[1]) {
                    begin = go.ip[2];
                    if (begin < 0) begin += size;
                } else {
                    begin = stride > 0 ? 0 : size - 1;
                }
                if (go.ip[3]) {
                    end = go.ip[4];
                    if (end < 0) end += size;
                } else {
                    end = stride > 0 ? size : -1;
                }
                if (stride < 0) {
                    sanm_check(begin > end && end >= -1 && begin < size);
                } else {
                    sanm_check(begin < end && begin >= 0 && end <= size);
                }
                s.dim[axis] = (std::abs(end - begin) - 1) / std::abs(stride) + 1;
                o0.shape = s;
                break;
            }
            case OP_CONCAT: {
                Shape s = in_shape(0);
                int axis = go.ip[0];
                sanm_check(axis >= 1 && axis < s.rank,
                           "concat axis %d unsupported (only per-item axes)", axis);
                uint64_t tot = 0;
                for (size_t i = 0; i < lo.in.size(); ++i) {
                    const Shape& si = in_shape(i);
                    sanm_check(si.rank == s.rank, "concat rank mismatch");
                    for (int d = 0; d < s.rank; ++d) {
                        sanm_check(d == axis || si.dim[d] == s.dim[d],
                                   "concat shape mismatch");
                    }
                    tot += si.dim[axis];
                }
                s.dim[axis] = tot;
                o0.shape = s;
                break;
            }
        }
        for (int v : lo.out) {
            auto& lv = L.vars[v];
            sanm_check(lv.shape.rank >= 1 && lv.shape.dim[0] == L.batch,
                       "var shape %s: batch mismatch", lv.shape.str().c_str());
            set_dims(lv);
            sanm_check(lv.e <= MAXE, "per-item size %d above %d", lv.e, MAXE);
            lv.is_const = all_const;
        }
        lo.const_op = all_const;
        L.ops.emplace_back(std::move(lo));
    }
    sanm_check(L.input >= 0, "input not found");
    L.output = L.gvar2lvar[out_var];
    sanm_check(!L.vars[L.output].is_const || true);

    // stack / cache requirements
    bool elemwise = true;
    for (auto& lo : L.ops) {
        const GraphOp& go = g.ops[lo.gop];
        auto nc = [&](int i) { return !L.vars[lo.in[i]].is_const; };
        auto eo = [&](int i) { return L.vars[lo.out[i]].e; };
        if (lo.const_op && lo.type != OP_CONSTANT) {
            // still evaluated at order 0; caches of that evaluation are kept
        }
        switch (lo.type) {
            case OP_PLACEHOLDER:
            case OP_CONSTANT:
                break;
            case OP_LINCOMB:
                for (size_t i = 0; i < lo.in.size(); ++i) {
                    if (L.vars[lo.in[i]].e != eo(0)) elemwise = false;
                }
                break;
            case OP_MULTIPLY:
                if (nc(0) && nc(1)) {
                    L.vars[lo.in[0]].needs_stack = true;
                    L.vars[lo.in[1]].needs_stack = true;
                    lo.esb = eo(0);
                }
                for (int i = 0; i < 2; ++i) {
                    if (L.vars[lo.in[i]].e != eo(0)) elemwise = false;
                }
                break;
            case OP_POW:
            case OP_LOG:
                lo.eaux0 = eo(0);
                if (nc(0)) {
                    L.vars[lo.in[0]].needs_stack = true;
                    L.vars[lo.out[0]].needs_stack = true;
                    lo.esb = eo(0);
                }
                lo.need_flag = lo.type == OP_POW;
                (void)go;
                break;
            case OP_MATMUL:
                elemwise = false;
                if (nc(0) && nc(1)) {
                    L.vars[lo.in[0]].needs_stack = true;
                    L.vars[lo.in[1]].needs_stack = true;
                    lo.esb = eo(0);
                }
                break;
            case OP_DET:
                elemwise = false;
                lo.eaux0 = L.vars[lo.in[0]].e;
                if (nc(0)) {
                    L.vars[lo.in[0]].needs_stack = true;
                    lo.esb = 1;
                    if (L.vars[lo.in[0]].d1 == 3) lo.eauxk = 3;
                    // m >= 4: [G_j, T_j, D_j] per order (taylor_dev.h op_det)
                    if (L.vars[lo.in[0]].d1 >= 4) lo.eauxk = L.vars[lo.in[0]].e + 2;
                }
                break;
            case OP_MATINVMUL:
                elemwise = false;
                if (lo.in.size() > 1) lo.eaux0 = eo(0);
                if (nc(0)) {
                    L.vars[lo.in[0]].needs_stack = true;
                    L.vars[lo.out[0]].needs_stack = true;
                    lo.esb = eo(0);
                }
                break;
            case OP_SVDW:
                elemwise = false;
                // [U0, S0]; self-bias [D, Bpw]; hidden stack P_k (n x n items)
                lo.eaux0 = eo(0) + eo(1);
                if (nc(0)) {
                    L.vars[lo.in[0]].needs_stack = true;
                    L.vars[lo.out[2]].needs_stack = true;
                    lo.esb = 2 * eo(0);
                    lo.eauxk = eo(0);
                }
                break;
            default:
                elemwise = false;
                break;
        }
    }
    L.elemwise = elemwise &&
                 L.vars[L.input].e == L.vars[L.output].e;

    // SVD-W mode (libsanm/oprs/linalg.cpp:529-543): polar when neither U nor S
    // of an SVD-W that depends on the input is read, full otherwise
    std::vector<int> readers(L.vars.size(), 0);
    for (auto& lo : L.ops) {
        for (int v : lo.in) ++readers[v];
    }
    ++readers[L.output];
    for (auto& lo : L.ops) {
        if (lo.type == OP_SVDW && !lo.const_op &&
            (readers[lo.out[0]] || readers[lo.out[1]])) {
            sanm_check(L.vars[lo.out[0]].e == 9,
                       "full-mode SVD-W (U or S read) is implemented for 3x3 "
                       "matrices only; the polar mode (W only) for any size");
            lo.svdw_full = true;
            for (int q = 0; q < 3; ++q) L.vars[lo.out[q]].needs_stack = true;
            lo.esb = 27;    // Bu, Bw, Mbias
            lo.eauxk = 18;  // (U S)_k, (U S U^T)_k
        }
    }
    return L;
}

}  // namespace sanm_b200
