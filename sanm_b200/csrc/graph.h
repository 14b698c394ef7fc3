/* sanm_b200/csrc/graph.h -- host-side symbolic graph (the libsanm
 * ComputingGraph / SymbolVar surface, libsanm/symbolic.h:283-293 and
 * libsanm/oprs.h:14-105) kept as a flat operator list, plus the lowering of
 * the sub-graph reachable from one output variable to a device program.
 */
#pragma once

#include "rt.h"
#include "taylor_dev.h"

#include <memory>

namespace sanm_b200 {

struct Shape {
    int rank = 0;
    uint64_t dim[3] = {0, 0, 0};

    Shape() = default;
    Shape(std::initializer_list<uint64_t> d) {
        rank = 0;
        for (auto i : d) dim[rank++] = i;
    }
    Shape(int r, const uint64_t* d) : rank{r} {
        sanm_check(r >= 0 && r <= 3, "rank=%d", r);
        for (int i = 0; i < r; ++i) dim[i] = d[i];
    }
    uint64_t total() const {
        uint64_t t = rank ? 1 : 0;
        for (int i = 0; i < rank; ++i) t *= dim[i];
        return t;
    }
    uint64_t per_item() const {
        uint64_t t = 1;
        for (int i = 1; i < rank; ++i) t *= dim[i];
        return t;
    }
    bool is_batched_scalar() const {
        return rank == 1 || (rank == 2 && dim[1] == 1);
    }
    bool operator==(const Shape& o) const {
        if (rank != o.rank) return false;
        for (int i = 0; i < rank; ++i) {
            if (dim[i] != o.dim[i]) return false;
        }
        return true;
    }
    bool operator!=(const Shape& o) const { return !(*this == o); }
    std::string str() const;
};

struct GraphOp {
    taylor::OpType type;
    std::vector<int> in;
    std::vector<int> out;
    std::vector<double> f;  //!< LINCOMB coeffs + bias, POW exponent
    int ip[6] = {0, 0, 0, 0, 0, 0};
    //! CONSTANT payload, row-major [N, ...]
    std::vector<double> cdata;
    Shape cshape;
};

class Graph {
public:
    std::vector<GraphOp> ops;
    std::vector<int> var_owner;  //!< var id -> op index

    int add_op(GraphOp op, int nout);
    int placeholder();
    int constant(const double* data, const Shape& shape);
    int linear_combine(int n, const double* coeffs, const int* vars,
                       double bias);
    int multiply(int a, int b);
    int pow(int x, double e);
    int log(int x);
    int reduce_sum(int x, int axis, bool keepdim);
    int batched_matmul(int a, int b);
    int batched_transpose(int x);
    int batched_det(int x);
    int batched_mul_eye(int x, uint64_t dim);
    int batched_mat_inv_mul(int x, int a, bool is_left);
    void batched_svd_w(int x, bool require_rotation, int out[3]);
    int slice(int x, int axis, bool has_begin, int begin, bool has_end,
              int end, int stride);
    int concat(int n, const int* vars, int axis);

    void check_var(int v) const {
        sanm_check(v >= 0 && v < (int)var_owner.size(), "invalid var id %d", v);
    }
};

/*!
 * The program of one output variable for a fixed placeholder shape: operator
 * list in topological order, shapes, stack/aux layout.  Host description only;
 * taylor.cu turns it into device tables.
 */
struct LoweredVar {
    Shape shape;
    int e = 0, d1 = 1, d2 = 1;
    bool is_const = false;
    bool needs_stack = false;  //!< read by a convolution => keep orders 1..K
    bool live = false;
};

struct LoweredOp {
    taylor::OpType type;
    int gop;  //!< index into Graph::ops
    std::vector<int> in, out;  //!< lowered var ids
    bool const_op = false;
    int esb = 0;    //!< doubles of cached self-bias per item
    int eaux0 = 0;  //!< doubles of order-0 cache per item
    int eauxk = 0;  //!< doubles of hidden per-order stack per item
    bool need_flag = false;
    bool svdw_full = false;  //!< SVD-W: U or S has readers (full mode)
};

struct Lowered {
    std::vector<LoweredVar> vars;
    std::vector<LoweredOp> ops;
    std::vector<int> gvar2lvar;  //!< graph var id -> lowered var id (-1: unused)
    int input = -1, output = -1;
    uint64_t batch = 0;
    bool elemwise = false;  //!< every op is elementwise with no broadcast
};

Lowered lower_graph(const Graph& g, int out_var, const Shape& x_shape);

}  // namespace sanm_b200
