// cxx/sanm_host/tensor.hpp -- host tensor currency of the libsanm-compatible
// layer: TensorShape / TensorND / TensorArray / StSparseLinearTrans with the
// interface of libsanm/tensor.h:17-604.
//
// These are the HOST views that application code (fea/*.cpp, tests) reads and
// writes through ptr()/woptr()/rwptr(); the batched Taylor data itself lives
// in HBM behind the C ABI (include/sanm_b200.h) and is only copied into a
// TensorND when a caller asks for it.  The small dense helpers below
// (batched_mm, batched_matinv, ...) are set-up / post-processing utilities of
// the application (e.g. Dm^-1 once per mesh, fea/mesh_template.h:206), not
// the propagation path.
#pragma once

#include "sanm_host/base.hpp"

#include <array>
#include <cstdio>
#include <initializer_list>
#include <memory>
#include <vector>

namespace sanm {

struct TensorShape {
    static constexpr size_t MAX_RANK = 3;
    leastsize_t rank = 0;  //!< 0 = uninitialised; scalars are {1}
    leastsize_t dim[MAX_RANK] = {0, 0, 0};

    TensorShape() = default;
    TensorShape(std::initializer_list<size_t> sizes);

    TensorShape flatten_batched() const;  //!< {dim0, prod(rest)}
    TensorShape flatten_as_vec() const;   //!< {prod(all)}
    TensorShape with_batch_replaced(size_t batch) const;
    size_t total_nr_elems() const;
    size_t total_nr_elems_per_batch() const;
    bool operator==(const TensorShape& rhs) const;
    bool operator!=(const TensorShape& rhs) const { return !(*this == rhs); }
    size_t operator[](size_t i) const {
        sanm_assert(i < rank);
        return dim[i];
    }
    bool is_batched_scalar() const {
        return rank == 1 || (rank == 2 && dim[1] == 1);
    }
    bool is_single_scalar() const { return rank == 1 && dim[0] == 1; }
    TensorShape add_axis(size_t axis) const;
    TensorShape remove_axis(size_t axis) const;
    std::string str() const;
};

// A block of doubles.  All-zero and all-one tensors share immutable pool
// blocks (kind != PLAIN), which is how is_zero()/is_one() are answered
// without looking at the data (libsanm/tensor.h:149-159 answers by pointer
// identity against one growing pool block; here every pool block carries its
// kind, so growing the pool never moves memory other tensors point to).
class TensorStorage : public NonCopyable {
public:
    enum class Constant { ZERO, ONE };
    enum class Kind : uint8_t { PLAIN, ZERO, ONE };

    TensorStorage() = default;
    ~TensorStorage();
    void ensure_size(size_t size);
    fp_t* ptr() const { return m_data; }
    size_t size() const { return m_cap; }
    Kind kind() const { return m_kind; }

    //! shared immutable block of at least \p size zeros / ones
    template <Constant val>
    static const std::shared_ptr<const TensorStorage>& constant(
            size_t size = 0);

private:
    fp_t* m_data = nullptr;
    size_t m_cap = 0;
    Kind m_kind = Kind::PLAIN;
};

class TensorND {
public:
    TensorND() = default;
    TensorND(const TensorShape& shape) { set_shape(shape); }

    void clear() {
        m_shape.rank = 0;
        detach(false);
    }
    TensorND& set_shape(const TensorShape& shape);
    bool same_storage(const TensorND& o) const { return m_buf == o.m_buf; }
    bool is_zero() const {
        return m_buf && m_buf->kind() == TensorStorage::Kind::ZERO;
    }
    bool is_one() const {
        return m_buf && m_buf->kind() == TensorStorage::Kind::ONE;
    }
    bool empty() const { return m_shape.rank == 0; }
    const TensorShape& shape() const { return m_shape; }
    size_t shape(size_t i) const { return m_shape[i]; }
    size_t shape(size_t i, size_t default_) const {
        return i < m_shape.rank ? m_shape.dim[i] : default_;
    }
    size_t rank() const { return m_shape.rank; }

    fp_t* rwptr() {
        detach(true);
        sanm_assert(m_buf);
        return m_buf->ptr();
    }
    fp_t* woptr() {
        detach(false);
        sanm_assert(m_buf);
        return m_buf->ptr();
    }
    const fp_t* ptr() const {
        sanm_assert(m_buf);
        return m_buf->ptr();
    }

    TensorND& as_reshaped(const TensorND& src, const TensorShape& shape);
    TensorND reshape(const TensorShape& shape) const {
        TensorND r;
        r.as_reshaped(*this, shape);
        return r;
    }
    TensorND& reshape_inplace(const TensorShape& shape) {
        return as_reshaped(*this, shape);
    }
    TensorND& incr_sub_batch(const TensorND& src, size_t dbatch, size_t sbatch,
                             size_t size, fp_t alpha, fp_t beta);
    TensorND& copy_from_sub_batch(const TensorND& src, size_t dbatch,
                                  size_t sbatch, size_t size) {
        return incr_sub_batch(src, dbatch, sbatch, size, 0, 1);
    }
    TensorND take_sub(const TensorShape& shape) const;
    TensorND flatten_batched() const {
        return reshape(m_shape.flatten_batched());
    }
    TensorND flatten_as_vec() const {
        return reshape(m_shape.flatten_as_vec());
    }
    TensorND fill_with(fp_t value) const {
        return make_same_shape().fill_with_inplace(value);
    }
    TensorND& fill_with_inplace(fp_t value);
    TensorND make_same_shape() const { return TensorND{}.set_shape(shape()); }

    TensorND operator*(fp_t scale) const;
    TensorND& operator*=(fp_t scale);
    TensorND operator/(fp_t x) const { return operator*(1 / x); }
    TensorND& operator/=(fp_t x) { return operator*=(1 / x); }
    TensorND operator+(const TensorND& rhs) const;
    TensorND operator-(const TensorND& rhs) const;
    TensorND operator*(const TensorND& rhs) const;
    TensorND operator/(const TensorND& rhs) const;
    TensorND& operator+=(const TensorND& rhs);
    TensorND& operator-=(const TensorND& rhs);
    TensorND& operator*=(const TensorND& rhs);
    TensorND& operator/=(const TensorND& rhs);
    template <char op>
    TensorND& as_elem(const TensorND& lhs, const TensorND& rhs);
    TensorND operator-() const {
        TensorND r;
        r.as_neg(*this);
        return r;
    }
    TensorND& as_fma(const TensorND& x, const TensorND& y, const TensorND& b);
    TensorND& accum_mul(const TensorND& lhs, const TensorND& rhs,
                        fp_t scale = 1);
    TensorND& accum_mul(const TensorND& lhs, fp_t rhs);
    TensorND& as_neg(const TensorND& x);
    TensorND& inplace_neg() { return as_neg(*this); }
    TensorND& as_pow(const TensorND& x, fp_t exp);
    TensorND pow(fp_t exp) const {
        TensorND r;
        r.as_pow(*this, exp);
        return r;
    }
    TensorND& as_log(const TensorND& x);
    TensorND log() const {
        TensorND r;
        r.as_log(*this);
        return r;
    }

    TensorND& as_mm(const TensorND& lhs, const TensorND& rhs,
                    bool accum = false, bool trans_lhs = false,
                    bool trans_rhs = false);
    TensorND& as_batched_mm(const TensorND& lhs, const TensorND& rhs,
                            bool accum = false, bool trans_lhs = false,
                            bool trans_rhs = false);
    TensorND batched_mm(const TensorND& rhs, bool trans_self = false,
                        bool trans_rhs = false) const {
        return TensorND{}.as_batched_mm(*this, rhs, false, trans_self,
                                        trans_rhs);
    }
    TensorND& as_batched_transpose(const TensorND& src);
    TensorND& as_transpose(const TensorND& src);
    TensorND batched_transpose() const {
        return TensorND{}.as_batched_transpose(*this);
    }
    TensorND& as_batched_matinv(const TensorND& src);
    TensorND batched_matinv() const {
        return TensorND{}.as_batched_matinv(*this);
    }
    TensorND& as_batched_determinant(const TensorND& src);
    TensorND batched_determinant() const {
        return TensorND{}.as_batched_determinant(*this);
    }
    TensorND& as_batched_cofactor(const TensorND& src);
    TensorND batched_cofactor() const {
        return TensorND{}.as_batched_cofactor(*this);
    }
    TensorND& as_batched_mm_vecitem_left(const TensorND& lhs,
                                         const TensorND& rhs,
                                         bool accum = false);
    //! M = U diag(S) U^T W (libsanm/tensor.h:391-392)
    const TensorND& compute_batched_svd_w(TensorND& u, TensorND& s, TensorND& w,
                                          bool require_rotation = false) const;
    std::array<TensorND, 3> batched_svd_w(bool require_rotation = false) const {
        TensorND u, s, w;
        compute_batched_svd_w(u, s, w, require_rotation);
        return {u, s, w};
    }
    TensorND& as_diag(const TensorND& src);
    TensorND& as_batched_diag(const TensorND& src);
    TensorND& as_batched_diag_scalar(const TensorND& src, size_t dim);
    TensorND& as_batched_trace(const TensorND& src);
    TensorND& as_reduce_sum(const TensorND& src, int axis, bool keepdim);
    TensorND& as_broadcast(const TensorND& src, size_t axis, size_t size);
    TensorND reduce_sum(int axis, bool keepdim = true) const {
        return TensorND{}.as_reduce_sum(*this, axis, keepdim);
    }
    fp_t norm_l2() const;
    fp_t squared_norm_l2() const;
    fp_t norm_rms() const;
    fp_t flat_dot(const TensorND& rhs) const;
    void assert_allclose(const char* msg, const TensorND& rhs,
                         fp_t eps = 1e-4) const;

private:
    std::shared_ptr<TensorStorage> m_buf;
    TensorShape m_shape;

    //! make the block private to this tensor (copy the content if \p keep)
    void detach(bool keep);
};
using TensorArray = std::vector<TensorND>;

TensorShape infer_reduce_shape(const TensorShape& src, int axis, bool keepdim);

//! coefficient of a^order in det(sum_i a^i A_i), libsanm/tensor.h:493-502
TensorND compute_polymat_det_coeff(const TensorArray& a, size_t order);

class StSparseLinearTrans {
public:
    enum Type { INVALID, ELEMWISE, FULL };

    StSparseLinearTrans() = default;
    StSparseLinearTrans(Type type, bool is_batched, const TensorND& coeff) {
        reset(type, is_batched, coeff);
    }
    bool valid() const { return m_type != INVALID; }
    Type type() const { return m_type; }
    bool is_batched() const {
        sanm_assert(valid());
        return m_batched;
    }
    size_t batch() const {
        sanm_assert(is_batched());
        return m_batch;
    }
    size_t out_dim() const {
        sanm_assert(valid());
        return m_odim;
    }
    size_t inp_dim() const {
        sanm_assert(valid());
        return m_idim;
    }
    const TensorND& coeff() const {
        sanm_assert(valid());
        return m_coeff;
    }
    StSparseLinearTrans& reset(Type type, bool is_batched,
                               const TensorND& coeff);
    StSparseLinearTrans compose_with_elemwise(const TensorND& rhs) const;
    StSparseLinearTrans compose_with_full(const TensorND& rhs) const;
    StSparseLinearTrans compose_with_scaling(fp_t scale) const;
    StSparseLinearTrans& operator+=(const StSparseLinearTrans& rhs);
    StSparseLinearTrans& accum_full(bool is_batched, const TensorND& coeff) {
        return (*this) += StSparseLinearTrans{}.reset(FULL, is_batched, coeff);
    }
    TensorND as_full() const;
    TensorND apply(const TensorND& x) const;
    const StSparseLinearTrans& check_batched(bool req) const {
        sanm_assert(m_batched == req, "require batched %d, got %d", req,
                    m_batched);
        return *this;
    }

private:
    Type m_type = INVALID;
    bool m_batched = false;
    size_t m_batch = 0, m_odim = 0, m_idim = 0;
    TensorND m_coeff;
};

//! the device library is single-stream; these only record the request
//! (libsanm/tensor.h:601-602) and size the host-side analysis loops
int get_num_threads();
void set_num_threads(int nr);

namespace capi {
//! (rank, dims[]) as the C ABI wants them
struct Dims {
    int rank;
    uint64_t d[3];
    Dims(const TensorShape& s) : rank(s.rank), d{s.dim[0], s.dim[1], s.dim[2]} {}
};
TensorShape shape_from(int rank, const uint64_t* dims);
}  // namespace capi

}  // namespace sanm
