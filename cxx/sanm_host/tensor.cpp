// cxx/sanm_host/tensor.cpp -- bodies of the host tensor currency
// (interface: libsanm/tensor.h:17-604; semantics checked against
// libsanm/tensor.cpp, tensor_elemwise.cpp, tensor_linalg.cpp).  Plain loops:
// nothing here is on the propagation path, which runs in libsanm_b200.so.
#include <sys/mman.h>
#include "sanm_host/tensor.hpp"

#include "sanm_b200.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <numeric>

namespace sanm {

// ---------------------------------------------------------------- utilities
std::string svsprintf(const char* fmt, va_list ap0) {
    va_list ap;
    va_copy(ap, ap0);
    int n = vsnprintf(nullptr, 0, fmt, ap);
    va_end(ap);
    if (n < 0)
        throw std::runtime_error{"vsnprintf failed"};
    std::string out(static_cast<size_t>(n) + 1, '\0');
    va_copy(ap, ap0);
    vsnprintf(out.data(), out.size(), fmt, ap);
    va_end(ap);
    out.resize(static_cast<size_t>(n));
    return out;
}

std::string ssprintf(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    std::string r;
    try {
        r = svsprintf(fmt, ap);
    } catch (...) {
        va_end(ap);
        throw;
    }
    va_end(ap);
    return r;
}

void __assertion_failed(const char* expr, const char* file, const char* func,
                        int line, const char* fmt, ...) {
    std::string msg = ssprintf("assertion `%s' failed at %s:%s@%d", expr, file,
                               func, line);
    if (fmt) {
        va_list ap;
        va_start(ap, fmt);
        std::string extra;
        try {
            extra = svsprintf(fmt, ap);
        } catch (...) {
        }
        va_end(ap);
        msg += "\nmessage: " + extra;
    }
    throw SANMAssertionError{std::move(msg)};
}

namespace {
int g_threads = 1;
}
int get_num_threads() {
    return g_threads;
}
void set_num_threads(int nr) {
    sanm_assert(nr >= 1);
    g_threads = nr;
    // The reference brings its compute backend up before main() runs (Eigen /
    // MKL threading in a static constructor, libsanm/tensor.cpp:24-32).  The
    // equivalent here: create the CUDA context and load the kernel images now
    // (fea/main.cpp:1058 calls this first thing) rather than inside the first
    // timed solve.  SANM_DEVICE selects the GPU of this process.
    static bool device_up = false;
    if (!device_up) {
        const char* dev = getenv("SANM_DEVICE");
        capi::check(sanm_set_device(dev ? atoi(dev) : 0));
        device_up = true;
        // the reference prints its profiler tags at exit; so does this layer
        // (host wall time and the library's device time per tag)
        ScopedProfiler::get_node();
    }
}

TensorShape capi::shape_from(int rank, const uint64_t* dims) {
    TensorShape s;
    sanm_assert(rank >= 0 && rank <= 3);
    s.rank = rank;
    for (int i = 0; i < rank; ++i)
        s.dim[i] = static_cast<leastsize_t>(dims[i]);
    return s;
}

// -------------------------------------------------------------- TensorShape
TensorShape::TensorShape(std::initializer_list<size_t> sizes) {
    sanm_assert(sizes.size() <= MAX_RANK);
    for (size_t v : sizes) {
        sanm_assert(v <= std::numeric_limits<leastsize_t>::max());
        dim[rank++] = static_cast<leastsize_t>(v);
    }
}

size_t TensorShape::total_nr_elems() const {
    if (!rank)
        return 0;
    size_t n = 1;
    for (size_t i = 0; i < rank; ++i)
        n *= dim[i];
    return n;
}

size_t TensorShape::total_nr_elems_per_batch() const {
    if (!rank)
        return 0;
    size_t n = 1;
    for (size_t i = 1; i < rank; ++i)
        n *= dim[i];
    return n;
}

TensorShape TensorShape::flatten_batched() const {
    sanm_assert(rank);
    return {dim[0], total_nr_elems_per_batch()};
}

TensorShape TensorShape::flatten_as_vec() const {
    sanm_assert(rank);
    return {total_nr_elems()};
}

TensorShape TensorShape::with_batch_replaced(size_t batch) const {
    sanm_assert(rank);
    TensorShape r = *this;
    r.dim[0] = static_cast<leastsize_t>(batch);
    return r;
}

bool TensorShape::operator==(const TensorShape& o) const {
    if (rank != o.rank)
        return false;
    for (size_t i = 0; i < rank; ++i)
        if (dim[i] != o.dim[i])
            return false;
    return true;
}

TensorShape TensorShape::add_axis(size_t axis) const {
    sanm_assert(rank < MAX_RANK && axis <= rank, "rank=%zu axis=%zu",
                static_cast<size_t>(rank), axis);
    TensorShape r;
    r.rank = rank + 1;
    for (size_t i = 0, j = 0; i < r.rank; ++i)
        r.dim[i] = (i == axis) ? 1 : dim[j++];
    return r;
}

TensorShape TensorShape::remove_axis(size_t axis) const {
    sanm_assert(rank >= 2 && axis < rank && dim[axis] == 1,
                "can not remove axis %zu from shape %s", axis, str().c_str());
    TensorShape r;
    r.rank = rank - 1;
    for (size_t i = 0, j = 0; i < rank; ++i)
        if (i != axis)
            r.dim[j++] = dim[i];
    return r;
}

std::string TensorShape::str() const {
    std::string s = "{";
    for (size_t i = 0; i < rank; ++i) {
        if (i)
            s += ',';
        s += std::to_string(dim[i]);
    }
    return s + "}";
}

TensorShape infer_reduce_shape(const TensorShape& src, int axis,
                               bool keepdim) {
    if (axis == -1)
        return {src.dim[0], 1};
    sanm_assert(axis >= 0 && axis < static_cast<int>(src.rank),
                "invalid reduce axis: %d (shape %s)", axis, src.str().c_str());
    if (keepdim) {
        TensorShape r = src;
        r.dim[axis] = 1;
        return r;
    }
    TensorShape r;
    r.rank = src.rank - 1;
    for (size_t i = 0, j = 0; i < src.rank; ++i)
        if (static_cast<int>(i) != axis)
            r.dim[j++] = src.dim[i];
    return r;
}

// ------------------------------------------------------------ TensorStorage
TensorStorage::~TensorStorage() {
    if (m_kind == Kind::PLAIN)
        std::free(m_data);
    // pool storages live until exit (their address range is never returned)
}

void TensorStorage::ensure_size(size_t size) {
    if (size <= m_cap)
        return;
    // geometric growth, capped at 1 Mi elements of slack
    size_t want = std::min(std::max(m_cap * 2, size), size + (size_t(1) << 20));
    void* p = nullptr;
    if (int err = posix_memalign(&p, 64, want * sizeof(fp_t)))
        throw SANMError{ssprintf("host allocation of %zu bytes failed (%d)",
                                 want * sizeof(fp_t), err)};
    std::free(m_data);  // content is not preserved (callers rewrite it)
    m_data = static_cast<fp_t*>(p);
    m_cap = want;
}

/* The shared zero / one pools (libsanm/tensor.cpp:251-293).  As in the
 * reference there is ONE storage object per constant, so every all-zero tensor
 * has the same storage and the same pointer, whatever its size
 * (tests/tensor.cpp "Tensor.Constant").  The reference grows that storage by
 * reallocating it, which moves memory under concurrent readers (SURVEY section
 * 0).  Here each pool is a reserved address range (64 GiB of virtual memory,
 * committed page by page as it is touched) that never moves: growth only
 * extends the initialised prefix, under a mutex, and readers of a shorter
 * prefix are unaffected.  Anonymous pages read as zero, so the zero pool costs
 * nothing until someone reads it. */
template <TensorStorage::Constant val>
const std::shared_ptr<const TensorStorage>& TensorStorage::constant(
        size_t size) {
    static size_t RESERVE = size_t(1) << 33;  // elements
    static std::mutex mtx;
    static const std::shared_ptr<TensorStorage> pool = [] {
        auto blk = std::make_shared<TensorStorage>();
        void* p = MAP_FAILED;
        // strict-overcommit hosts may refuse the large range: halve until taken
        for (; RESERVE >= (size_t(1) << 24); RESERVE >>= 1) {
            p = mmap(nullptr, RESERVE * sizeof(fp_t), PROT_READ | PROT_WRITE,
                     MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
            if (p != MAP_FAILED)
                break;
        }
        if (p == MAP_FAILED)
            throw SANMError{"can not reserve the address range of a constant pool"};
        blk->m_data = static_cast<fp_t*>(p);
        blk->m_cap = 0;
        blk->m_kind = val == Constant::ZERO ? Kind::ZERO : Kind::ONE;
        return blk;
    }();
    static const std::shared_ptr<const TensorStorage> cpool = pool;
    if (size > pool->m_cap) {
        std::lock_guard<std::mutex> lk{mtx};
        if (size > pool->m_cap) {
            sanm_assert(size <= RESERVE, "constant tensor of %zu elements", size);
            size_t want = std::max(size, pool->m_cap * 2);
            want = std::min(want, RESERVE);
            if (val == Constant::ONE)
                std::fill(pool->m_data + pool->m_cap, pool->m_data + want, 1.);
            pool->m_cap = want;
        }
    }
    return cpool;
}
template const std::shared_ptr<const TensorStorage>&
        TensorStorage::constant<TensorStorage::Constant::ZERO>(size_t);
template const std::shared_ptr<const TensorStorage>&
        TensorStorage::constant<TensorStorage::Constant::ONE>(size_t);

// ----------------------------------------------------------------- TensorND
void TensorND::detach(bool keep) {
    if (!m_buf)
        return;
    // pool blocks (zero / one) are immutable even when nobody else holds them
    const bool shared = m_buf.use_count() > 1 ||
                        m_buf->kind() != TensorStorage::Kind::PLAIN;
    if (!shared)
        return;
    size_t n = m_shape.total_nr_elems();
    auto fresh = std::make_shared<TensorStorage>();
    fresh->ensure_size(n);
    if (keep && n)
        std::memcpy(fresh->ptr(), m_buf->ptr(), n * sizeof(fp_t));
    m_buf = std::move(fresh);
}

TensorND& TensorND::set_shape(const TensorShape& shape) {
    if (!m_buf || m_buf.use_count() > 1 ||
        m_buf->kind() != TensorStorage::Kind::PLAIN)
        m_buf = std::make_shared<TensorStorage>();
    m_buf->ensure_size(shape.total_nr_elems());
    m_shape = shape;
    return *this;
}

TensorND& TensorND::as_reshaped(const TensorND& src, const TensorShape& shape) {
    sanm_assert(src.m_shape.total_nr_elems() == shape.total_nr_elems(),
                "can not reshape from %s to %s", src.m_shape.str().c_str(),
                shape.str().c_str());
    if (this != &src)
        m_buf = src.m_buf;
    m_shape = shape;
    return *this;
}

TensorND TensorND::take_sub(const TensorShape& shape) const {
    sanm_assert(shape.total_nr_elems() <= m_shape.total_nr_elems(),
                "can not take subtensor of shape %s from %s",
                shape.str().c_str(), m_shape.str().c_str());
    TensorND r;
    r.m_buf = m_buf;
    r.m_shape = shape;
    return r;
}

TensorND& TensorND::incr_sub_batch(const TensorND& src, size_t dbatch,
                                   size_t sbatch, size_t size, fp_t alpha,
                                   fp_t beta) {
    sanm_assert(this != &src, "inplace not allowed");
    sanm_assert(sbatch + size <= src.shape(0));
    if (!dbatch && empty()) {
        set_shape(src.shape().with_batch_replaced(size));
    } else {
        sanm_assert(m_shape == src.shape().with_batch_replaced(shape(0)));
        sanm_assert(dbatch + size <= m_shape.dim[0]);
    }
    size_t per = m_shape.total_nr_elems_per_batch(), n = size * per;
    fp_t* d = rwptr() + dbatch * per;
    const fp_t* s = src.ptr() + sbatch * per;
    if (alpha == 0 && beta == 1) {
        std::memcpy(d, s, n * sizeof(fp_t));
    } else {
        for (size_t i = 0; i < n; ++i)
            d[i] = d[i] * alpha + s[i] * beta;
    }
    return *this;
}

TensorND& TensorND::fill_with_inplace(fp_t value) {
    size_t n = m_shape.total_nr_elems();
    using C = TensorStorage::Constant;
    if (value == 0.) {
        m_buf = std::const_pointer_cast<TensorStorage>(
                TensorStorage::constant<C::ZERO>(n));
    } else if (value == 1.) {
        m_buf = std::const_pointer_cast<TensorStorage>(
                TensorStorage::constant<C::ONE>(n));
    } else {
        std::fill_n(woptr(), n, value);
    }
    return *this;
}

TensorND TensorND::operator*(fp_t scale) const {
    if (scale == 0.)
        return fill_with(0);
    if (scale == 1. || is_zero())
        return *this;
    TensorND r = make_same_shape();
    size_t n = m_shape.total_nr_elems();
    const fp_t* s = ptr();
    fp_t* d = r.woptr();
    for (size_t i = 0; i < n; ++i)
        d[i] = s[i] * scale;
    return r;
}

TensorND& TensorND::operator*=(fp_t scale) {
    if (scale == 0.)
        return fill_with_inplace(0);
    if (is_zero() || scale == 1.)
        return *this;
    size_t n = m_shape.total_nr_elems();
    fp_t* d = rwptr();
    for (size_t i = 0; i < n; ++i)
        d[i] *= scale;
    return *this;
}

TensorND& TensorND::as_neg(const TensorND& x) {
    if (x.is_zero()) {
        *this = x;
        return *this;
    }
    size_t n = x.m_shape.total_nr_elems();
    if (this == &x) {
        fp_t* d = rwptr();
        for (size_t i = 0; i < n; ++i)
            d[i] = -d[i];
    } else {
        const fp_t* s = x.ptr();
        fp_t* d = set_shape(x.shape()).woptr();
        for (size_t i = 0; i < n; ++i)
            d[i] = -s[i];
    }
    return *this;
}

TensorND& TensorND::as_pow(const TensorND& x, fp_t exp) {
    if (x.is_zero()) {
        sanm_assert(exp > 0);
        *this = x;
        return *this;
    }
    size_t n = x.m_shape.total_nr_elems();
    TensorND keep = x;  // keeps the source alive when this == &x
    const fp_t* s = keep.ptr();
    fp_t* d = (this == &x) ? rwptr() : set_shape(x.shape()).woptr();
    s = (this == &x) ? d : s;
    auto loop = [&](auto f) {
        for (size_t i = 0; i < n; ++i)
            d[i] = f(s[i]);
    };
    if (exp == 2.)
        loop([](fp_t v) { return v * v; });
    else if (exp == 3.)
        loop([](fp_t v) { return v * v * v; });
    else if (exp == 1.)
        loop([](fp_t v) { return v; });
    else if (exp == -1.)
        loop([](fp_t v) { return 1 / v; });
    else if (exp == -2.)
        loop([](fp_t v) { return 1 / (v * v); });
    else if (exp == -3.)
        loop([](fp_t v) { return 1 / (v * v * v); });
    else if (exp == .5)
        loop([](fp_t v) { return std::sqrt(v); });
    else if (exp == -.5)
        loop([](fp_t v) { return 1 / std::sqrt(v); });
    else
        loop([exp](fp_t v) { return std::pow(v, exp); });
    return *this;
}

TensorND& TensorND::as_log(const TensorND& x) {
    if (x.is_one()) {
        set_shape(x.shape());
        return fill_with_inplace(0);
    }
    size_t n = x.m_shape.total_nr_elems();
    TensorND keep = x;
    fp_t* d = (this == &x) ? rwptr() : set_shape(x.shape()).woptr();
    const fp_t* s = (this == &x) ? d : keep.ptr();
    for (size_t i = 0; i < n; ++i)
        d[i] = std::log(s[i]);
    return *this;
}

// ---- broadcasting elementwise binary ops -----------------------------------
namespace {

//! output shape of a broadcasting binary op (scalar, batched scalar, or
//! per-axis 1 -> n), as libsanm/tensor_elemwise.cpp:12-50 accepts them
TensorShape bcast_shape(const TensorShape& a, const TensorShape& b) {
    sanm_assert(a.rank && b.rank);
    if (a == b)
        return a;
    if (a.is_single_scalar())
        return b;
    if (b.is_single_scalar())
        return a;
    if (a.dim[0] == b.dim[0]) {
        if (a.is_batched_scalar())
            return b;
        if (b.is_batched_scalar())
            return a;
    }
    if (a.rank == b.rank) {
        TensorShape r;
        r.rank = a.rank;
        bool ok = true;
        for (size_t i = 0; i < a.rank && ok; ++i) {
            if (a.dim[i] == b.dim[i] || b.dim[i] == 1)
                r.dim[i] = a.dim[i];
            else if (a.dim[i] == 1)
                r.dim[i] = b.dim[i];
            else
                ok = false;
        }
        if (ok)
            return r;
    }
    throw SANMError{ssprintf("tensor shape mismatch for elemwise op: %s vs %s",
                             a.str().c_str(), b.str().c_str())};
}

//! element strides of operand \p s when broadcast to \p out (rank padded to 3)
void bcast_strides(const TensorShape& s, const TensorShape& out,
                   size_t strides[3]) {
    size_t d[3] = {1, 1, 1};
    if (s.is_batched_scalar() && s.rank != out.rank) {
        d[0] = s.dim[0];  // {b} or {b,1} against {b,...}; {1} against anything
    } else {
        sanm_assert(s.rank == out.rank);
        for (size_t i = 0; i < s.rank; ++i)
            d[i] = s.dim[i];
    }
    size_t mul = 1;
    for (int i = 2; i >= 0; --i) {
        strides[i] = d[i] == 1 ? 0 : mul;
        mul *= d[i];
    }
}

template <bool accum, class F>
void bcast_apply(fp_t* dst, const TensorShape& out, const fp_t* l,
                 const TensorShape& ls, const fp_t* r, const TensorShape& rs,
                 F f) {
    size_t n = out.total_nr_elems();
    if (ls == out && rs == out) {
        for (size_t i = 0; i < n; ++i) {
            fp_t v = f(l[i], r[i]);
            if (accum)
                dst[i] += v;
            else
                dst[i] = v;
        }
        return;
    }
    size_t o[3] = {1, 1, 1}, sl[3], sr[3];
    for (size_t i = 0; i < out.rank; ++i)
        o[i] = out.dim[i];
    bcast_strides(ls, out, sl);
    bcast_strides(rs, out, sr);
    size_t p = 0;
    for (size_t i = 0; i < o[0]; ++i)
        for (size_t j = 0; j < o[1]; ++j)
            for (size_t k = 0; k < o[2]; ++k, ++p) {
                fp_t v = f(l[i * sl[0] + j * sl[1] + k * sl[2]],
                           r[i * sr[0] + j * sr[1] + k * sr[2]]);
                if (accum)
                    dst[p] += v;
                else
                    dst[p] = v;
            }
}

template <char op>
struct BinOp {
    fp_t scale = 1;
    fp_t operator()(fp_t a, fp_t b) const {
        if constexpr (op == '+')
            return a + b;
        else if constexpr (op == '-')
            return a - b;
        else if constexpr (op == '*')
            return a * b;
        else if constexpr (op == '/')
            return a / b;
        else
            return a * b * scale;  // 's': scaled product
    }
};

template <char op, bool accum>
void binary(TensorND& dst, const TensorND& lhs, const TensorND& rhs,
            fp_t scale = 1) {
    TensorShape out = bcast_shape(lhs.shape(), rhs.shape());
    const bool alias = (&dst == &lhs) || (&dst == &rhs);
    if (alias) {
        sanm_assert((&dst != &lhs || out == lhs.shape()) &&
                            (&dst != &rhs || out == rhs.shape()),
                    "inplace elemwise shape mismatch: %s vs %s",
                    lhs.shape().str().c_str(), rhs.shape().str().c_str());
    } else if (accum) {
        if (dst.shape() != out) {
            TensorShape wide = bcast_shape(dst.shape(), out);
            sanm_assert(dst.shape() == wide,
                        "accum dst shape mismatch: %s vs %s",
                        dst.shape().str().c_str(), out.str().c_str());
        }
    }

    // structural zero shortcuts keep the zero flag alive
    bool lz = lhs.is_zero(), rz = rhs.is_zero();
    constexpr bool additive = (op == '+' || op == '-');
    const TensorShape& final_shape = (accum || alias) ? dst.shape() : out;
    if (additive) {
        if (lz && final_shape != rhs.shape())
            lz = false;
        if (rz && final_shape != lhs.shape())
            rz = false;
    }
    if (lz || rz) {
        if constexpr (op == '+') {
            TensorND v = lz ? rhs : lhs;
            if (accum)
                dst += v;
            else
                dst = v;
            return;
        } else if constexpr (op == '-') {
            TensorND v = rz ? lhs : -rhs;
            if (accum)
                dst += v;
            else
                dst = v;
            return;
        } else {
            if constexpr (op == '/')
                sanm_assert(!rz, "division by zero");
            if (!accum) {
                if (!alias)
                    dst.set_shape(out);
                dst.fill_with_inplace(0);
            }
            return;
        }
    }

    // keep the sources alive across detach -- but not the aliased one: a
    // second reference to dst's own buffer would force rwptr() to copy it, and
    // `a -= b` must keep a's storage (tests/tensor.cpp "Tensor.SimpleArith")
    TensorND l, r;
    if (&dst != &lhs)
        l = lhs;
    if (&dst != &rhs)
        r = rhs;
    fp_t* d;
    if (alias || accum) {
        d = dst.rwptr();
    } else {
        d = dst.set_shape(out).woptr();
    }
    const fp_t* lp = (&dst == &lhs) ? d : l.ptr();
    const fp_t* rp = (&dst == &rhs) ? d : r.ptr();
    BinOp<op> f;
    f.scale = scale;
    bcast_apply<accum>(d, dst.shape(), lp, lhs.shape(), rp, rhs.shape(), f);
}

}  // namespace

#define SANM_HOST_BINOP(ch, sym)                                           \
    template <>                                                            \
    TensorND& TensorND::as_elem<ch>(const TensorND& l, const TensorND& r) { \
        binary<ch, false>(*this, l, r);                                    \
        return *this;                                                      \
    }                                                                      \
    TensorND TensorND::operator sym(const TensorND& rhs) const {           \
        TensorND out;                                                      \
        out.as_elem<ch>(*this, rhs);                                       \
        return out;                                                        \
    }                                                                      \
    TensorND& TensorND::operator sym##=(const TensorND& rhs) {             \
        return as_elem<ch>(*this, rhs);                                    \
    }
SANM_HOST_BINOP('+', +)
SANM_HOST_BINOP('-', -)
SANM_HOST_BINOP('*', *)
SANM_HOST_BINOP('/', /)
#undef SANM_HOST_BINOP

TensorND& TensorND::accum_mul(const TensorND& lhs, const TensorND& rhs,
                              fp_t scale) {
    if (empty()) {
        *this = lhs * rhs * scale;
    } else if (scale == 1.) {
        binary<'*', true>(*this, lhs, rhs);
    } else {
        if (lhs.is_zero() || rhs.is_zero())
            return *this;
        binary<'s', true>(*this, lhs, rhs, scale);
    }
    return *this;
}

TensorND& TensorND::accum_mul(const TensorND& lhs, fp_t rhs) {
    if (empty() || (is_zero() && m_shape == lhs.shape()))
        return *this = lhs * rhs;
    TensorShape wide = bcast_shape(m_shape, lhs.shape());
    sanm_assert(wide == m_shape, "can not accum tensor of shape %s into %s",
                lhs.shape().str().c_str(), m_shape.str().c_str());
    if (lhs.is_zero() || rhs == 0.)
        return *this;
    TensorND l = lhs;
    fp_t* d = rwptr();
    TensorShape one{1};
    bcast_apply<true>(d, m_shape, l.ptr(), l.shape(), &rhs, one, BinOp<'*'>{});
    return *this;
}

TensorND& TensorND::as_fma(const TensorND& x, const TensorND& y,
                           const TensorND& b) {
    sanm_assert(x.shape() == y.shape() && x.shape() == b.shape(),
                "shape mismatch in FMA: %s %s %s", x.shape().str().c_str(),
                y.shape().str().c_str(), b.shape().str().c_str());
    if (x.is_zero() || y.is_zero()) {
        *this = b;
        return *this;
    }
    if (x.is_one())
        return as_elem<'+'>(y, b);
    if (y.is_one())
        return as_elem<'+'>(x, b);
    TensorND xs = x, ys = y, bs = b;
    size_t n = xs.shape().total_nr_elems();
    fp_t* d = set_shape(xs.shape()).woptr();
    const fp_t *px = xs.ptr(), *py = ys.ptr(), *pb = bs.ptr();
    for (size_t i = 0; i < n; ++i)
        d[i] = px[i] * py[i] + pb[i];
    return *this;
}

// ---- reductions -------------------------------------------------------------
TensorND& TensorND::as_reduce_sum(const TensorND& src, int axis, bool keepdim) {
    sanm_assert(this != &src, "inplace not allowed");
    if (axis == -2) {
        sanm_assert(src.rank());
        set_shape({1});
        if (src.is_zero())
            return fill_with_inplace(0);
        const fp_t* s = src.ptr();
        fp_t acc = 0;
        for (size_t i = 0, n = src.shape().total_nr_elems(); i < n; ++i)
            acc += s[i];
        woptr()[0] = acc;
        return *this;
    }
    TensorShape oshp = infer_reduce_shape(src.shape(), axis, keepdim);
    set_shape(oshp);
    if (src.is_zero())
        return fill_with_inplace(0);
    size_t before = 1, ax = 1, after = 1;
    if (axis == -1) {
        before = src.shape(0);
        ax = src.shape().total_nr_elems_per_batch();
    } else {
        ax = src.shape(axis);
        for (int i = 0; i < axis; ++i)
            before *= src.shape(i);
        for (size_t i = axis + 1; i < src.rank(); ++i)
            after *= src.shape(i);
    }
    const fp_t* s = src.ptr();
    fp_t* d = woptr();
    for (size_t i = 0; i < before; ++i)
        for (size_t k = 0; k < after; ++k) {
            fp_t acc = 0;
            for (size_t j = 0; j < ax; ++j)
                acc += s[(i * ax + j) * after + k];
            d[i * after + k] = acc;
        }
    return *this;
}

TensorND& TensorND::as_broadcast(const TensorND& src, size_t axis,
                                 size_t size) {
    sanm_assert(this != &src, "inplace not allowed");
    sanm_assert(axis < src.rank() && src.shape(axis) == 1 && size > 0);
    if (size == 1) {
        *this = src;
        return *this;
    }
    TensorShape oshp = src.shape();
    oshp.dim[axis] = static_cast<leastsize_t>(size);
    set_shape(oshp);
    if (src.is_zero())
        return fill_with_inplace(0);
    size_t before = 1, after = 1;
    for (size_t i = 0; i < axis; ++i)
        before *= src.shape(i);
    for (size_t i = axis + 1; i < src.rank(); ++i)
        after *= src.shape(i);
    const fp_t* s = src.ptr();
    fp_t* d = woptr();
    for (size_t i = 0; i < before; ++i)
        for (size_t j = 0; j < size; ++j)
            std::memcpy(d + (i * size + j) * after, s + i * after,
                        after * sizeof(fp_t));
    return *this;
}

// Pairwise (cascade) summation: error grows with log n instead of n, and
// the result does not depend on any SIMD width.
namespace {
fp_t pairwise_dot(const fp_t* a, const fp_t* b, size_t n) {
    if (n <= 64) {
        fp_t s = 0;
        for (size_t i = 0; i < n; ++i)
            s += a[i] * b[i];
        return s;
    }
    size_t h = n / 2;
    return pairwise_dot(a, b, h) + pairwise_dot(a + h, b + h, n - h);
}
}  // namespace

fp_t TensorND::squared_norm_l2() const {
    if (is_zero())
        return 0;
    return pairwise_dot(ptr(), ptr(), m_shape.total_nr_elems());
}

fp_t TensorND::norm_l2() const {
    return std::sqrt(squared_norm_l2());
}

fp_t TensorND::norm_rms() const {
    return std::sqrt(squared_norm_l2() / m_shape.total_nr_elems());
}

fp_t TensorND::flat_dot(const TensorND& rhs) const {
    sanm_assert(m_shape.total_nr_elems() == rhs.shape().total_nr_elems());
    if (is_zero() || rhs.is_zero())
        return 0;
    return pairwise_dot(ptr(), rhs.ptr(), m_shape.total_nr_elems());
}

void TensorND::assert_allclose(const char* msg, const TensorND& rhs,
                               fp_t eps) const {
    sanm_assert(shape() == rhs.shape(), "%s: shape mismatch: %s vs %s", msg,
                shape().str().c_str(), rhs.shape().str().c_str());
    const fp_t *a = ptr(), *b = rhs.ptr();
    for (size_t i = 0, n = m_shape.total_nr_elems(); i < n; ++i) {
        fp_t lim = eps * std::max<fp_t>(
                                 std::min(std::fabs(a[i]), std::fabs(b[i])), 1);
        sanm_assert(std::fabs(a[i] - b[i]) < lim,
                    "%s: value mismatch at index %zu on shape %s: v0=%g v1=%g "
                    "(eps=%g)",
                    msg, i, shape().str().c_str(), a[i], b[i], eps);
    }
}

// ---- small dense helpers ----------------------------------------------------
namespace {

//! C (m x n) (+)= op(A) op(B), row-major, op = transpose when flagged
void gemm(fp_t* c, const fp_t* a, const fp_t* b, size_t m, size_t k, size_t n,
          bool ta, bool tb, bool accum) {
    for (size_t i = 0; i < m; ++i)
        for (size_t j = 0; j < n; ++j) {
            fp_t s = accum ? c[i * n + j] : 0;
            for (size_t p = 0; p < k; ++p) {
                fp_t av = ta ? a[p * m + i] : a[i * k + p];
                fp_t bv = tb ? b[j * k + p] : b[p * n + j];
                s += av * bv;
            }
            c[i * n + j] = s;
        }
}

//! LU with partial pivoting of an n x n row-major matrix held in \p a;
//! returns the sign of the permutation, 0 when a pivot vanishes exactly
int lu_inplace(fp_t* a, int* piv, size_t n) {
    int sign = 1;
    for (size_t c = 0; c < n; ++c) {
        size_t best = c;
        for (size_t r = c + 1; r < n; ++r)
            if (std::fabs(a[r * n + c]) > std::fabs(a[best * n + c]))
                best = r;
        piv[c] = static_cast<int>(best);
        if (best != c) {
            for (size_t j = 0; j < n; ++j)
                std::swap(a[c * n + j], a[best * n + j]);
            sign = -sign;
        }
        fp_t p = a[c * n + c];
        if (p == 0)
            return 0;
        for (size_t r = c + 1; r < n; ++r) {
            fp_t f = a[r * n + c] / p;
            a[r * n + c] = f;
            for (size_t j = c + 1; j < n; ++j)
                a[r * n + j] -= f * a[c * n + j];
        }
    }
    return sign;
}

fp_t det3(const fp_t* m) {
    return m[0] * (m[4] * m[8] - m[5] * m[7]) -
           m[1] * (m[3] * m[8] - m[5] * m[6]) +
           m[2] * (m[3] * m[7] - m[4] * m[6]);
}

//! cofactor matrix (cof[i][j] = (-1)^(i+j) minor_ij) of a 3x3
void cof3(const fp_t* m, fp_t* c) {
    c[0] = m[4] * m[8] - m[5] * m[7];
    c[1] = m[5] * m[6] - m[3] * m[8];
    c[2] = m[3] * m[7] - m[4] * m[6];
    c[3] = m[2] * m[7] - m[1] * m[8];
    c[4] = m[0] * m[8] - m[2] * m[6];
    c[5] = m[1] * m[6] - m[0] * m[7];
    c[6] = m[1] * m[5] - m[2] * m[4];
    c[7] = m[2] * m[3] - m[0] * m[5];
    c[8] = m[0] * m[4] - m[1] * m[3];
}

fp_t det_general(const fp_t* m, size_t n, std::vector<fp_t>& tmp,
                 std::vector<int>& piv) {
    if (n == 1)
        return m[0];
    if (n == 2)
        return m[0] * m[3] - m[1] * m[2];
    if (n == 3)
        return det3(m);
    tmp.assign(m, m + n * n);
    piv.resize(n);
    int sign = lu_inplace(tmp.data(), piv.data(), n);
    fp_t d = sign;
    for (size_t i = 0; i < n && sign; ++i)
        d *= tmp[i * n + i];
    return d;
}

void inverse_general(const fp_t* m, fp_t* out, size_t n, std::vector<fp_t>& lu,
                     std::vector<int>& piv) {
    if (n == 3) {
        fp_t c[9];
        cof3(m, c);
        fp_t d = m[0] * c[0] + m[1] * c[1] + m[2] * c[2];
        fp_t id = 1 / d;
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j)
                out[i * 3 + j] = c[j * 3 + i] * id;
        return;
    }
    if (n == 2) {
        fp_t id = 1 / (m[0] * m[3] - m[1] * m[2]);
        out[0] = m[3] * id;
        out[1] = -m[1] * id;
        out[2] = -m[2] * id;
        out[3] = m[0] * id;
        return;
    }
    if (n == 1) {
        out[0] = 1 / m[0];
        return;
    }
    lu.assign(m, m + n * n);
    piv.resize(n);
    int sign = lu_inplace(lu.data(), piv.data(), n);
    if (!sign)
        throw SANMNumericalError{"batched_matinv: singular matrix"};
    // solve A X = I column by column (apply P, L, then U)
    std::vector<fp_t> col(n);
    for (size_t j = 0; j < n; ++j) {
        for (size_t i = 0; i < n; ++i)
            col[i] = (i == j);
        for (size_t i = 0; i < n; ++i)
            std::swap(col[i], col[piv[i]]);
        for (size_t i = 0; i < n; ++i)
            for (size_t p = 0; p < i; ++p)
                col[i] -= lu[i * n + p] * col[p];
        for (size_t ii = n; ii-- > 0;) {
            for (size_t p = ii + 1; p < n; ++p)
                col[ii] -= lu[ii * n + p] * col[p];
            col[ii] /= lu[ii * n + ii];
        }
        for (size_t i = 0; i < n; ++i)
            out[i * n + j] = col[i];
    }
}

}  // namespace

TensorND& TensorND::as_mm(const TensorND& lhs, const TensorND& rhs, bool accum,
                          bool tl, bool tr) {
    sanm_assert(this != &lhs && this != &rhs, "inplace not allowed");
    sanm_assert(lhs.rank() == 2 && rhs.rank() == 2);
    size_t m = lhs.shape(tl ? 1 : 0), k = lhs.shape(tl ? 0 : 1),
           k2 = rhs.shape(tr ? 1 : 0), n = rhs.shape(tr ? 0 : 1);
    sanm_assert(k == k2, "matmul shape mismatch: %s vs %s",
                lhs.shape().str().c_str(), rhs.shape().str().c_str());
    TensorShape oshp{m, n};
    if (accum && !empty()) {
        sanm_assert(m_shape == oshp);
        gemm(rwptr(), lhs.ptr(), rhs.ptr(), m, k, n, tl, tr, true);
    } else {
        gemm(set_shape(oshp).woptr(), lhs.ptr(), rhs.ptr(), m, k, n, tl, tr,
             false);
    }
    return *this;
}

TensorND& TensorND::as_batched_mm(const TensorND& lhs, const TensorND& rhs,
                                  bool accum, bool tl, bool tr) {
    sanm_assert(this != &lhs && this != &rhs, "inplace not allowed");
    sanm_assert(lhs.rank() == 3 && rhs.rank() == 3 &&
                        lhs.shape(0) == rhs.shape(0),
                "batched matmul shape mismatch: %s vs %s",
                lhs.shape().str().c_str(), rhs.shape().str().c_str());
    size_t nb = lhs.shape(0), m = lhs.shape(tl ? 2 : 1),
           k = lhs.shape(tl ? 1 : 2), k2 = rhs.shape(tr ? 2 : 1),
           n = rhs.shape(tr ? 1 : 2);
    sanm_assert(k == k2, "batched matmul shape mismatch: %s vs %s",
                lhs.shape().str().c_str(), rhs.shape().str().c_str());
    TensorShape oshp{nb, m, n};
    bool acc = accum && !empty();
    if (acc)
        sanm_assert(m_shape == oshp);
    if (lhs.is_zero() || rhs.is_zero()) {
        if (!acc)
            set_shape(oshp).fill_with_inplace(0);
        return *this;
    }
    fp_t* d = acc ? rwptr() : set_shape(oshp).woptr();
    const fp_t *a = lhs.ptr(), *b = rhs.ptr();
    for (size_t i = 0; i < nb; ++i)
        gemm(d + i * m * n, a + i * m * k, b + i * k * n, m, k, n, tl, tr, acc);
    return *this;
}

TensorND& TensorND::as_batched_mm_vecitem_left(const TensorND& lhs,
                                               const TensorND& rhs,
                                               bool accum) {
    // out[b,(m,n),p] = sum_k l[b,(m,k),p] r[b,k,n]
    sanm_assert(this != &lhs && this != &rhs, "inplace not allowed");
    sanm_assert(lhs.rank() == 3 && rhs.rank() == 3 &&
                lhs.shape(0) == rhs.shape(0));
    size_t nb = lhs.shape(0), k = rhs.shape(1), n = rhs.shape(2),
           p = lhs.shape(2);
    sanm_assert(lhs.shape(1) % k == 0);
    size_t m = lhs.shape(1) / k;
    TensorShape oshp{nb, m * n, p};
    bool acc = accum && !empty();
    if (acc)
        sanm_assert(m_shape == oshp);
    fp_t* d = acc ? rwptr() : set_shape(oshp).woptr();
    const fp_t *a = lhs.ptr(), *b = rhs.ptr();
    for (size_t ib = 0; ib < nb; ++ib)
        for (size_t im = 0; im < m; ++im)
            for (size_t in = 0; in < n; ++in)
                for (size_t ip = 0; ip < p; ++ip) {
                    fp_t s = acc ? d[((ib * m + im) * n + in) * p + ip] : 0;
                    for (size_t ik = 0; ik < k; ++ik)
                        s += a[((ib * m + im) * k + ik) * p + ip] *
                             b[(ib * k + ik) * n + in];
                    d[((ib * m + im) * n + in) * p + ip] = s;
                }
    return *this;
}

TensorND& TensorND::as_batched_transpose(const TensorND& src) {
    sanm_assert(this != &src, "inplace not allowed");
    sanm_assert(src.rank() == 3);
    size_t nb = src.shape(0), m = src.shape(1), n = src.shape(2);
    set_shape({nb, n, m});
    if (src.is_zero())
        return fill_with_inplace(0);
    const fp_t* s = src.ptr();
    fp_t* d = woptr();
    for (size_t b = 0; b < nb; ++b)
        for (size_t i = 0; i < m; ++i)
            for (size_t j = 0; j < n; ++j)
                d[(b * n + j) * m + i] = s[(b * m + i) * n + j];
    return *this;
}

TensorND& TensorND::as_transpose(const TensorND& src) {
    sanm_assert(this != &src, "inplace not allowed");
    sanm_assert(src.rank() == 2);
    size_t m = src.shape(0), n = src.shape(1);
    const fp_t* s = src.ptr();
    fp_t* d = set_shape({n, m}).woptr();
    for (size_t i = 0; i < m; ++i)
        for (size_t j = 0; j < n; ++j)
            d[j * m + i] = s[i * n + j];
    return *this;
}

TensorND& TensorND::as_batched_matinv(const TensorND& src) {
    sanm_assert(this != &src, "inplace not allowed");
    sanm_assert(src.rank() == 3 && src.shape(1) == src.shape(2),
                "batched_matinv needs square matrices, got %s",
                src.shape().str().c_str());
    size_t nb = src.shape(0), n = src.shape(1);
    const fp_t* s = src.ptr();
    fp_t* d = set_shape(src.shape()).woptr();
    std::vector<fp_t> lu;
    std::vector<int> piv;
    for (size_t b = 0; b < nb; ++b)
        inverse_general(s + b * n * n, d + b * n * n, n, lu, piv);
    return *this;
}

TensorND& TensorND::as_batched_determinant(const TensorND& src) {
    sanm_assert(this != &src, "inplace not allowed");
    sanm_assert(src.rank() == 3 && src.shape(1) == src.shape(2));
    size_t nb = src.shape(0), n = src.shape(1);
    const fp_t* s = src.ptr();
    fp_t* d = set_shape({nb, 1}).woptr();
    std::vector<fp_t> tmp;
    std::vector<int> piv;
    for (size_t b = 0; b < nb; ++b)
        d[b] = det_general(s + b * n * n, n, tmp, piv);
    return *this;
}

TensorND& TensorND::as_batched_cofactor(const TensorND& src) {
    sanm_assert(this != &src, "inplace not allowed");
    sanm_assert(src.rank() == 3 && src.shape(1) == src.shape(2));
    size_t nb = src.shape(0), n = src.shape(1);
    const fp_t* s = src.ptr();
    fp_t* d = set_shape(src.shape()).woptr();
    std::vector<fp_t> minor((n - 1) * (n - 1)), tmp;
    std::vector<int> piv;
    for (size_t b = 0; b < nb; ++b) {
        const fp_t* m = s + b * n * n;
        fp_t* c = d + b * n * n;
        if (n == 3) {
            cof3(m, c);
            continue;
        }
        if (n == 1) {
            c[0] = 1;
            continue;
        }
        for (size_t i = 0; i < n; ++i)
            for (size_t j = 0; j < n; ++j) {
                size_t q = 0;
                for (size_t r = 0; r < n; ++r)
                    for (size_t cc = 0; cc < n; ++cc)
                        if (r != i && cc != j)
                            minor[q++] = m[r * n + cc];
                fp_t v = det_general(minor.data(), n - 1, tmp, piv);
                c[i * n + j] = ((i + j) & 1) ? -v : v;
            }
    }
    return *this;
}

TensorND& TensorND::as_diag(const TensorND& src) {
    sanm_assert(this != &src && src.rank() == 1);
    size_t n = src.shape(0);
    const fp_t* s = src.ptr();
    fp_t* d = set_shape({n, n}).woptr();
    std::fill_n(d, n * n, 0.);
    for (size_t i = 0; i < n; ++i)
        d[i * n + i] = s[i];
    return *this;
}

TensorND& TensorND::as_batched_diag(const TensorND& src) {
    sanm_assert(this != &src && src.rank() == 2);
    size_t nb = src.shape(0), n = src.shape(1);
    const fp_t* s = src.ptr();
    fp_t* d = set_shape({nb, n, n}).woptr();
    std::fill_n(d, nb * n * n, 0.);
    for (size_t b = 0; b < nb; ++b)
        for (size_t i = 0; i < n; ++i)
            d[(b * n + i) * n + i] = s[b * n + i];
    return *this;
}

TensorND& TensorND::as_batched_diag_scalar(const TensorND& src, size_t dim) {
    sanm_assert(this != &src && src.shape().is_batched_scalar());
    size_t nb = src.shape(0);
    const fp_t* s = src.ptr();
    fp_t* d = set_shape({nb, dim, dim}).woptr();
    std::fill_n(d, nb * dim * dim, 0.);
    for (size_t b = 0; b < nb; ++b)
        for (size_t i = 0; i < dim; ++i)
            d[(b * dim + i) * dim + i] = s[b];
    return *this;
}

TensorND& TensorND::as_batched_trace(const TensorND& src) {
    sanm_assert(this != &src && src.rank() == 3 &&
                src.shape(1) == src.shape(2));
    size_t nb = src.shape(0), n = src.shape(1);
    const fp_t* s = src.ptr();
    fp_t* d = set_shape({nb, 1}).woptr();
    for (size_t b = 0; b < nb; ++b) {
        fp_t t = 0;
        for (size_t i = 0; i < n; ++i)
            t += s[(b * n + i) * n + i];
        d[b] = t;
    }
    return *this;
}

// SVD-W by one-sided Jacobi: A V = U S, then W = U V^T and M = U S U^T W.
const TensorND& TensorND::compute_batched_svd_w(TensorND& u, TensorND& s,
                                                TensorND& w,
                                                bool require_rotation) const {
    sanm_assert(rank() == 3 && shape(1) == shape(2));
    size_t nb = shape(0), n = shape(1);
    const fp_t* src = ptr();
    fp_t* pu = u.set_shape({nb, n, n}).woptr();
    fp_t* ps = s.set_shape({nb, n}).woptr();
    fp_t* pw = w.set_shape({nb, n, n}).woptr();
    std::vector<fp_t> a(n * n), v(n * n), tmp;
    std::vector<size_t> ord(n);
    std::vector<int> piv;
    for (size_t b = 0; b < nb; ++b) {
        std::copy_n(src + b * n * n, n * n, a.begin());
        std::fill(v.begin(), v.end(), 0.);
        for (size_t i = 0; i < n; ++i)
            v[i * n + i] = 1;
        for (int sweep = 0; sweep < 60; ++sweep) {
            fp_t off = 0;
            for (size_t p = 0; p + 1 < n; ++p)
                for (size_t q = p + 1; q < n; ++q) {
                    fp_t app = 0, aqq = 0, apq = 0;
                    for (size_t i = 0; i < n; ++i) {
                        app += a[i * n + p] * a[i * n + p];
                        aqq += a[i * n + q] * a[i * n + q];
                        apq += a[i * n + p] * a[i * n + q];
                    }
                    if (std::fabs(apq) <=
                        1e-300 + 1e-17 * std::sqrt(app * aqq))
                        continue;
                    off = std::max(off, std::fabs(apq) / std::sqrt(app * aqq));
                    fp_t zeta = (aqq - app) / (2 * apq);
                    fp_t t = (zeta >= 0 ? 1. : -1.) /
                             (std::fabs(zeta) + std::sqrt(1 + zeta * zeta));
                    fp_t c = 1 / std::sqrt(1 + t * t), sn = c * t;
                    for (size_t i = 0; i < n; ++i) {
                        fp_t x = a[i * n + p], y = a[i * n + q];
                        a[i * n + p] = c * x - sn * y;
                        a[i * n + q] = sn * x + c * y;
                        x = v[i * n + p];
                        y = v[i * n + q];
                        v[i * n + p] = c * x - sn * y;
                        v[i * n + q] = sn * x + c * y;
                    }
                }
            if (off < 1e-15)
                break;
        }
        // singular values = column norms; sort descending
        std::vector<fp_t> sv(n);
        for (size_t j = 0; j < n; ++j) {
            fp_t t = 0;
            for (size_t i = 0; i < n; ++i)
                t += a[i * n + j] * a[i * n + j];
            sv[j] = std::sqrt(t);
        }
        std::iota(ord.begin(), ord.end(), size_t(0));
        std::stable_sort(ord.begin(), ord.end(),
                         [&](size_t x, size_t y) { return sv[x] > sv[y]; });
        fp_t* U = pu + b * n * n;
        fp_t* W = pw + b * n * n;
        std::vector<fp_t> Vs(n * n);
        for (size_t jj = 0; jj < n; ++jj) {
            size_t j = ord[jj];
            ps[b * n + jj] = sv[j];
            for (size_t i = 0; i < n; ++i) {
                U[i * n + jj] = sv[j] > 0 ? a[i * n + j] / sv[j] : (i == jj);
                Vs[i * n + jj] = v[i * n + j];
            }
        }
        // (u_j, v_j) is only defined up to a common sign, and the Jacobi sweeps
        // can land on either one for nearby inputs (the rotation angle jumps
        // where two column norms cross).  Fix it by the largest component of
        // u_j, so that U is a continuous function of M (tests/tensor.cpp takes
        // finite differences of it); W = U V^T does not depend on the choice.
        for (size_t j = 0; j < n; ++j) {
            size_t im = 0;
            for (size_t i = 1; i < n; ++i)
                if (std::fabs(U[i * n + j]) > std::fabs(U[im * n + j]))
                    im = i;
            if (U[im * n + j] < 0) {
                for (size_t i = 0; i < n; ++i) {
                    U[i * n + j] = -U[i * n + j];
                    Vs[i * n + j] = -Vs[i * n + j];
                }
            }
        }
        auto make_w = [&]() { gemm(W, U, Vs.data(), n, n, n, false, true, false); };
        make_w();
        if (require_rotation && det_general(W, n, tmp, piv) < 0) {
            // negate singular values (and their U columns) so that det W = 1,
            // chosen as libsanm/tensor_svd.cpp:93-121 chooses them: the first
            // group of (nearly) equal values with the fewest members -- a
            // single value, or all of an odd-sized group
            constexpr fp_t EPS = 1e-3;
            fp_t* S = ps + b * n;
            int best_idx = -1, best_nr = int(n) + 1;
            for (size_t i = 0; i < n; ++i) {
                size_t j = i + 1;
                while (j < n && std::fabs(S[i] - S[j]) < EPS)
                    ++j;
                int nr = int(j - i);
                if (nr <= best_nr || (nr == best_nr + 1 && nr % 2 == 1)) {
                    best_idx = int(i);
                    best_nr = nr;
                    if (nr == 1)
                        break;
                }
                i = j;  // and the loop increments once more, as the reference's does
            }
            int lo = best_idx, hi = best_idx + 1;
            if (!(best_nr == 1 || best_nr % 2 == 0))
                hi = best_idx + best_nr;
            for (int j = lo; j < hi; ++j) {
                S[j] = -S[j];
                for (size_t i = 0; i < n; ++i)
                    U[i * n + j] = -U[i * n + j];
            }
            make_w();
        }
    }
    return *this;
}

// coefficient of a^order in det(sum_i a^i A_i): Leibniz expansion with
// truncated polynomial products (host utility for small matrices)
TensorND compute_polymat_det_coeff(const TensorArray& a, size_t order) {
    sanm_assert(!a.empty() && a[0].rank() == 3 && a[0].shape(1) == a[0].shape(2));
    size_t nb = a[0].shape(0), n = a[0].shape(1), nt = std::min(a.size(), order + 1);
    sanm_assert(n <= 8, "polymat det on the host layer: dim %zu too large", n);
    TensorND out{TensorShape{nb, 1}};
    fp_t* o = out.woptr();
    std::vector<size_t> perm(n);
    std::vector<fp_t> acc(order + 1), nxt(order + 1);
    for (size_t b = 0; b < nb; ++b) {
        std::iota(perm.begin(), perm.end(), size_t(0));
        fp_t total = 0;
        do {
            int inv = 0;
            for (size_t i = 0; i < n; ++i)
                for (size_t j = i + 1; j < n; ++j)
                    inv += perm[i] > perm[j];
            std::fill(acc.begin(), acc.end(), 0.);
            acc[0] = 1;
            for (size_t r = 0; r < n; ++r) {
                std::fill(nxt.begin(), nxt.end(), 0.);
                for (size_t t = 0; t < nt; ++t) {
                    if (a[t].is_zero())
                        continue;
                    fp_t e = a[t].ptr()[(b * n + r) * n + perm[r]];
                    for (size_t d = 0; d + t <= order; ++d)
                        nxt[d + t] += acc[d] * e;
                }
                acc.swap(nxt);
            }
            total += (inv & 1) ? -acc[order] : acc[order];
        } while (std::next_permutation(perm.begin(), perm.end()));
        o[b] = total;
    }
    return out;
}

// ------------------------------------------------------ StSparseLinearTrans
StSparseLinearTrans& StSparseLinearTrans::reset(Type type, bool is_batched,
                                                const TensorND& coeff) {
    size_t want_rank, ax_out = is_batched ? 1 : 0, ax_in;
    if (type == ELEMWISE) {
        want_rank = is_batched ? 2 : 1;
        ax_in = ax_out;
    } else if (type == FULL) {
        want_rank = is_batched ? 3 : 2;
        ax_in = ax_out + 1;
    } else {
        throw SANMError{ssprintf("invalid StSparseLinearTrans type: %d",
                                 static_cast<int>(type))};
    }
    sanm_assert(coeff.rank() == want_rank, "expected rank=%zu, shape=%s",
                want_rank, coeff.shape().str().c_str());
    m_type = type;
    m_batched = is_batched;
    m_batch = is_batched ? coeff.shape(0) : 0;
    m_odim = coeff.shape(ax_out);
    m_idim = coeff.shape(ax_in);
    m_coeff = coeff;
    return *this;
}

StSparseLinearTrans StSparseLinearTrans::compose_with_elemwise(
        const TensorND& rhs) const {
    sanm_assert(valid());
    TensorND f;
    if (m_batched) {
        f = rhs.flatten_batched();
        sanm_assert(f.shape(0) == m_batch && f.shape(1) == m_idim,
                    "shape mismatch: batch=%zu idim=%zu rhs=%s", m_batch,
                    m_idim, rhs.shape().str().c_str());
        if (m_type == FULL)
            f.reshape_inplace({m_batch, 1, m_idim});
    } else {
        f = rhs.flatten_as_vec();
        sanm_assert(f.shape(0) == m_idim, "shape mismatch: idim=%zu rhs=%s",
                    m_idim, rhs.shape().str().c_str());
        if (m_type == FULL)
            f.reshape_inplace({1, m_idim});
    }
    return {m_type, m_batched, m_coeff * f};
}

StSparseLinearTrans StSparseLinearTrans::compose_with_full(
        const TensorND& rhs) const {
    sanm_assert(valid());
    TensorND l = as_full(), out;
    sanm_assert(rhs.rank() == l.rank(), "shape mismatch: %s",
                rhs.shape().str().c_str());
    if (m_batched)
        out.as_batched_mm(l, rhs);
    else
        out.as_mm(l, rhs);
    return {FULL, m_batched, out};
}

StSparseLinearTrans StSparseLinearTrans::compose_with_scaling(
        fp_t scale) const {
    sanm_assert(valid());
    return {m_type, m_batched, m_coeff * scale};
}

StSparseLinearTrans& StSparseLinearTrans::operator+=(
        const StSparseLinearTrans& rhs) {
    sanm_assert(rhs.valid());
    if (!valid())
        return *this = rhs;
    sanm_assert(m_batched == rhs.m_batched,
                "batch setting mismatches in add-assign: %d vs %d", m_batched,
                rhs.m_batched);
    sanm_assert(m_batch == rhs.m_batch && m_idim == rhs.m_idim &&
                m_odim == rhs.m_odim);
    if (m_type == rhs.m_type) {
        m_coeff += rhs.m_coeff;
    } else {
        m_coeff = as_full();
        m_coeff += rhs.as_full();
        m_type = FULL;
    }
    return *this;
}

TensorND StSparseLinearTrans::as_full() const {
    sanm_assert(valid());
    if (m_type == FULL)
        return m_coeff;
    return m_batched ? TensorND{}.as_batched_diag(m_coeff)
                     : TensorND{}.as_diag(m_coeff);
}

TensorND StSparseLinearTrans::apply(const TensorND& x) const {
    sanm_assert(valid());
    if (m_batched) {
        TensorND f = x.flatten_batched();
        sanm_assert(f.shape(0) == m_batch && f.shape(1) == m_idim,
                    "shape mismatch: batch=%zu inp_dim=%zu x=%s", m_batch,
                    m_idim, x.shape().str().c_str());
        if (m_type == FULL) {
            f.reshape_inplace({m_batch, m_idim, 1});
            return TensorND{}.as_batched_mm(m_coeff, f).reshape(
                    {m_batch, m_odim});
        }
        return m_coeff * f;
    }
    TensorND f = x.flatten_as_vec();
    sanm_assert(f.shape(0) == m_idim, "shape mismatch: inp_dim=%zu x=%s",
                m_idim, x.shape().str().c_str());
    if (m_type == FULL) {
        f.reshape_inplace({m_idim, 1});
        return TensorND{}.as_mm(m_coeff, f).reshape({m_odim});
    }
    return m_coeff * f;
}

}  // namespace sanm
