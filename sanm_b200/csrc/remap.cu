/* sanm_b200/csrc/remap.cu -- see remap.h. */
#include "remap.h"

#include <mutex>

#include "capi_util.h"

namespace sanm_b200 {

namespace {
struct ApplyKernel {
    const uint32_t* rowptr;
    const uint32_t* idx;
    const double* coeff;
    const double* x;
    double* y;
    VecLayout xl, yl;
    uint64_t nout;
    int ye;
    SANM_D void operator()(size_t i) const {
        // enumerate outputs in their storage order so that stores coalesce
        size_t row = i;
        if (ye) {
            size_t n = nout / ye;
            size_t c = i / n, t = i % n;
            row = t * ye + c;
        }
        double s = 0;
        for (uint32_t p = rowptr[row], pe = rowptr[row + 1]; p < pe; ++p) {
            s += x[xl.addr(idx[p])] * coeff[p];
        }
        y[yl.addr(row)] = s;
    }
};
}  // namespace

Remap::Remap(const Shape& out_shape, const Shape& in_shape,
             const uint64_t* rowptr, const uint64_t* idx, const double* coeff)
        : m_out{out_shape}, m_in{in_shape} {
    const uint64_t n = out_shape.total(), nin = in_shape.total();
    sanm_check(n > 0 && nin > 0);
    const uint64_t nnz = rowptr[n];
    sanm_check(nnz < (1ull << 32) && nin < (1ull << 32),
               "remap too large for 32-bit indices");
    m_h_rowptr.resize(n + 1);
    for (uint64_t i = 0; i <= n; ++i) {
        sanm_check(i == 0 || rowptr[i] >= rowptr[i - 1], "rowptr not monotone");
        m_h_rowptr[i] = (uint32_t)rowptr[i];
    }
    m_h_idx.resize(nnz);
    for (uint64_t i = 0; i < nnz; ++i) {
        sanm_check(idx[i] < nin, "remap index %zu out of range %zu",
                   (size_t)idx[i], (size_t)nin);
        m_h_idx[i] = (uint32_t)idx[i];
    }
    m_h_coeff.assign(coeff, coeff + nnz);
    m_hash_struct = hash_vec(m_h_idx, hash_vec(m_h_rowptr, n * 31 + nin));
    m_hash_full = hash_vec(m_h_coeff, m_hash_struct);
    m_rowptr.upload(m_h_rowptr);
    m_idx.upload(m_h_idx);
    m_coeff.upload(m_h_coeff);
}

std::shared_ptr<Remap> Remap::make_identity(const Shape& in_shape,
                                            const Shape& out_shape) {
    // libsanm/anm.cpp:19-49,77-81
    const uint64_t n = out_shape.total();
    sanm_check(n <= in_shape.total());
    std::vector<uint64_t> rowptr(n + 1), idx(n);
    std::vector<double> coeff(n, 1.0);
    for (uint64_t i = 0; i < n; ++i) {
        rowptr[i] = i;
        idx[i] = i;
    }
    rowptr[n] = n;
    return std::make_shared<Remap>(out_shape, in_shape, rowptr.data(),
                                   idx.data(), coeff.data());
}

void Remap::apply(double* y, VecLayout ylay, const double* x,
                  VecLayout xlay) const {
    rt::for_each(nr_out(), ApplyKernel{m_rowptr.ptr(), m_idx.ptr(),
                                       m_coeff.ptr(), x, y, xlay, ylay,
                                       nr_out(), ylay.e});
}

/* ---- allocation-free apply (sanm_remap_apply_static) -----------------------
 * sanm_remap_apply stages the caller's host vectors in two device buffers that
 * it takes from the stream-ordered pool on every call.  Callers that must not
 * allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc) get the same gather with that scratch in a file-scope
 * device array instead: SANM_REMAP_STATIC_MAX_DOUBLES doubles (16 MiB), held by
 * every GPU that loads the library until exit.  Layout: x at offset 0, y at the
 * next 256-byte boundary after x.
 *
 * g_apply_scratch is a DEVICE symbol: in host code its name is only a host-side
 * stand-in, never a device address.  The address the kernel gets comes from
 * cudaGetSymbolAddress, asked once per device with that device current.
 * All calls share the array and it keeps the last call's contents; the y part
 * is cleared before every apply.  Calls that overlap on the GPU (several
 * streams or host threads) race on it: one call at a time per device. */
namespace {
constexpr uint64_t STATIC_DOUBLES = SANM_REMAP_STATIC_MAX_DOUBLES;
#ifdef SANM_HOSTEMU
alignas(256) double g_apply_scratch[STATIC_DOUBLES];
double* apply_scratch() { return g_apply_scratch; }
#else
static __device__ __align__(256) double g_apply_scratch[STATIC_DOUBLES];

double* apply_scratch() {
    rt::check_device();
    constexpr int MAX_DEV = 64;
    static double* ptr[MAX_DEV] = {};
    static std::mutex mu;
    int dev = -1;
    SANM_CUDA_CHECK(cudaGetDevice(&dev));
    sanm_check(dev >= 0 && dev < MAX_DEV, "device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lk{mu};
    if (!ptr[dev]) {
        double* p = nullptr;
        SANM_CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_apply_scratch));
        sanm_check(p != nullptr, "cudaGetSymbolAddress returned a null address");
        ptr[dev] = p;
    }
    return ptr[dev];
}
#endif
}  // namespace

void Remap::apply_static(double* y_host, const double* x_host) const {
    const uint64_t nin = nr_in(), nout = nr_out();
    const uint64_t yoff = (nin + 31) / 32 * 32;  // 256-byte boundary
    if (yoff + nout > STATIC_DOUBLES) {
        sanm_throw(SANM_ERR_ASSERT,
                   "sanm_remap_apply_static: %zu input + %zu output doubles exceed "
                   "the static scratch of %zu doubles "
                   "(SANM_REMAP_STATIC_MAX_DOUBLES); use sanm_remap_apply",
                   (size_t)nin, (size_t)nout, (size_t)STATIC_DOUBLES);
    }
    double* sx = apply_scratch();
    double* sy = sx + yoff;
    // the array keeps the previous call's contents: rows without entries and
    // every other output start from zero
    rt::dzero(sy, nout * sizeof(double));
    rt::h2d(sx, x_host, nin * sizeof(double));
    apply(sy, VecLayout{}, sx, VecLayout{});
    rt::d2h(y_host, sy, nout * sizeof(double));
}

}  // namespace sanm_b200

extern "C" int sanm_remap_apply_static(sanm_remap_t* r, const double* x, double* y) {
    using namespace sanm_b200;
    return capi::guard([&] {
        sanm_check(r && x && y);
        r->r->apply_static(y, x);
    });
}
