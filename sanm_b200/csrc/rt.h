/* sanm_b200/csrc/rt.h -- thin runtime layer under every kernel in this library.
 *
 * Product build (nvcc, sm_100a): device memory, one CUDA stream, kernels.
 * SANM_HOSTEMU build (g++ -x c++): the SAME kernel bodies run as plain loops
 * on the host.  That build exists only so that the CPU-only development
 * container can check kernel *logic* against the oracle (tests/, `-m "not gpu"`);
 * it is compiled by tests/hostemu_build.py into tests/_hostemu/ and is never
 * loaded by the sanm_b200 package -- the product library has no CPU fallback
 * and every entry point fails with SANM_ERR_CUDA when no device is present.
 *
 * Kernel-writing discipline that makes the dual build possible:
 *   - per-item kernels are functors run through rt::for_each (one thread per
 *     item, item index fastest-varying in every array => coalesced);
 *   - block-cooperative kernels are functors run through rt::for_each_block;
 *     inside, work is written as `for (i = tid; i < n; i += nthr)` phases
 *     separated by ctx.sync(); threads communicate only through shared memory
 *     across a sync().  In the host build nthr == 1 and sync() is a no-op.
 */
#pragma once

#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include "../../include/sanm_b200.h"

#ifdef SANM_HOSTEMU
#define SANM_HD
#define SANM_D
#include <cmath>
#else
#include <cuda_runtime.h>
#include <cooperative_groups.h>
#define SANM_HD __host__ __device__
#define SANM_D __device__
#endif

//! compiler-only barrier: loads written before it are issued before any
//! arithmetic written after it (memory-level parallelism by construction)
#define SANM_LOADS_ISSUED() asm volatile("" ::: "memory")
//! read-only global load (ld.global.nc): for data no kernel writes while it
//! runs.  Besides the cache policy this tells ptxas the access is GLOBAL:
//! pointers fetched from descriptors are generic, and generic loads are kept
//! in program order next to their uses instead of being issued together.
#ifdef __CUDA_ARCH__
#define SANM_LDG(p) __ldg(p)
#else
#define SANM_LDG(p) (*(p))
#endif
//! the same for loads the compiler may treat as invariant (const __restrict__):
//! an empty instruction that consumes and redefines 8 loaded doubles, so all
//! 8 loads are issued before the first use
#ifdef __CUDA_ARCH__
#define SANM_TIE8(a)                                                         \
    asm volatile("" : "+d"(a[0]), "+d"(a[1]), "+d"(a[2]), "+d"(a[3]),         \
                      "+d"(a[4]), "+d"(a[5]), "+d"(a[6]), "+d"(a[7]))
#else
#define SANM_TIE8(a) (void)0
#endif

namespace sanm_b200 {

// status codes (SANM_OK, SANM_ERR_*) come from the public header

struct Error : public std::runtime_error {
    int code;
    Error(int c, const std::string& msg) : std::runtime_error{msg}, code{c} {}
};

std::string ssprintf(const char* fmt, ...) __attribute__((format(printf, 1, 2)));

#define sanm_throw(code, ...) \
    throw ::sanm_b200::Error { code, ::sanm_b200::ssprintf(__VA_ARGS__) }

#define sanm_check(expr, ...)                                               \
    do {                                                                    \
        if (!(expr)) {                                                      \
            throw ::sanm_b200::Error{                                       \
                    SANM_ERR_ASSERT,                           \
                    std::string{"assertion `" #expr "' failed at "} +       \
                            __FILE__ + ":" + std::to_string(__LINE__) +     \
                            ": " + ::sanm_b200::ssprintf("" __VA_ARGS__)};  \
        }                                                                   \
    } while (0)

SANM_HD inline void rt_atomic_inc(int* p) {
#if defined(__CUDA_ARCH__)
    atomicAdd(p, 1);
#else
    ++*p;
#endif
}

//! 64-bit content hash (splitmix-style mixing of 8-byte words); keys of the
//! topology caches (symbolic analysis, assembly lists) across solver instances
inline uint64_t hash_bytes(const void* data, size_t bytes,
                           uint64_t seed = 0x9E3779B97F4A7C15ull) {
    const unsigned char* p = static_cast<const unsigned char*>(data);
    uint64_t h = seed ^ (bytes * 0xBF58476D1CE4E5B9ull);
    size_t i = 0;
    for (; i + 8 <= bytes; i += 8) {
        uint64_t w;
        memcpy(&w, p + i, 8);
        h ^= w;
        h *= 0xBF58476D1CE4E5B9ull;
        h ^= h >> 29;
    }
    uint64_t w = 0;
    if (i < bytes) memcpy(&w, p + i, bytes - i);
    h ^= w;
    h *= 0x94D049BB133111EBull;
    h ^= h >> 31;
    return h;
}
template <typename T>
inline uint64_t hash_vec(const std::vector<T>& v, uint64_t seed) {
    return hash_bytes(v.data(), v.size() * sizeof(T), seed);
}

namespace rt {

#ifdef SANM_HOSTEMU
using stream_t = int;
inline void check_device() {}
#define SANM_CUDA_CHECK(x) (void)(x)
#else
using stream_t = cudaStream_t;
#define SANM_CUDA_CHECK(x)                                                  \
    do {                                                                    \
        cudaError_t err__ = (x);                                            \
        if (err__ != cudaSuccess) {                                         \
            throw ::sanm_b200::Error{                                       \
                    SANM_ERR_CUDA,                             \
                    std::string{"CUDA error at "} + __FILE__ + ":" +        \
                            std::to_string(__LINE__) + ": " +               \
                            cudaGetErrorString(err__)};                     \
        }                                                                   \
    } while (0)
void check_device();
#endif

//! the single stream every kernel of this library is launched on
stream_t stream();
//! run on the caller's stream (nullptr: back to the library's own)
void set_stream(void* cuda_stream);
void sync();

void* dmalloc(size_t bytes);
void dfree(void* p);
void h2d(void* dst, const void* src, size_t bytes);
void d2h(void* dst, const void* src, size_t bytes);
void d2d(void* dst, const void* src, size_t bytes);
void dzero(void* dst, size_t bytes);
/* Copy pipes: two extra streams (0: host->device, 1: device->host) so that
 * transfers of one order overlap the element kernels of another.  Ordering is
 * expressed with fences: a fence taken on one stream can be awaited by the
 * other without blocking the host. */
constexpr int PIPE_H2D = 0, PIPE_D2H = 1;
//! third side stream: exchange steps (comm.cu) that run beside the element
//! kernels of the next order
constexpr int PIPE_COMM = 2;
//! side stream of the factorisation's look-ahead (sparse_lu.cu)
constexpr int PIPE_AUX = 3;
//! while alive, stream() is the given pipe: kernels and collectives launched
//! in the scope go to the side stream (the caller orders it with fences)
struct ScopedPipe {
    explicit ScopedPipe(int pipe);
    ~ScopedPipe();
    ScopedPipe(const ScopedPipe&) = delete;
    ScopedPipe& operator=(const ScopedPipe&) = delete;
};
void pipe_h2d(void* dst, const void* src, size_t bytes);  //!< async on pipe 0
void pipe_d2h(void* dst, const void* src, size_t bytes);  //!< async on pipe 1
//! opaque fence = everything queued so far on the main stream / on a pipe
void* fence_main();
void* fence_pipe(int pipe);
void main_wait(void* fence);           //!< main stream waits; consumes the fence
void pipe_wait(int pipe, void* fence); //!< pipe waits; consumes the fence
void pipe_sync(int pipe);              //!< host waits for the pipe
void fence_drop(void* fence);          //!< release a fence nobody waited for
//! pinned host staging memory (plain malloc in the host build)
void* hmalloc_pinned(size_t bytes);
void hfree_pinned(void* p);

//! threads of the host-side parallel loops (analysis, list building, Pade):
//! SANM_HOST_THREADS or min(cores, 16); independent of omp_set_num_threads()
//! calls made by other libraries in the process
int host_threads();

//! number of kernels launched by this library since process start
uint64_t launch_count();
void count_launch();

//! owning device buffer
template <typename T>
class DBuf {
    T* m_p = nullptr;
    size_t m_n = 0;

public:
    DBuf() = default;
    explicit DBuf(size_t n) { alloc(n); }
    DBuf(const DBuf&) = delete;
    DBuf& operator=(const DBuf&) = delete;
    DBuf(DBuf&& o) noexcept : m_p{o.m_p}, m_n{o.m_n} {
        o.m_p = nullptr;
        o.m_n = 0;
    }
    DBuf& operator=(DBuf&& o) noexcept {
        if (this != &o) {
            release();
            m_p = o.m_p;
            m_n = o.m_n;
            o.m_p = nullptr;
            o.m_n = 0;
        }
        return *this;
    }
    ~DBuf() { release(); }
    void release() {
        if (m_p) dfree(m_p);
        m_p = nullptr;
        m_n = 0;
    }
    void alloc(size_t n) {
        release();
        if (n) m_p = static_cast<T*>(dmalloc(n * sizeof(T)));
        m_n = n;
    }
    //! grow-only allocation; contents are NOT preserved
    void ensure(size_t n) {
        if (n > m_n) alloc(n);
    }
    void upload(const T* src, size_t n) {
        ensure(n);
        if (n) h2d(m_p, src, n * sizeof(T));
    }
    void upload(const std::vector<T>& v) { upload(v.data(), v.size()); }
    void download(T* dst, size_t n) const {
        if (n) d2h(dst, m_p, n * sizeof(T));
    }
    std::vector<T> to_host() const {
        std::vector<T> v(m_n);
        download(v.data(), m_n);
        return v;
    }
    void zero() {
        if (m_n) dzero(m_p, m_n * sizeof(T));
    }
    T* ptr() const { return m_p; }
    size_t size() const { return m_n; }
};

/*! A block functor may split off a prologue: `auto pre(const BlockCtx&) const`
 * reads ONLY data that no kernel writes (descriptors and lists uploaded at
 * analysis time) and its result is passed to `operator()(c, pre)`.  The CUDA
 * build runs it before the programmatic-dependent-launch wait, i.e. while the
 * previous kernel of the stream is still computing. */
template <class F, class = void>
struct HasPre : std::false_type {};
template <class F>
struct HasPre<F, std::void_t<decltype(&F::pre)>> : std::true_type {};

#ifdef SANM_HOSTEMU
/* ---------------------------------------------------------------- host */
template <class F>
void for_each(size_t n, const F& f) {
    count_launch();
    for (size_t i = 0; i < n; ++i) f(i);
}

struct BlockCtx {
    int tid = 0, nthr = 1;
    size_t bid = 0, nblk = 1;
    char* smem = nullptr;
    void sync() const {}
};

template <int MAXT = 0, int MINB = 1, class F>
void for_each_block(size_t nblocks, int /*threads*/, size_t smem_bytes,
                    const F& f) {
    count_launch();
    std::vector<char> sm(smem_bytes + 16);
    for (size_t b = 0; b < nblocks; ++b) {
        BlockCtx c;
        c.bid = b;
        c.nblk = nblocks;
        c.smem = sm.data();
        if constexpr (HasPre<F>::value) {
            f(c, f.pre(c));
        } else {
            f(c);
        }
    }
}

/*! Persistent launch: a fixed grid of co-resident blocks runs f(c, phase) for
 * phase = 0 .. nphases-1 with a grid-wide barrier (and memory fence) between
 * phases; inside a phase the blocks share the work by striding over it with
 * c.bid / c.nblk.  Replaces a chain of dependent kernel launches.  Host
 * build: three emulated blocks, phases as plain loops. */
template <int MAXT = 1024, class F>
void for_each_block_phased(int /*threads*/, size_t smem_bytes, uint32_t nphases,
                           const F& f) {
    count_launch();
    std::vector<char> sm(smem_bytes + 16);
    constexpr size_t NBLK = 3;
    for (uint32_t p = 0; p < nphases; ++p) {
        for (size_t b = 0; b < NBLK; ++b) {
            BlockCtx c;
            c.bid = b;
            c.nblk = NBLK;
            c.smem = sm.data();
            f(c, p);
        }
    }
}
#else
/* ---------------------------------------------------------------- CUDA */
template <class F>
__global__ void k_for_each(F f, size_t n) {
    // programmatic dependent launch, see k_for_each_block_b
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;");
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < n) f(i);
}

//! launch attribute of the kernels that start with griddepcontrol.wait
inline void pdl_config(cudaLaunchConfig_t& cfg, cudaLaunchAttribute* attr) {
    // SANM_NO_PDL=1: plain stream order (the kernels' griddepcontrol
    // instructions are then no-ops)
    static const bool pdl = getenv("SANM_NO_PDL") == nullptr;
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
}

template <class F, int THREADS = 128>
void for_each(size_t n, const F& f) {
    if (!n) return;
    count_launch();
    size_t blocks = (n + THREADS - 1) / THREADS;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)blocks);
    cfg.blockDim = dim3(THREADS);
    cfg.stream = stream();
    cudaLaunchAttribute attr[1];
    pdl_config(cfg, attr);
    SANM_CUDA_CHECK(cudaLaunchKernelEx(&cfg, k_for_each<F>, f, n));
}

struct BlockCtx {
    int tid, nthr;
    size_t bid, nblk;
    char* smem;
    __device__ void sync() const { __syncthreads(); }
};

/* MAXT = compile-time bound of the block size (__launch_bounds__).  It matters
 * beyond register allocation: without it ptxas schedules for 1024-thread
 * blocks and serialises independent global loads (load, use, load, use ...)
 * even where the PTX issues them back to back, which turns every latency-
 * bound kernel of the solver into a chain of memory round trips. */
template <class F>
__device__ __forceinline__ void run_block(F& f) {
    extern __shared__ __align__(16) char smem_raw[];
    BlockCtx c;
    c.tid = threadIdx.x;
    c.nthr = blockDim.x;
    c.bid = blockIdx.x;
    c.nblk = gridDim.x;
    c.smem = smem_raw;
    f(c);
}
//! MAXT = 0: no bound (the register-heavy element kernels size themselves)
template <class F, int MAXT>
__global__ void k_for_each_block(F f) {
    run_block(f);
}
//! bounded, and "one block per SM is enough": lets ptxas spend registers on
//! keeping loads in flight
template <class F, int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT > 0 ? MAXT : 1024, MINB) k_for_each_block_b(F f) {
    // Programmatic dependent launch (for_each_block below): wait for the
    // kernel before this one in the stream to finish and flush, THEN allow the
    // one after it to be staged -- its blocks become resident while this grid
    // computes and sit in their own wait, so a chain of short dependent
    // kernels (the solver's levels and pivot steps) does not pay the launch
    // latency per link.  Nothing is read or written before the wait.
    extern __shared__ __align__(16) char smem_raw[];
    BlockCtx c;
    c.tid = threadIdx.x;
    c.nthr = blockDim.x;
    c.bid = blockIdx.x;
    c.nblk = gridDim.x;
    c.smem = smem_raw;
    if constexpr (HasPre<F>::value) {
        // descriptor loads in flight across the wait (see HasPre)
        auto p = f.pre(c);
        asm volatile("griddepcontrol.wait;" ::: "memory");
        asm volatile("griddepcontrol.launch_dependents;");
        f(c, p);
    } else {
        asm volatile("griddepcontrol.wait;" ::: "memory");
        asm volatile("griddepcontrol.launch_dependents;");
        f(c);
    }
}

//! see the host build above; cooperative launch, grid = all co-resident blocks
template <class F, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) k_phased(F f, uint32_t nphases) {
    extern __shared__ __align__(16) char smem_raw[];
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    BlockCtx c;
    c.tid = threadIdx.x;
    c.nthr = blockDim.x;
    c.bid = blockIdx.x;
    c.nblk = gridDim.x;
    c.smem = smem_raw;
    for (uint32_t p = 0; p < nphases; ++p) {
        f(c, p);
        if (p + 1 < nphases) grid.sync();
    }
}

int sm_count();

template <int MAXT = 1024, class F>
void for_each_block_phased(int threads, size_t smem_bytes, uint32_t nphases,
                           const F& f) {
    if (!nphases) return;
    if (threads > MAXT) {
        throw ::sanm_b200::Error{SANM_ERR_ASSERT, "block size above its bound"};
    }
    count_launch();
    if (smem_bytes > 48 * 1024) {
        SANM_CUDA_CHECK(cudaFuncSetAttribute(
                k_phased<F, MAXT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                (int)smem_bytes));
    }
    int per_sm = 0;
    SANM_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(
            &per_sm, k_phased<F, MAXT>, threads, smem_bytes));
    if (per_sm < 1) {
        throw ::sanm_b200::Error{SANM_ERR_CUDA,
                                 "persistent kernel does not fit on an SM"};
    }
    F fcopy = f;
    uint32_t np = nphases;
    void* args[] = {&fcopy, &np};
    SANM_CUDA_CHECK(cudaLaunchCooperativeKernel(
            (const void*)k_phased<F, MAXT>, dim3((unsigned)(per_sm * sm_count())),
            dim3((unsigned)threads), args, smem_bytes, stream()));
}

template <int MAXT = 0, int MINB = 1, class F>
void for_each_block(size_t nblocks, int threads, size_t smem_bytes,
                    const F& f) {
    if (!nblocks) return;
    if (MAXT > 0 && threads > MAXT) {
        throw ::sanm_b200::Error{SANM_ERR_ASSERT, "block size above its bound"};
    }
    count_launch();
    void (*kern)(F);
    if constexpr (MAXT > 0) {
        kern = k_for_each_block_b<F, MAXT, MINB>;
    } else {
        kern = k_for_each_block<F, MAXT>;
    }
    if (smem_bytes > 48 * 1024) {
        SANM_CUDA_CHECK(cudaFuncSetAttribute(
                kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                (int)smem_bytes));
    }
    if constexpr (MAXT > 0) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)nblocks);
        cfg.blockDim = dim3((unsigned)threads);
        cfg.dynamicSmemBytes = smem_bytes;
        cfg.stream = stream();
        cudaLaunchAttribute attr[1];
        pdl_config(cfg, attr);
        SANM_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, f));
    } else {
        kern<<<(unsigned)nblocks, threads, smem_bytes, stream()>>>(f);
        SANM_CUDA_CHECK(cudaGetLastError());
    }
}
#endif

}  // namespace rt
}  // namespace sanm_b200
