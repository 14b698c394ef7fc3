/* sanm_b200/csrc/rt.cu -- runtime layer implementation (see rt.h). */
#include "rt.h"

#include <omp.h>
#include <cstdlib>

#include <atomic>

namespace sanm_b200 {
namespace rt {

int host_threads() {
    static const int n = [] {
        if (const char* e = getenv("SANM_HOST_THREADS")) {
            int v = atoi(e);
            if (v > 0) return v;
        }
        int p = omp_get_num_procs();
        return p < 1 ? 1 : p > 16 ? 16 : p;
    }();
    return n;
}

namespace {
std::atomic<uint64_t> g_launches{0};
}

uint64_t launch_count() { return g_launches.load(); }
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

#ifdef SANM_HOSTEMU

stream_t stream() { return 0; }
void set_stream(void*) {}
void sync() {}
void* dmalloc(size_t bytes) {
    void* p = malloc(bytes ? bytes : 1);
    if (!p) sanm_throw(SANM_ERR_GENERIC, "host-emulation malloc(%zu) failed", bytes);
    return p;
}
void dfree(void* p) { free(p); }
void h2d(void* dst, const void* src, size_t bytes) { memcpy(dst, src, bytes); }
void d2h(void* dst, const void* src, size_t bytes) { memcpy(dst, src, bytes); }
void d2d(void* dst, const void* src, size_t bytes) { memmove(dst, src, bytes); }
void dzero(void* dst, size_t bytes) { memset(dst, 0, bytes); }
void pipe_h2d(void* dst, const void* src, size_t bytes) { memcpy(dst, src, bytes); }
void pipe_d2h(void* dst, const void* src, size_t bytes) { memcpy(dst, src, bytes); }
void* fence_main() { return nullptr; }
void* fence_pipe(int) { return nullptr; }
void main_wait(void*) {}
void pipe_wait(int, void*) {}
void pipe_sync(int) {}
void fence_drop(void*) {}
ScopedPipe::ScopedPipe(int) {}
ScopedPipe::~ScopedPipe() {}
void* hmalloc_pinned(size_t bytes) { return malloc(bytes ? bytes : 1); }
void hfree_pinned(void* p) { free(p); }

#else

namespace {
cudaStream_t g_stream = nullptr, g_user_stream = nullptr;
bool g_device_ok = false, g_use_user_stream = false;
cudaStream_t g_override = nullptr;  //!< ScopedPipe
}  // namespace

void check_device() {
    if (g_device_ok) return;
    int n = 0;
    cudaError_t err = cudaGetDeviceCount(&n);
    if (err != cudaSuccess || n <= 0) {
        // no CPU fallback: the product path needs a CUDA device
        throw Error{SANM_ERR_CUDA,
                    std::string{"sanm_b200 requires a CUDA device (sm_100a): "} +
                            (err != cudaSuccess ? cudaGetErrorString(err)
                                                : "no device found")};
    }
    SANM_CUDA_CHECK(cudaStreamCreateWithFlags(&g_stream, cudaStreamNonBlocking));
    {
        // device memory comes from the stream-ordered pool and stays cached in
        // it: solvers are created and destroyed once per ANM iteration
        // (libsanm/anm.cpp:205), cudaMalloc/cudaFree of their 0.1-1 GB buffers
        // would cost 1-100 ms each time
        int dev = 0;
        SANM_CUDA_CHECK(cudaGetDevice(&dev));
        cudaMemPool_t pool;
        SANM_CUDA_CHECK(cudaDeviceGetDefaultMemPool(&pool, dev));
        uint64_t keep = UINT64_MAX;
        SANM_CUDA_CHECK(cudaMemPoolSetAttribute(
                pool, cudaMemPoolAttrReleaseThreshold, &keep));
    }
    g_device_ok = true;
}

stream_t stream() {
    check_device();
    if (g_override) return g_override;
    return g_use_user_stream ? g_user_stream : g_stream;
}

void set_stream(void* s) {
    check_device();
    // finish what is queued on the old stream before switching
    SANM_CUDA_CHECK(cudaStreamSynchronize(stream()));
    g_user_stream = static_cast<cudaStream_t>(s);
    g_use_user_stream = s != nullptr;
}

void sync() { SANM_CUDA_CHECK(cudaStreamSynchronize(stream())); }

int sm_count() {
    static const int n = [] {
        check_device();
        int dev = 0, v = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
        return v > 0 ? v : 1;
    }();
    return n;
}

void* dmalloc(size_t bytes) {
    check_device();
    void* p = nullptr;
    SANM_CUDA_CHECK(cudaMallocAsync(&p, bytes ? bytes : 1, stream()));
    return p;
}

void dfree(void* p) {
    // stream-ordered: kernels queued on the library stream that still use p
    // run before the memory is handed out again (the copy pipes are fenced
    // against the main stream before their buffers can be released)
    if (p) cudaFreeAsync(p, g_device_ok ? stream() : nullptr);
}

void h2d(void* dst, const void* src, size_t bytes) {
    SANM_CUDA_CHECK(
            cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, stream()));
    // the source may be pageable / reused by the caller right away
    SANM_CUDA_CHECK(cudaStreamSynchronize(stream()));
}

void d2h(void* dst, const void* src, size_t bytes) {
    SANM_CUDA_CHECK(
            cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, stream()));
    SANM_CUDA_CHECK(cudaStreamSynchronize(stream()));
}

void d2d(void* dst, const void* src, size_t bytes) {
    SANM_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice,
                                    stream()));
}

void dzero(void* dst, size_t bytes) {
    SANM_CUDA_CHECK(cudaMemsetAsync(dst, 0, bytes, stream()));
}

namespace {
cudaStream_t g_pipe[4] = {nullptr, nullptr, nullptr, nullptr};
cudaStream_t pipe_stream(int pipe) {
    check_device();
    if (!g_pipe[pipe]) {
        SANM_CUDA_CHECK(cudaStreamCreateWithFlags(&g_pipe[pipe], cudaStreamNonBlocking));
    }
    return g_pipe[pipe];
}
void* fence_on(cudaStream_t s) {
    cudaEvent_t ev;
    SANM_CUDA_CHECK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    SANM_CUDA_CHECK(cudaEventRecord(ev, s));
    return ev;
}
void wait_on(cudaStream_t s, void* fence) {
    cudaEvent_t ev = static_cast<cudaEvent_t>(fence);
    SANM_CUDA_CHECK(cudaStreamWaitEvent(s, ev, 0));
    // the wait has been queued; the event's resources are released by the
    // runtime once it has completed
    SANM_CUDA_CHECK(cudaEventDestroy(ev));
}
}  // namespace

void pipe_h2d(void* dst, const void* src, size_t bytes) {
    SANM_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice,
                                    pipe_stream(PIPE_H2D)));
}
void pipe_d2h(void* dst, const void* src, size_t bytes) {
    SANM_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost,
                                    pipe_stream(PIPE_D2H)));
}
void* fence_main() { return fence_on(stream()); }
void* fence_pipe(int pipe) { return fence_on(pipe_stream(pipe)); }
void main_wait(void* fence) { wait_on(stream(), fence); }
void pipe_wait(int pipe, void* fence) { wait_on(pipe_stream(pipe), fence); }
void pipe_sync(int pipe) { SANM_CUDA_CHECK(cudaStreamSynchronize(pipe_stream(pipe))); }
void fence_drop(void* fence) {
    if (fence) cudaEventDestroy(static_cast<cudaEvent_t>(fence));
}
ScopedPipe::ScopedPipe(int pipe) {
    sanm_check(!g_override, "nested ScopedPipe");
    g_override = pipe_stream(pipe);
}
ScopedPipe::~ScopedPipe() { g_override = nullptr; }

void* hmalloc_pinned(size_t bytes) {
    check_device();
    void* p = nullptr;
    SANM_CUDA_CHECK(cudaMallocHost(&p, bytes ? bytes : 1));
    return p;
}

void hfree_pinned(void* p) {
    if (p) cudaFreeHost(p);
}

#endif

}  // namespace rt
}  // namespace sanm_b200
