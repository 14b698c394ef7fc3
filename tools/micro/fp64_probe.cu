// tools/micro/fp64_probe.cu -- FP64 throughput probes on the B200:
//   dfma        register-only chains of fma.rn.f64 (the FMA pipe)
//   dmma884     mma.sync.aligned.m8n8k4.row.col.f64  (legacy FP64 tensor path)
//   dmma16816   mma.sync.aligned.m16n8k16.row.col.f64
// Prints TFLOP/s of each (2 flops per multiply-add), whole GPU, CUDA events.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fp64_probe fp64_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { \
    printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

constexpr int ITERS = 4096;

__global__ void __launch_bounds__(256) k_dfma(double* out, double seed) {
    double a[16];
    const double x = seed + threadIdx.x * 1e-9, y = 1.0 - 1e-9;
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = i * 0.5;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) a[i] = fma(a[i], y, x);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) k_dmma884(double* out, double seed) {
    double c[8][2];
    const double a = seed + threadIdx.x * 1e-9, b = 1e-3;
#pragma unroll
    for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = i;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            asm volatile(
                "mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) k_dmma16816(double* out, double seed) {
    double c[4][4];
    double a[8], b[4];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = seed + threadIdx.x * 1e-9 + i;
#pragma unroll
    for (int i = 0; i < 4; ++i) b[i] = 1e-3 * (i + 1);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) c[i][j] = i + j;
    for (int it = 0; it < ITERS / 4; ++it) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            asm volatile(
                "mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 "
                "{%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
                : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]),
                  "d"(a[6]), "d"(a[7]), "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) s += c[i][j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class K>
static int run(const char* name, K kernel, double flops_per_thread, int blocks_per_sm) {
    int dev = 0, sms = 0;
    CK(cudaGetDevice(&dev));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int blocks = sms * blocks_per_sm, threads = 256;
    double* out;
    CK(cudaMalloc(&out, sizeof(double) * blocks * threads));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    kernel<<<blocks, threads>>>(out, 1.0);
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        CK(cudaEventRecord(e0));
        kernel<<<blocks, threads>>>(out, 1.0 + rep);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        best = ms < best ? ms : best;
    }
    const double tf = flops_per_thread * blocks * threads / (best * 1e-3) / 1e12;
    printf("%-10s blocks/SM=%d  %.3f ms  %.2f TFLOP/s\n", name, blocks_per_sm, best, tf);
    CK(cudaFree(out));
    return 0;
}

int main() {
    for (int bps : {1, 2, 4, 8}) {
        // dfma: 16 chains * ITERS * 2 flops per thread
        if (run("dfma", k_dfma, 16.0 * ITERS * 2, bps)) return 1;
    }
    for (int bps : {1, 2, 4, 8}) {
        // m8n8k4: 8*8*4*2 = 512 flops per warp-instruction = 16 per thread
        if (run("dmma884", k_dmma884, 8.0 * ITERS * 16, bps)) return 1;
    }
    for (int bps : {1, 2, 4, 8}) {
        // m16n8k16: 16*8*16*2 = 4096 flops per warp-instruction = 128 per thread
        if (run("dmma16816", k_dmma16816, 4.0 * (ITERS / 4) * 128, bps)) return 1;
    }
    return 0;
}
