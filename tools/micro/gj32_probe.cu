// Micro-probe (tools only; not part of the library): 32x32 Gauss-Jordan with
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o gj32_probe gj32_probe.cu
// partial pivoting by 4 warps, each owning 8 columns; checks inv(P A) and
// prints the cycles per phase.
#include <cstdio>
#include <cstdint>
#include <cmath>
#include <cstdlib>
#include <cuda_runtime.h>
constexpr int NB = 32;
#define SANM_GJ_PROBE
#include "../../sanm_b200/csrc/gj32.cuh"
__global__ void probe(const double* A, double* out, uint32_t* lperm_out, uint32_t nb, long long* cyc) {
    __shared__ double D[NB * NB];
    __shared__ double acol[2 * NB];
    __shared__ uint32_t lperm[NB];
    __shared__ int npert;
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int j = w; j < NB; j += 4) D[j * NB + lane] = (lane < nb && j < (int)nb) ? A[j * NB + lane] : 0.0;
    __syncthreads();
    long long c0 = clock64();
    sanm_b200::block_gj32(D, nb, lperm, acol, 1e-300, &npert);
    long long c1 = clock64();
    for (int j = w; j < NB; j += 4) out[j * NB + lane] = D[j * NB + lane];
    if (threadIdx.x < NB) lperm_out[threadIdx.x] = lperm[threadIdx.x];
    if (threadIdx.x == 0) cyc[0] = c1 - c0;
}
int main() {
    for (uint32_t nb : {32u, 32u, 32u, 19u, 1u}) {
        double h[NB * NB], o[NB * NB];
        uint32_t lp[NB];
        srand(7); for (int i = 0; i < NB * NB; ++i) h[i] = rand() / (double)RAND_MAX - 0.5;
        double *dA, *dO; long long* dc; uint32_t* dl;
        cudaMalloc(&dA, sizeof h); cudaMalloc(&dO, sizeof h); cudaMalloc(&dc, 64); cudaMalloc(&dl, 128);
        cudaMemcpy(dA, h, sizeof h, cudaMemcpyHostToDevice);
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        cudaEventRecord(e0);
        probe<<<1, 128>>>(dA, dO, dl, nb, dc);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        long long c; cudaMemcpy(&c, dc, 8, cudaMemcpyDeviceToHost);
        cudaMemcpy(o, dO, sizeof o, cudaMemcpyDeviceToHost);
        cudaMemcpy(lp, dl, sizeof lp, cudaMemcpyDeviceToHost);
        // check: (P A) * out == I, row i of P A = row lperm[i] of A
        double err = 0;
        for (uint32_t i = 0; i < nb; ++i)
            for (uint32_t j = 0; j < nb; ++j) {
                double s = 0;
                for (uint32_t m = 0; m < nb; ++m) s += h[m * NB + lp[i]] * o[j * NB + m];
                err = fmax(err, fabs(s - (i == j)));
            }
        printf("nb %u: %.1f us, %lld cycles, max |PA*inv - I| = %.3g (%s)\n", nb, ms * 1e3, c, err,
               cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
