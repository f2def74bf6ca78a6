// Microbenchmark 6: the production filter scans (srt_pairs.cuh) in isolation, no hits: issue cycles
// per sphere per warp pass (64 rays).  Floors by the 2-cycle FFMA2 model: secondary 18, primary 11.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o scan_bench scan_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../../simd-ray-tracer_b200/csrc/srt_pairs.cuh"
using namespace srt;

template <int MODE, int SUB>
__global__ void __launch_bounds__(256, 2) kern(const float4 *tab, uint32_t n_pad, int reps, uint32_t *out, float one) {
    extern __shared__ float4 s_tab[];
    for (uint32_t i = threadIdx.x; i < n_pad; i += blockDim.x) s_tab[i] = tab[i];
    __syncthreads();
    const float o = 0.1f * threadIdx.x;
    pk2 ox = pk_make(o, o + 0.5f), oy = pk_make(0.2f, 0.3f), oz = pk_make(0.3f, 0.1f);
    pk2 dx = pk_make(0.6f * one, 0.6f * one), dy = pk_make(0.64f * one, 0.64f * one), dz = pk_make(0.48f * one, 0.48f * one);
    uint32_t total = 0;
    for (int r = 0; r < reps; ++r) {
        CandQueue queue;
        uint32_t qn = 0;
        if (MODE == 0) {
            const ScanRays rays = make_scan_rays(ox, oy, oz, dx, dy, dz);
            scan_general_tile<SUB>(s_tab, n_pad, 0u, rays, queue, qn);
        } else {
            const ScanPrimary rays = make_scan_primary(0.1f, 0.2f, 0.3f, dx, dy, dz);
            scan_primary_tile<SUB>(s_tab, n_pad, 0u, rays, queue, qn);
        }
        for (uint32_t i = 0; i < qn; ++i) total += queue[i].x;
        ox = pk_fma(pk_make(1e-3f, 1e-3f), pk_make(one, one), ox);
        dx = pk_mul(dx, pk_make(one, one));
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = total;
}

template <int MODE, int SUB> void run(const char *name, int blocks_per_sm) {
    const uint32_t n = 1024;
    float4 *h = new float4[n];
    for (uint32_t i = 0; i < n; ++i) {
        const float cx = 100.f + i, cy = 200.f + i, cz = -50.f - (float)i, r2 = 1e-6f;
        h[i] = make_float4(cx, cy, cz, MODE == 0 ? (float)(r2 - ((double)cx * cx + (double)cy * cy + (double)cz * cz)) : -1.0e9f);
    }
    float4 *d; uint32_t *out;
    cudaMalloc(&d, n * 16); cudaMemcpy(d, h, n * 16, cudaMemcpyHostToDevice);
    const int blocks = 148 * blocks_per_sm, reps = 300;
    cudaMalloc(&out, blocks * 256 * 4);
    kern<MODE, SUB><<<blocks, 256, n * 16>>>(d, n, 2, out, 1.0f);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    kern<MODE, SUB><<<blocks, 256, n * 16>>>(d, n, reps, out, 1.0f);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double warps_per_smsp = blocks_per_sm * 8 / 4.0;
    printf("%-28s SUB=%2d warps/SMSP=%.0f  cycles per sphere per warp pass = %.2f  (%s)\n", name, SUB, warps_per_smsp,
           ms * 1e-3 * 1.965e9 / ((double)reps * n * warps_per_smsp), cudaGetErrorString(cudaGetLastError()));
    cudaFree(d); cudaFree(out); delete[] h;
}
int main() {
    for (int b : {1, 2}) {
        run<0, 32>("secondary scan (7 FFMA2)", b);
        run<0, 16>("secondary scan (7 FFMA2)", b);
        run<1, 32>("primary scan (4 FFMA2)", b);
    }
}
