// Microbenchmark 6: the production filter scans (srt_pairs.cuh) in isolation, no hits: issue cycles
// per sphere per warp pass (64 rays).  Floors by the 2-cycle FFMA2 model: secondary 18, primary 11.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o scan_bench scan_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../../simd-ray-tracer_b200/csrc/srt_pairs.cuh"
using namespace srt;

// candidate iff |a| <= r' for either ray of the pair; a = p.up + q.uq - ou  (2 FFMA2 + 2 FSETP + @p LOP3)
__device__ __forceinline__ void slab_sphere(uint32_t &mask, uint32_t bit, float p, float q, float rr, pk2 up, pk2 uq, pk2 nou) {
    const pk2 a = pk_fma(pk_make(p, p), up, pk_fma(pk_make(q, q), uq, nou));
    float aa, ab;
    pk_split(a, aa, ab);
    asm("{\n\t.reg .pred p;\n\t.reg .f32 t, u;\n\tabs.f32 t, %1;\n\tabs.f32 u, %2;\n\tsetp.le.f32 p, t, %3;\n\tsetp.le.or.f32 p, u, %3, p;\n\t@p or.b32 %0, %0, %4;\n\t}"
        : "+r"(mask) : "f"(aa), "f"(ab), "f"(rr), "r"(bit));
}
// table: per 4 spheres {p0..3}, {q0..3}, {r0..3}
template <int LAYOUT>
__device__ __forceinline__ void slab_scan(const float4 *__restrict__ tab, uint32_t count, pk2 up, pk2 uq, pk2 nou, CandQueue &queue, uint32_t &qn) {
#pragma unroll 1
    for (uint32_t c = 0; c < (count >> 5); ++c) {
        uint32_t m = 0;
        if (LAYOUT == 0) {
            const float4 *blk = tab + c * 24;
#pragma unroll
            for (int g = 0; g < 8; ++g) {
                const float4 P = blk[3 * g], Q = blk[3 * g + 1], R = blk[3 * g + 2];
                slab_sphere(m, 0x80000000u >> (4 * g + 0), P.x, Q.x, R.x, up, uq, nou);
                slab_sphere(m, 0x80000000u >> (4 * g + 1), P.y, Q.y, R.y, up, uq, nou);
                slab_sphere(m, 0x80000000u >> (4 * g + 2), P.z, Q.z, R.z, up, uq, nou);
                slab_sphere(m, 0x80000000u >> (4 * g + 3), P.w, Q.w, R.w, up, uq, nou);
            }
        } else {
            const float4 *blk = tab + c * 32;
#pragma unroll
            for (int j = 0; j < 32; ++j) { const float4 S = blk[j]; slab_sphere(m, 0x80000000u >> j, S.x, S.y, S.z, up, uq, nou); }
        }
        if (m) { queue[qn] = make_uint2(m, c << 5); qn += 1; }
    }
}

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
        } else if (MODE == 2 || MODE == 3) {
            slab_scan<MODE - 2>(s_tab, n_pad, dx, dz, oy, queue, qn);
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
        if (MODE >= 2) { h[i] = make_float4(1e6f + i, 2e6f + i, 0.5f, 0.f); continue; }
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
    for (int b : {1, 2, 3}) {
        run<0, 32>("secondary scan (7 FFMA2)", b);
        run<0, 16>("secondary scan (7 FFMA2)", b);
        run<1, 32>("primary scan (4 FFMA2)", b);
        run<2, 32>("slab scan SoA4 (2 FFMA2)", b);
        run<3, 32>("slab scan AoS LDS.128 (2 FFMA2)", b);
    }
}
