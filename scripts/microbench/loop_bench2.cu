// Microbenchmark 4: sphere-pair loop (srt_packed.cuh) vs ray-pair filter+select loop (srt_pairs.cuh)
// in isolation.  Gtests/s is the number to read (592 SMSPs x 1.965 GHz x 32 lanes / Gtests/s = issue
// cycles per 32-ray sphere test; 15.0 is the floor of the reference's own instruction sequence).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o loop_bench2 loop_bench2.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../../simd-ray-tracer_b200/csrc/srt_pairs.cuh"
using namespace srt;

template <int MODE, int THREADS>
__global__ void __launch_bounds__(THREADS) kern(const float4 *sph, uint32_t n_pad, int reps, float *out, long long *cycles, float neg1, float one) {
    extern __shared__ float4 s_sph[];
    float4 *s_scan = s_sph + n_pad;
    for (uint32_t i = threadIdx.x; i < n_pad; i += blockDim.x) {
        const float4 s = sph[i];
        s_sph[i] = s;
        if (MODE == 1) {   // filter table, as srt_set_scene builds it
            const double u = 5.9604644775390625e-08, c2 = (double)s.x * s.x + (double)s.y * s.y + (double)s.z * s.z;
            s_scan[i] = make_float4(s.x, s.y, s.z, (float)((double)s.w * (1.0 + 128.0 * u) + 128.0 * u * c2 + 1e-30 - c2));
        }
    }
    __syncthreads();
    f3 o = mk3(0.1f * threadIdx.x, 0.2f, 0.3f + blockIdx.x * 1e-3f), d = mk3(0.6f * one, 0.64f * one, 0.48f * one);
    float acc = 0.f;
    long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
        if (MODE == 0) { Hit h = intersect_lanes_packed<8, true, 4>(s_sph, n_pad, o, d, neg1, one); acc += h.t; }
        if (MODE == 1) {
            Hit h0, h1;
            intersect_pair<8, true>(s_scan, (const float *)s_sph, sph, n_pad, n_pad,
                                    pk_make(o.x, o.x + 0.5f), pk_make(o.y, o.y), pk_make(o.z, o.z),
                                    pk_make(d.x, d.x), pk_make(d.y, d.y), pk_make(d.z, d.z), neg1, one, h0, h1);
            acc += h0.t + h1.t;
        }
        o.x += 1e-3f;
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int MODE, int THREADS> void run(const char *name, int blocks_per_sm, bool hits) {
    const uint32_t n = 1024;
    float4 *h = new float4[n];
    // never hit: far away spheres; "hits": a 32x32 grid in front of the rays with ~2 candidates per ray
    for (uint32_t i = 0; i < n; ++i)
        h[i] = hits ? make_float4(2.f + (i % 32) * 1.5f, 2.f + (i / 32) * 1.5f, 3.f + (i % 7), 0.09f)
                    : make_float4(100.f + i, 200.f + i, -50.f - i, 1e-6f);
    if (MODE == 0) {   // pair layout for the sphere-pair loop
        float4 *p = new float4[n];
        for (uint32_t k = 0; k < n / 2; ++k) {
            p[2 * k] = make_float4(h[2 * k].x, h[2 * k + 1].x, h[2 * k].y, h[2 * k + 1].y);
            p[2 * k + 1] = make_float4(h[2 * k].z, h[2 * k + 1].z, h[2 * k].w, h[2 * k + 1].w);
        }
        delete[] h; h = p;
    }
    float4 *d; float *out; long long *cyc, hc;
    cudaMalloc(&d, n * 16); cudaMemcpy(d, h, n * 16, cudaMemcpyHostToDevice);
    int blocks = 148 * blocks_per_sm, reps = 200;
    size_t smem = n * 32;
    cudaMalloc(&out, blocks * THREADS * 4); cudaMalloc(&cyc, 8);
    cudaFuncSetAttribute(kern<MODE, THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int occ = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern<MODE, THREADS>, THREADS, smem);
    kern<MODE, THREADS><<<blocks, THREADS, smem>>>(d, n, 2, out, cyc, -1.f, 1.f);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    kern<MODE, THREADS><<<blocks, THREADS, smem>>>(d, n, reps, out, cyc, -1.f, 1.f);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    cudaMemcpy(&hc, cyc, 8, cudaMemcpyDeviceToHost);
    const int rays_per_thread = MODE == 1 ? 2 : 1;
    double warps_per_smsp = blocks_per_sm * (THREADS / 32) / 4.0;
    double cyc_per_test = (double)hc / ((double)reps * n * warps_per_smsp * rays_per_thread);
    double gtests = (double)blocks * THREADS * rays_per_thread * reps * n / (ms * 1e6);
    printf("%-26s %s thr=%d blocks/SM=%d (occ %d) warps/SMSP=%.1f cycles/warp-test/SMSP=%.2f Gtests/s=%.0f %s\n", name, hits ? "hits  " : "nohits",
           THREADS, blocks_per_sm, occ, warps_per_smsp, cyc_per_test, gtests, cudaGetErrorString(cudaGetLastError()));
    cudaFree(d); cudaFree(out); cudaFree(cyc); delete[] h;
}
int main() {
    for (int hits = 0; hits < 2; ++hits) {
        for (int b : {1, 2}) run<0, 256>("sphere-pair (production)", b, hits);
        for (int b : {1, 2, 3, 4}) run<1, 256>("ray-pair + queue", b, hits);
        for (int b : {2, 4, 6, 8}) run<1, 128>("ray-pair + queue", b, hits);
    }
}
