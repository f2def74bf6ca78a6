// Microbenchmark 3: the production intersection loop in isolation (no hits, no shading).
// Reports SMSP cycles per warp-level sphere test for several variants.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../simd-ray-tracer_b200/csrc/srt_packed.cuh"
using namespace srt;

template <int MODE>
__global__ void __launch_bounds__(256) kern(const float4 *sph, uint32_t n_pad, int reps, float *out, long long *cycles, float neg1, float one) {
    extern __shared__ float4 s_sph[];
    for (uint32_t i = threadIdx.x; i < n_pad; i += blockDim.x) s_sph[i] = sph[i];
    __syncthreads();
    f3 o = mk3(0.1f * threadIdx.x, 0.2f, 0.3f + blockIdx.x * 1e-3f), d = mk3(0.6f, 0.64f, 0.48f);
    float acc = 0.f;
    long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
        Hit h;
        if (MODE == 0) h = intersect_lanes_packed<8, true, 4>(s_sph, n_pad, o, d, neg1, one);
        if (MODE == 1) h = intersect_lanes<8, true>(s_sph, n_pad, o, d);
        if (MODE == 2) h = intersect_lanes_packed<8, true, 2>(s_sph, n_pad, o, d, neg1, one);
        if (MODE == 3) h = intersect_lanes_packed<8, false, 4>(s_sph, n_pad, o, d, neg1, one);
        acc += h.t;
        o.x += 1e-3f;
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int MODE> void run(const char *name, int blocks_per_sm) {
    const uint32_t n = 1024;
    float4 *h = new float4[n];
    for (uint32_t i = 0; i < n; ++i) h[i] = make_float4(100.f + i, 200.f + i, -50.f - i, 1e-6f);   // never hit (also fine as pair layout)
    float4 *d; float *out; long long *cyc, hc;
    cudaMalloc(&d, n * 16); cudaMemcpy(d, h, n * 16, cudaMemcpyHostToDevice);
    int blocks = 148 * blocks_per_sm, reps = 200;
    cudaMalloc(&out, blocks * 256 * 4); cudaMalloc(&cyc, 8);
    cudaFuncSetAttribute(kern<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    int occ = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern<MODE>, 256, n * 16);
    kern<MODE><<<blocks, 256, n * 16>>>(d, n, 2, out, cyc, -1.f, 1.f);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    kern<MODE><<<blocks, 256, n * 16>>>(d, n, reps, out, cyc, -1.f, 1.f);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    cudaMemcpy(&hc, cyc, 8, cudaMemcpyDeviceToHost);
    int warps_per_smsp = blocks_per_sm * 2;
    double cyc_per_test = (double)hc / ((double)reps * n * warps_per_smsp);
    double gtests = (double)blocks * 256 * reps * n / (ms * 1e6);
    printf("%-34s blocks/SM=%d (occ limit %d) cycles/warp-test/SMSP=%.2f  Gtests/s=%.0f  err=%s\n", name, blocks_per_sm, occ, cyc_per_test, gtests, cudaGetErrorString(cudaGetLastError()));
    cudaFree(d); cudaFree(out); cudaFree(cyc); delete[] h;
}
int main() {
    for (int b : {1, 2, 3}) {
        run<0>("packed W8 FMA, branch per 8", b);
        run<2>("packed W8 FMA, branch per 4", b);
        run<1>("scalar-issue W8 FMA", b);
        run<3>("packed W8 IEEE(no FMA), per 8", b);
    }
}
