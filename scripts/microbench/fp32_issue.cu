// Microbenchmark: FP32 issue/pipe throughput on sm_100a for scalar FFMA vs packed FFMA2 and
// mixes.  Reports warp-instructions per cycle per SMSP and FMA lanes per cycle per SM.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp32_issue fp32_issue.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ float fma1(float a, float b, float c) { float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }

template <int NS, int NP>   // NS scalar chains, NP packed chains, all independent
__global__ void kern(float *out, int iters, float seed, long long *cycles) {
    float s[NS > 0 ? NS : 1]; u64 p[NP > 0 ? NP : 1];
    float a = seed + threadIdx.x * 1e-7f, b = 0.999f;
    u64 a2, b2;
    asm("mov.b64 %0, {%1,%2};" : "=l"(a2) : "f"(a), "f"(a));
    asm("mov.b64 %0, {%1,%2};" : "=l"(b2) : "f"(b), "f"(b));
#pragma unroll
    for (int i = 0; i < NS; ++i) s[i] = a + i;
#pragma unroll
    for (int i = 0; i < NP; ++i) p[i] = a2 + i;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int i = 0; i < (NS > NP ? NS : NP); ++i) {
                if (i < NP) p[i] = fma2(p[i], b2, a2);
                if (i < NS) s[i] = fma1(s[i], b, a);
            }
        }
    }
    long long t1 = clock64();
    float acc = 0; 
#pragma unroll
    for (int i = 0; i < NS; ++i) acc += s[i];
#pragma unroll
    for (int i = 0; i < NP; ++i) { float lo, hi; asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[i])); acc += lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int NS, int NP>
void run(const char *name, int warps_per_smsp) {
    int threads = 128 * warps_per_smsp, blocks = 148, iters = 20000;
    float *out; long long *cyc, h;
    cudaMalloc(&out, (size_t)threads * blocks * 4); cudaMalloc(&cyc, 8);
    kern<NS, NP><<<blocks, threads>>>(out, 100, 1.0f, cyc);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    kern<NS, NP><<<blocks, threads>>>(out, iters, 1.0f, cyc);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    double insts_per_warp = (double)iters * 4 * (NS + NP);
    double ipc = insts_per_warp * warps_per_smsp / (double)h;                 // warp-insts / cycle / SMSP
    double lanes = (double)iters * 4 * (NS + 2.0 * NP) * warps_per_smsp * 32 * 4 / (double)h;   // FMA lanes / cycle / SM
    printf("%-28s warps/SMSP=%d  cycles=%lld  ms=%.3f  warp-inst/cyc/SMSP=%.3f  FMA lanes/cyc/SM=%.1f\n", name, warps_per_smsp, h, ms, ipc, lanes);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    for (int w : {1, 2, 4, 8}) {
        run<8, 0>("scalar FFMA x8", w);
        run<0, 8>("packed FFMA2 x8", w);
        run<4, 4>("mix 4 FFMA + 4 FFMA2", w);
        run<8, 4>("mix 8 FFMA + 4 FFMA2", w);
        run<4, 8>("mix 4 FFMA + 8 FFMA2", w);
    }
    return 0;
}
