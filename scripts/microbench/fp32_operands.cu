// Microbenchmark 2: does operand fetch (distinct register sources) limit FFMA/FFMA2 issue?
// Patterns over independent chains with all-distinct source registers.
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 d; asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ float fma1(float a, float b, float c) { float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }
__device__ __forceinline__ float mul1(float a, float b) { float d; asm volatile("mul.rn.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b)); return d; }
__device__ __forceinline__ u64 mk(float a, float b) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }

// MODE 0: scalar fma x = fma(x, y_i, z_i) (3 distinct regs)   1: scalar mul x = mul(x, y_i)
// MODE 2: packed fma 3 distinct pairs                          3: packed mul 2 distinct pairs
// MODE 4: packed fma, 2 pairs + same pair twice (x = fma(y,y,x))
template <int MODE>
__global__ void kern(float *out, int iters, float seed, long long *cycles) {
    constexpr int N = 8;
    float x[N], y[N], z[N]; u64 X[N], Y[N], Z[N];
#pragma unroll
    for (int i = 0; i < N; ++i) {
        x[i] = seed + i + threadIdx.x * 1e-6f; y[i] = 0.999f - i * 1e-4f; z[i] = 1e-3f * i;
        X[i] = mk(x[i], x[i] + 1); Y[i] = mk(y[i], y[i]); Z[i] = mk(z[i], z[i]);
    }
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int i = 0; i < N; ++i) {
                if (MODE == 0) x[i] = fma1(x[i], y[i], z[i]);
                if (MODE == 1) x[i] = mul1(x[i], y[i]);
                if (MODE == 2) X[i] = fma2(X[i], Y[i], Z[i]);
                if (MODE == 3) X[i] = mul2(X[i], Y[i]);
                if (MODE == 4) X[i] = fma2(Y[i], Y[i], X[i]);
            }
        }
    }
    long long t1 = clock64();
    float acc = 0;
#pragma unroll
    for (int i = 0; i < N; ++i) { float lo, hi; asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(X[i])); acc += x[i] + lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}
template <int MODE> void run(const char *name, int w) {
    int threads = 128 * w, blocks = 148, iters = 20000;
    float *out; long long *cyc, h;
    cudaMalloc(&out, (size_t)threads * blocks * 4); cudaMalloc(&cyc, 8);
    kern<MODE><<<blocks, threads>>>(out, 100, 1.0f, cyc);
    kern<MODE><<<blocks, threads>>>(out, iters, 1.0f, cyc);
    cudaDeviceSynchronize();
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-44s warps/SMSP=%d  cycles/warp-inst/SMSP=%.3f\n", name, w, (double)h / ((double)iters * 4 * 8 * w));
    cudaFree(out); cudaFree(cyc);
}
int main() {
    for (int w : {2, 6}) {
        run<0>("scalar FFMA 3 distinct regs", w);
        run<1>("scalar FMUL 2 distinct regs", w);
        run<2>("packed FFMA2 3 distinct pairs", w);
        run<3>("packed FMUL2 2 distinct pairs", w);
        run<4>("packed FFMA2 y,y,x", w);
    }
}
