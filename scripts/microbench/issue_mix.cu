// Microbenchmark 5: can the SM sub-partition issue ALU / LSU instructions in the shadow of a packed
// FFMA2 (which holds the FMA pipe for 2 cycles)?  Each variant runs a loop body of 8 independent
// FFMA2 plus K other instructions; cycles per body per SMSP tell whether the extras are free.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o issue_mix issue_mix.cu
#include <cstdio>
#include <cuda_runtime.h>

#define FFMA2(d, a, b) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b))

template <int MODE>
__global__ void __launch_bounds__(256) kern(int reps, unsigned long long *out, long long *cycles, unsigned long long a0, unsigned long long b0, float thr) {
    __shared__ float4 tab[256];
    tab[threadIdx.x] = make_float4(threadIdx.x, 1.f, 2.f, 3.f);
    __syncthreads();
    unsigned long long x0 = a0 + threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    const unsigned long long a = a0, b = b0;
    unsigned m0 = threadIdx.x, m1 = 1, m2 = 2, m3 = 3;
    float f0 = thr * 0.5f, f1 = thr * 2.f;
    float4 acc4 = make_float4(0, 0, 0, 0);
    long long t0 = clock64();
#pragma unroll 1
    for (int r = 0; r < reps; ++r) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            FFMA2(x0, a, b); FFMA2(x1, a, b); FFMA2(x2, a, b); FFMA2(x3, a, b);
            if (MODE == 1 || MODE == 5) {          // 4 integer ALU ops (LOP3)
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(m0) : "r"(m1), "r"(m2));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(m3) : "r"(m1), "r"(m2));
            }
            if (MODE == 2 || MODE == 5) {          // FSETP + predicated OR, as the scan loop
                asm volatile("{ .reg .pred p; setp.gt.f32 p, %1, %2; setp.gt.or.f32 p, %3, %2, p; @p or.b32 %0, %0, 1; }" : "+r"(m1) : "f"(f0), "f"(thr), "f"(f1));
            }
            if (MODE == 3 || MODE == 5) {          // one LDS.128 broadcast
                float4 v;
                asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"((unsigned)__cvta_generic_to_shared(&tab[(r + u) & 255])));
                acc4.x += 0.f * v.x;   // keep
                if (v.y == 12345.f) acc4.y += v.z + v.w;
            }
            if (MODE == 4) {                        // 4 scalar FMUL on the FMA pipe (control: must cost)
                asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(f0) : "f"(thr));
                asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(f1) : "f"(thr));
            }
            FFMA2(x4, a, b); FFMA2(x5, a, b); FFMA2(x6, a, b); FFMA2(x7, a, b);
            if (MODE == 1 || MODE == 5) {
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(m0) : "r"(m1), "r"(m3));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(m2) : "r"(m1), "r"(m3));
            }
            if (MODE == 2) {
                asm volatile("{ .reg .pred p; setp.gt.f32 p, %1, %2; setp.gt.or.f32 p, %3, %2, p; @p or.b32 %0, %0, 2; }" : "+r"(m2) : "f"(f1), "f"(thr), "f"(f0));
            }
            if (MODE == 4) {
                asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(f0) : "f"(thr));
                asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(f1) : "f"(thr));
            }
        }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7 + m0 + m1 + m2 + m3 + (unsigned long long)(f0 + f1 + acc4.x + acc4.y);
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int MODE> void run(const char *name, int extra, int blocks_per_sm) {
    unsigned long long *out; long long *cyc, hc;
    cudaMalloc(&out, 148 * 8 * 256 * 8); cudaMalloc(&cyc, 8);
    const int reps = 4000;
    float one = 1.0f; unsigned long long a, b; float av[2] = {one, one}, bv[2] = {0.5f, 0.25f};
    memcpy(&a, av, 8); memcpy(&b, bv, 8);
    kern<MODE><<<148 * blocks_per_sm, 256>>>(10, out, cyc, a, b, 1.0f);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    kern<MODE><<<148 * blocks_per_sm, 256>>>(reps, out, cyc, a, b, 1.0f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    cudaMemcpy(&hc, cyc, 8, cudaMemcpyDeviceToHost);
    const double warps_per_smsp = blocks_per_sm * 8 / 4.0;
    const double per_body = (double)hc / reps / 8 / warps_per_smsp;   // cycles per (8 FFMA2 + extras) per warp slot on one SMSP
    printf("%-44s warps/SMSP=%.0f  cycles per 8 FFMA2 + %d extra = %.2f  [events: %.2f]  (%s)\n", name, warps_per_smsp, extra, per_body, ms * 1e-3 * 1.965e9 / reps / 8 / warps_per_smsp, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out); cudaFree(cyc);
}

int main() {
    for (int bps : {1, 2, 4}) {
        run<0>("8 FFMA2", 0, bps);
        run<1>("8 FFMA2 + 4 LOP3", 4, bps);
        run<2>("8 FFMA2 + 2x(2 FSETP + @p OR)", 6, bps);
        run<3>("8 FFMA2 + 1 LDS.128 (+FMUL,FSETP)", 3, bps);
        run<4>("8 FFMA2 + 4 FMUL (control)", 4, bps);
        run<5>("8 FFMA2 + 4 LOP3 + 3 setp/or + LDS", 10, bps);
    }
    return 0;
}
