// Microbenchmark 7: a HALF-precision slab scan (2 HFMA2 + 1 HSETP2 + @p LOP3 per sphere per ray
// pair, both rays of a thread in the halves of one register) in isolation, no hits: issue cycles
// per sphere per warp pass (64 rays).  Question: does HFMA2 hold the issue port for 1 cycle
// (then the floor is ~4.5) or for 2 like FFMA2 (then nothing is gained over the f32x2 slab scan)?
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o hslab_bench hslab_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

// candidate iff |a| <= r' for either half; a = p.up + q.uq + nou
// acc |= sel & fail, fail = 0xffff per half whose |a| > r' (ordered: a NaN never fails)
__device__ __forceinline__ void hslab_sphere(uint32_t &acc, uint32_t sel, uint32_t pp, uint32_t qq, uint32_t rr, uint32_t up, uint32_t uq, uint32_t nou) {
    const __half2 a = __hfma2(*(__half2 *)&pp, *(__half2 *)&up, __hfma2(*(__half2 *)&qq, *(__half2 *)&uq, *(__half2 *)&nou));
    acc |= sel & __hgt2_mask(__habs2(a), *(__half2 *)&rr);
}
__device__ __forceinline__ uint32_t lo2(uint32_t x) { uint32_t d; asm("{\n\t.reg .b16 l, h;\n\tmov.b32 {l, h}, %1;\n\tmov.b32 %0, {l, l};\n\t}" : "=r"(d) : "r"(x)); return d; }
__device__ __forceinline__ uint32_t hi2(uint32_t x) { uint32_t d; asm("{\n\t.reg .b16 l, h;\n\tmov.b32 {l, h}, %1;\n\tmov.b32 %0, {h, h};\n\t}" : "=r"(d) : "r"(x)); return d; }

// LAYOUT 0: per 8 spheres three uint4 {p0..7}, {q0..7}, {r0..7} (halves), broadcast by PRMT / operand selector
// LAYOUT 1: per 4 spheres three uint4 of pre-duplicated half2 {p0p0,p1p1,p2p2,p3p3}, {q..}, {r..}
template <int LAYOUT>
__device__ __forceinline__ void hslab_scan(const uint4 *__restrict__ tab, uint32_t count, uint32_t up, uint32_t uq, uint32_t nou, uint2 *queue, uint32_t &qn) {
#pragma unroll 1
    for (uint32_t c = 0; c < (count >> 5); ++c) {
        uint32_t f[2] = {0, 0};               // fail bits: low 16 = ray a, high 16 = ray b, bit 15 - j = sphere j of the 16
        if (LAYOUT == 0) {
            const uint4 *blk = tab + c * 12;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                const uint4 P = blk[3 * g], Q = blk[3 * g + 1], R = blk[3 * g + 2];
                const uint32_t b = 0x80008000u >> (8 * (g & 1));
                uint32_t &m = f[g >> 1];
                hslab_sphere(m, b >> 0, lo2(P.x), lo2(Q.x), lo2(R.x), up, uq, nou);
                hslab_sphere(m, b >> 1, hi2(P.x), hi2(Q.x), hi2(R.x), up, uq, nou);
                hslab_sphere(m, b >> 2, lo2(P.y), lo2(Q.y), lo2(R.y), up, uq, nou);
                hslab_sphere(m, b >> 3, hi2(P.y), hi2(Q.y), hi2(R.y), up, uq, nou);
                hslab_sphere(m, b >> 4, lo2(P.z), lo2(Q.z), lo2(R.z), up, uq, nou);
                hslab_sphere(m, b >> 5, hi2(P.z), hi2(Q.z), hi2(R.z), up, uq, nou);
                hslab_sphere(m, b >> 6, lo2(P.w), lo2(Q.w), lo2(R.w), up, uq, nou);
                hslab_sphere(m, b >> 7, hi2(P.w), hi2(Q.w), hi2(R.w), up, uq, nou);
            }
        } else {
            const uint4 *blk = tab + c * 24;
#pragma unroll
            for (int g = 0; g < 8; ++g) {
                const uint4 P = blk[3 * g], Q = blk[3 * g + 1], R = blk[3 * g + 2];
                const uint32_t b = 0x80008000u >> (4 * (g & 3));
                uint32_t &m = f[g >> 2];
                hslab_sphere(m, b >> 0, P.x, Q.x, R.x, up, uq, nou);
                hslab_sphere(m, b >> 1, P.y, Q.y, R.y, up, uq, nou);
                hslab_sphere(m, b >> 2, P.z, Q.z, R.z, up, uq, nou);
                hslab_sphere(m, b >> 3, P.w, Q.w, R.w, up, uq, nou);
            }
        }
        // a sphere is a candidate of the pair unless it fails for both rays
        const uint32_t both0 = f[0] & (f[0] >> 16), both1 = f[1] & (f[1] >> 16);
        const uint32_t m = ~((both0 << 16) | (both1 & 0xffffu));
        if (m) { queue[qn] = make_uint2(m, c << 5); qn += 1; }
    }
}

template <int LAYOUT>
__global__ void __launch_bounds__(256, 2) kern(const uint4 *tab, uint32_t n16, uint32_t n_sph, int reps, uint32_t *out, uint32_t step, float f0, float f1) {
    extern __shared__ uint4 s_tab[];
    for (uint32_t i = threadIdx.x; i < n16; i += blockDim.x) s_tab[i] = tab[i];
    __syncthreads();
    __half2 up = __floats2half2_rn(f0, f1), uq = __floats2half2_rn(f1, -f0), nou = __floats2half2_rn(0.1f * threadIdx.x, 0.2f);
    uint32_t total = 0;
    for (int r = 0; r < reps; ++r) {
        uint2 queue[32];
        uint32_t qn = 0;
        hslab_scan<LAYOUT>(s_tab, n_sph, *(uint32_t *)&up, *(uint32_t *)&uq, *(uint32_t *)&nou, queue, qn);
        for (uint32_t i = 0; i < qn; ++i) total += queue[i].x;
        *(uint32_t *)&nou += step;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = total;
}

// pure HFMA2 / FFMA2 issue rate: 8 independent chains
template <int KIND>
__global__ void __launch_bounds__(256, 2) rate(int reps, uint32_t *out, uint32_t seed) {
    uint32_t a[8];
    unsigned long long f[8];
    for (int i = 0; i < 8; ++i) { a[i] = seed + i; f[i] = seed + i; }
    const uint32_t x = seed * 3 + threadIdx.x, y = seed * 5;
    const unsigned long long fx = x, fy = y;
    for (int r = 0; r < reps; ++r) {
#pragma unroll
        for (int k = 0; k < 8; ++k)
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (KIND == 0) asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(x), "r"(y));
                else asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(f[i]) : "l"(fx), "l"(fy));
            }
    }
    uint32_t t = 0;
    for (int i = 0; i < 8; ++i) t += a[i] + (uint32_t)f[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
}

template <int LAYOUT> void run(const char *name, int blocks_per_sm) {
    const uint32_t n = 1024, n16 = LAYOUT == 0 ? n / 8 * 3 : n / 4 * 3;
    uint4 *h = new uint4[n16];
    const __half2 big = __floats2half2_rn(3000.f, 3000.f), rad = __floats2half2_rn(0.5f, 0.5f);
    for (uint32_t i = 0; i < n16; ++i) { const uint32_t v = (i % 3 == 2) ? *(const uint32_t *)&rad : *(const uint32_t *)&big; h[i] = make_uint4(v, v, v, v); }
    uint4 *d; uint32_t *out;
    cudaMalloc(&d, n16 * 16); cudaMemcpy(d, h, n16 * 16, cudaMemcpyHostToDevice);
    const int blocks = 148 * blocks_per_sm, reps = 300;
    cudaMalloc(&out, blocks * 256 * 4);
    kern<LAYOUT><<<blocks, 256, n16 * 16>>>(d, n16, n, 2, out, 0, 0.6f, 0.8f);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    kern<LAYOUT><<<blocks, 256, n16 * 16>>>(d, n16, n, reps, out, 0, 0.6f, 0.8f);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double warps_per_smsp = blocks_per_sm * 8 / 4.0;
    printf("%-44s warps/SMSP=%.0f  cycles per sphere per warp pass = %.2f  (%s)\n", name, warps_per_smsp,
           ms * 1e-3 * 1.965e9 / ((double)reps * n * warps_per_smsp), cudaGetErrorString(cudaGetLastError()));
    cudaFree(d); cudaFree(out); delete[] h;
}
template <int KIND> void run_rate(const char *name) {
    uint32_t *out; cudaMalloc(&out, 148 * 2 * 256 * 4);
    const int reps = 2000;
    rate<KIND><<<296, 256>>>(2, out, 1);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    rate<KIND><<<296, 256>>>(reps, out, 1);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("%-44s warp-instructions / clk / SMSP = %.3f\n", name, (double)reps * 64 * 4 / (ms * 1e-3 * 1.965e9));
    cudaFree(out);
}
int main() {
    run_rate<0>("HFMA2 (8 chains x 4 warps/SMSP)");
    run_rate<1>("FFMA2 (8 chains x 4 warps/SMSP)");
    for (int b : {1, 2, 3}) {
        run<0>("half slab scan, packed halves + PRMT", b);
        run<1>("half slab scan, pre-duplicated half2", b);
    }
}
