// srt_packed.cuh -- the brute-force intersection loop on Blackwell's packed FP32 pipe.
//
// sm_100 adds add/mul/fma.rn.f32x2 (SASS FADD2 / FMUL2 / FFMA2): one issue slot does the
// same IEEE-754 round-to-nearest operation on two independent fp32 values held in an aligned
// register pair.  The loop below tests TWO spheres (2i, 2i+1) per instruction stream against
// one ray: 15 packed instructions per sphere pair in the FMA profile (19 in the IEEE profile)
// instead of 30 (38) scalar ones, which takes the loop from issue-bound to FP32-pipe-bound.
// Each half performs exactly the scalar sequence of srt_device.cuh, so results are
// bit-identical.
//
// ptxas 12.9 caveat (measured): it contracts `mul.rn.f32x2` feeding
// `add/sub.rn.f32x2` into one FFMA2 even though both carry an explicit .rn and -fmad=false
// is set -- which would change the rounding.  Wherever the reference has a separately
// rounded multiply followed by an add/sub, the add/sub is therefore issued as
// fma2(x, K, y) with K = +1 or -1 read from kernel parameters (opaque to the compiler):
// RN(x*(+-1) + y) == RN(y +- x) exactly, and ptxas cannot fuse through it.
#pragma once
#include "srt_device.cuh"

namespace srt {

typedef unsigned long long pk2;   // two fp32 in an aligned 64-bit register pair: {lo, hi}

__device__ __forceinline__ pk2 pk_make(float lo, float hi) {
    pk2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void pk_split(pk2 v, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ pk2 pk_sub(pk2 a, pk2 b) {
    pk2 r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ pk2 pk_mul(pk2 a, pk2 b) {
    pk2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ pk2 pk_fma(pk2 a, pk2 b, pk2 c) {
    pk2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}

// v3x::Dot / LengthSquared on two spheres at once.  one = {+1,+1} (opaque).
template <bool FMA>
__device__ __forceinline__ pk2 pk_dot(pk2 ax, pk2 ay, pk2 az, pk2 bx, pk2 by, pk2 bz, pk2 one) {
    if (FMA) return pk_fma(az, bz, pk_fma(ay, by, pk_mul(ax, bx)));            // x64_math.h:409-420
    const pk2 mx = pk_mul(ax, bx), my = pk_mul(ay, by), mz = pk_mul(az, bz);    // x64_math.h:299-309
    return pk_fma(pk_fma(mx, one, my), one, mz);                                // (x + y) + z, each rounded
}

// Shared-memory layout: sphere pair p = (2p, 2p+1) occupies two float4:
//   sph2[2p]   = { x0, x1, y0, y1 }      sph2[2p+1] = { z0, z1, r2_0, r2_1 }
// Same lane semantics as intersect_lanes (srt_device.cuh): sphere s is lane s % W of group
// s / W, per-lane MinT / group / sticky inside bit, first lane with the minimum wins.
template <int W, bool FMA, int PAIRS>
__device__ __forceinline__ Hit intersect_lanes_packed(const float4 *__restrict__ sph2, uint32_t n_pad, f3 o, f3 d,
                                                      float neg_one_f, float one_f) {
    static_assert((2 * PAIRS) % W == 0, "a loop body must cover whole SIMD groups: the lane of sphere e is e % W");
    float min_t[W];
    uint32_t group[W];
    uint32_t inside_bits = 0;
#pragma unroll
    for (int l = 0; l < W; ++l) { min_t[l] = kMax; group[l] = 0; }
    const pk2 ox = pk_make(o.x, o.x), oy = pk_make(o.y, o.y), oz = pk_make(o.z, o.z);
    const pk2 dx = pk_make(d.x, d.x), dy = pk_make(d.y, d.y), dz = pk_make(d.z, d.z);
    const pk2 neg_one = pk_make(neg_one_f, neg_one_f), one = pk_make(one_f, one_f);

#pragma unroll 1
    for (uint32_t base = 0; base < n_pad; base += 2 * PAIRS) {
        // Always-executed part, branch-free for 8 spheres (4 independent dependency chains
        // for the scheduler to interleave): 15 packed FP32 instructions per sphere pair.
        pk2 t2[PAIRS], dd[PAIRS];
        float r2[2 * PAIRS];
        bool any = false;
#pragma unroll
        for (int p = 0; p < PAIRS; ++p) {
            const float4 a = sph2[base + 2 * p];           // warp-uniform LDS.128 broadcasts
            const float4 b = sph2[base + 2 * p + 1];
            const pk2 scx = pk_sub(pk_make(a.x, a.y), ox);                                   // main.cpp:401
            const pk2 scy = pk_sub(pk_make(a.z, a.w), oy);
            const pk2 scz = pk_sub(pk_make(b.x, b.y), oz);
            t2[p] = pk_dot<FMA>(scx, scy, scz, dx, dy, dz, one);                             // :402
            const pk2 vx = pk_fma(pk_mul(dx, t2[p]), neg_one, scx);                          // :403,407  sc - d*t
            const pk2 vy = pk_fma(pk_mul(dy, t2[p]), neg_one, scy);
            const pk2 vz = pk_fma(pk_mul(dz, t2[p]), neg_one, scz);
            dd[p] = pk_dot<FMA>(vx, vy, vz, vx, vy, vz, one);
            float d2_0, d2_1;
            pk_split(dd[p], d2_0, d2_1);
            r2[2 * p] = b.z;
            r2[2 * p + 1] = b.w;
            any |= (d2_0 < b.z) | (d2_1 < b.w);                                              // :409 HitMask
        }
        if (any) {                                                                           // :411 (rare)
#pragma unroll
            for (int e = 0; e < 2 * PAIRS; ++e) {
                float lo, hi;
                pk_split(dd[e >> 1], lo, hi);
                const float d2 = (e & 1) ? hi : lo;
                if (d2 < r2[e]) {
                    pk_split(t2[e >> 1], lo, hi);
                    const float t = (e & 1) ? hi : lo;
                    constexpr int kLaneMask = W - 1;
                    const int l = e & kLaneMask;          // static when 2*PAIRS is a multiple of W
                    const float x = __fsqrt_rn(__fsub_rn(r2[e], d2));                        // :413
                    float it = __fsub_rn(t, x);                                              // :415
                    const bool test = it < kEps;                                             // :416
                    if (test) it = __fadd_rn(t, x);                                          // :417
                    if (it < min_t[l] && it > kEps) {                                        // :419-421
                        min_t[l] = it;                                                       // :427
                        group[l] = (base + e) / W;                                           // :426
                        if (test) inside_bits |= (1u << l);                                  // :425 sticky
                    }
                }
            }
        }
    }
    float mv = min_t[0];
#pragma unroll
    for (int l = 1; l < W; ++l) mv = (min_t[l] < mv) ? min_t[l] : mv;                        // HorizontalMin :432
    Hit h;
    h.t = mv;
    h.hit = !(mv == kMax);                                                                   // :433
    int idx = W - 1;
#pragma unroll
    for (int l = W - 2; l >= 0; --l) idx = (min_t[l] == mv) ? l : idx;                       // FindFirstIndex :441
    uint32_t g = group[0];
#pragma unroll
    for (int l = 1; l < W; ++l) g = (idx == l) ? group[l] : g;
    h.prim = h.hit ? g * W + (uint32_t)idx : 0xFFFFFFFFu;                                    // :443
    h.inside = (inside_bits >> idx) & 1u;                                                    // :455
    return h;
}

// Centre + r^2 of sphere `idx` from the paired layout.
__device__ __forceinline__ float4 paired_sphere(const float4 *__restrict__ sph2, uint32_t idx) {
    const float4 a = sph2[(idx >> 1) * 2], b = sph2[(idx >> 1) * 2 + 1];
    return (idx & 1u) ? make_float4(a.y, a.w, b.y, b.w) : make_float4(a.x, a.z, b.x, b.z);
}

} // namespace srt
