// srt_device.cuh -- device-side arithmetic of the path-tracing hot path (sm_100a).
//
// Every floating-point operation that reaches the image is written with an explicit
// round-to-nearest intrinsic (__fadd_rn / __fmul_rn / __fdiv_rn / __fsqrt_rn), which nvcc
// never contracts into FMA, and the file is compiled with -fmad=false on top of that.
// __fmaf_rn appears exactly where the reference's AVX2 profile has _mm256_fmadd_ps
// (x64_math.h:409-420) and nowhere else.  That is what makes images bit-identical to the
// reference's CPU path under the per-path RNG stream contract.
//
// Citations are file:line in the reference (Ne0nWinds/SIMD-Ray-Tracer).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace srt {

constexpr float kEps = 1e-4f;   // F32Epsilon base.h:889
constexpr float kMax = 1e30f;   // F32Max     base.h:891
constexpr int kMaxBounce = 5;   // MaxRayBounce main.cpp:387

struct f3 { float x, y, z; };

// Debug build (-DSRT_SPIN_LIMIT=N): every wait loop of the pool kernel -- ring pop / push, batch acquisition, the
// TMA stage barrier -- reports its site to pinned host memory and traps after N iterations, so that a protocol bug
// reads as "spin limit at site s" from srt_wait instead of a hung GPU (compute-sanitizer is not available on the pool).
enum : uint32_t { kSpinPop = 1, kSpinPush = 2, kSpinAcquire = 3, kSpinMbar = 4 };
#ifdef SRT_SPIN_LIMIT
__device__ unsigned long long *g_spin_report;     // pinned + mapped host word of the handle that launched last
__device__ __noinline__ void spin_fail(uint32_t site, unsigned long long value = 0ull) {
    if (g_spin_report) {
        *(volatile unsigned long long *)(g_spin_report + 1) = value;
        *(volatile unsigned long long *)g_spin_report = 0xDEAD000000000000ull | ((unsigned long long)site << 32) | (blockIdx.x << 12) | (threadIdx.x & 0xFFFu);
        __threadfence_system();
    }
    __trap();
}
#define SRT_SPIN(site, n) do { if (++(n) > (uint32_t)(SRT_SPIN_LIMIT)) spin_fail(site); } while (0)
#define SRT_CHECK(site, cond) do { if (!(cond)) spin_fail(site); } while (0)      // index assertions of the debug build (sites >= 16)
#define SRT_CHECKV(site, cond, value) do { if (!(cond)) spin_fail(site, (unsigned long long)(value)); } while (0)
#else
#define SRT_SPIN(site, n) do { (void)(n); } while (0)
#define SRT_CHECK(site, cond) do { } while (0)
#define SRT_CHECKV(site, cond, value) do { } while (0)
#endif

__device__ __forceinline__ f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ f3 add3(f3 a, f3 b) { return mk3(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z)); }
__device__ __forceinline__ f3 sub3(f3 a, f3 b) { return mk3(__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y), __fsub_rn(a.z, b.z)); }
__device__ __forceinline__ f3 mul3(f3 a, f3 b) { return mk3(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y), __fmul_rn(a.z, b.z)); }
__device__ __forceinline__ f3 scale3(float s, f3 a) { return mk3(__fmul_rn(s, a.x), __fmul_rn(s, a.y), __fmul_rn(s, a.z)); }
__device__ __forceinline__ f3 neg3(f3 a) { return mk3(-a.x, -a.y, -a.z); }   // sign flip, x64_math.h:219-222
// v3::Dot x64_math.h:224-227: mul,mul,mul then (x+y)+z
__device__ __forceinline__ float dot3(f3 a, f3 b) {
    return __fadd_rn(__fadd_rn(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y)), __fmul_rn(a.z, b.z));
}
// v3::Normalize x64_math.h:234-245 (NormalizeFast == Normalize, wasm_math.h:168-170)
__device__ __forceinline__ f3 normalize3(f3 a) {
    const float lsq = dot3(a, a);
    if (!(lsq > kEps)) return mk3(0.0f, 0.0f, 0.0f);
    const float len = __fsqrt_rn(lsq);
    return mk3(__fdiv_rn(a.x, len), __fdiv_rn(a.y, len), __fdiv_rn(a.z, len));
}

// Out-of-line copy: normalize3 is ~1 KB of SASS (3 IEEE divisions + sqrt) and a path step uses
// it 6 times.  The pair / pool kernels call this one to stay closer to the 32 KB L1.5
// instruction cache; the one-path kernel for tiny, shading-bound scenes keeps it inline.
__device__ __noinline__ f3 normalize3_call(f3 a) { return normalize3(a); }
template <bool OUTLINE>
__device__ __forceinline__ f3 normalize3_sel(f3 a) { return OUTLINE ? normalize3_call(a) : normalize3(a); }

// ---- PRNG: u32_random_state::PCG base.h:954-963, RandomFloat base.h:983-989 ------------
// InverseMaxInt = (f32)((f64)(Max-Min)/4294967295.0) is exactly 2^-32 for a unit range and
// 2^-31 for a range of 2 (checked on the host in tests/test_host_logic.py).
constexpr float kInvU32Range1 = 2.3283064365386963e-10f;   // 2^-32
constexpr float kInvU32Range2 = 4.6566128730773926e-10f;   // 2^-31

__device__ __forceinline__ uint32_t pcg_next(uint64_t &s) {
    const uint64_t old = s;
    s = old * 6364136223846793005ULL + 1442695040888963407ULL;
    const uint32_t v = (uint32_t)(old >> 32) ^ (uint32_t)old;
    return __funnelshift_r(v, v, (uint32_t)(old >> 59));    // RotateRight32 x64_math.h:168-170
}
__device__ __forceinline__ float rand_float(uint64_t &s, float inv_range, float mn) {
    return __fadd_rn(__fmul_rn(__uint2float_rn(pcg_next(s)), inv_range), mn);
}
// Per-path stream seed: main.cpp:667-676 evaluated on a 64-bit index, with ONE change: a shift count of zero
// becomes 32.  In the reference's expression `s ^= s >> 0` zeroes the seed, so every index with i % 64 == 57
// (first shift) or == 41 (second shift) ends on the SAME value.  The reference only ever feeds it thread numbers;
// fed path indices, 1/32 of all paths -- whole pixel columns when W*H % 64 == 0 -- would share one stream for
// every sample and never converge (stream contract).  All other indices keep the reference's value.
__host__ __device__ __forceinline__ uint64_t stream_seed(uint64_t i) {
    uint64_t s = 0x420247153476526ULL * i;
    s += 0x8442885C91A5C8DULL;
    const uint32_t k1 = (uint32_t)(7 + i) & 63u, k2 = (uint32_t)(0x29 ^ i) & 63u;
    s ^= s >> (k1 ? k1 : 32u);
    s ^= s << 23;
    s ^= s >> (k2 ? k2 : 32u);
    s = (s * 0x11C19226CEB4769AULL) + 0x1105404122082911ULL;
    s ^= s << 19;
    s ^= s >> 13;
    return s;
}

// ---- Reflectance main.cpp:292-300 ---------------------------------------------------------
__device__ __forceinline__ float reflectance(float c, float eta) {
    float r0 = __fdiv_rn(__fsub_rn(1.0f, eta), __fadd_rn(1.0f, eta));
    r0 = __fmul_rn(r0, r0);
    float r1 = __fsub_rn(1.0f, c);
    r1 = __fmul_rn(__fmul_rn(__fmul_rn(__fmul_rn(r1, r1), r1), r1), r1);
    return __fadd_rn(r0, __fmul_rn(__fsub_rn(1.0f, r0), r1));
}

// ---- Tonemap + pack main.cpp:312-346 ------------------------------------------------------
__device__ __forceinline__ float saturate_ref(float v) {   // x64_math.h:97-101 (NaN passes through)
    if (v < 0.0f) return 0.0f;
    if (v > 1.0f) return 1.0f;
    return v;
}
__device__ __forceinline__ float linear_to_srgb(float l) {
    l = saturate_ref(l);
    if (l < 0.0031308f) return __fmul_rn(l, 12.92f);
    return __fsqrt_rn(l);
}
__device__ __forceinline__ uint32_t to_u8(float v) {
    const float s = __fmul_rn(saturate_ref(v), 255.0f);
    return (uint32_t)__float2int_rz(s) & 0xFFu;             // NaN -> 0, like cvttss2si's low byte
}
__device__ __forceinline__ uint32_t pack_rgba8(float r, float g, float b) {
    return to_u8(linear_to_srgb(r)) | (to_u8(linear_to_srgb(g)) << 8) | (to_u8(linear_to_srgb(b)) << 16) | (255u << 24);
}

// ---- device-side scene records --------------------------------------------------------------
// Sphere: float4 {cx, cy, cz, r*r}.  r*r is the reference's own f32 product (main.cpp:406),
// taken once at upload instead of once per test; the radius itself is never used again.
struct DevMaterial {
    float4 color_spec;   // Color.xyz, Specular
    float4 emis_ior;     // Emissive.xyz, IndexOfRefraction
    float4 derived;      // x = (f32)(1.0 - (f64)Specular) (main.cpp:463), y = 1.0f / IOR (main.cpp:466)
};

struct DevCamera {
    float pos[3], x[3], y[3], film_center[3];
    float film_w, film_h;
};

// x / d for a divisor fixed per launch and x < 2^30: q = (x * m) >> sh with m = ceil(2^(30+l) / d), 2^l >= d
// (Granlund-Montgomery; 3 instructions instead of the ~20 of a 32-bit division -- the pool kernel's hot code
// has to stay inside the 32 KB instruction cache).
struct MagicDiv { uint32_t m, sh; };
__host__ __device__ __forceinline__ uint32_t magic_div(uint32_t x, MagicDiv d) { return (uint32_t)(((uint64_t)x * d.m) >> d.sh); }

struct KernelParams {
    const float4 *spheres;       // n_pad entries {cx,cy,cz,r2}, zero spheres as padding (never hit);
                                 // packed kernels: pair layout {x0,x1,y0,y1},{z0,z1,r2_0,r2_1}
    const float4 *scan_table;    // pair kernel: n_pad entries {cx,cy,cz,-kappa} (srt_pairs.cuh), padding -1e30
    const float4 *ptab;          // streamed scenes: the launch's slab table (3 float4 per 4 spheres, srt_pairs.cuh), built per launch
    int32_t slab_axis;           // coordinate axis the slab planes contain (0/1/2), chosen at scene upload
    float omax, omax2;           // bound on |ray origin| the slab radii were inflated for, and its square
    const DevMaterial *materials;
    uint32_t n_pad;              // multiple of 8
    uint32_t n_spheres;          // real sphere count
    int32_t use_sky;
    DevCamera cam;
    uint32_t width, height;
    uint32_t blocks_x;           // 8x4 pixel blocks per row
    MagicDiv div_bx;             // / blocks_x
    uint32_t total_items;        // blocks_x * blocks_y * 32
    uint32_t first_sample, n_samples, sample_stride, accum_mode;
    uint64_t stream_base;
    float4 *accum;
    uint32_t *rgba;
    uint32_t *tap_prim;          // optional
    float *tap_t;                // optional
    unsigned long long *rays;
    unsigned int *work_counter;
    unsigned int *ticket;        // CTAs that have left the launch; the last one publishes *rays to *host_rays and re-arms the counters
    unsigned long long *host_rays;   // pinned + mapped host memory: path segments cast by the launch (main.cpp:390)
    uint32_t part, n_parts;      // image partition over devices (srt_set_partition): local block row L is block row L * n_parts + part
    unsigned long long *debug;   // optional diagnostics (pool kernel)
    // pool kernel work items: (sample chunk, pixel), chunk-major, so that the end of a launch does not
    // wait for single pixels to run all their samples
    uint32_t *done;              // per pixel: done_tag | number of chunks folded into accum[pix] in this launch
    uint32_t done_tag;           // launch sequence number << 12
    uint32_t chunk_span;         // samples per chunk * sample_stride
    uint32_t items_per_gen;      // blocks_x * blocks_y * 32; pool: total_items = items_per_gen * chunks
    uint32_t regen_threshold;    // start new paths once this many lanes of a warp are idle
    // pool kernel, bundle mode: a work item is ONE sample; 2^bundle_log2 consecutive samples of a
    // pixel are consecutive items, so a primary warp pass holds all samples of one (or a few) pixels and the
    // level-1 slab test is done once per (pixel, sphere) for the whole bundle.  Finished samples are staged and
    // folded into the running mean in sample order by the slot that delivers the chunk's last one.
    uint32_t bundle_log2;        // 0: bundle mode off
    uint32_t n_chunks;           // sample chunks per pixel = ceil(n_samples / 2^bundle_log2)
    uint32_t band_items;         // pixels are processed in bands of whole 8x4-block rows: pixel items per band
    uint32_t band_last;          // pixel items of the last band (<= band_items)
    uint32_t n_bands;
    uint32_t band_pix;           // pixel-index distance of the same pixel position one band earlier (4 * block rows * width)
    MagicDiv div_full, div_band, div_last;   // / (band_items * n_chunks), / band_items, / band_last
    float bundle_rho;            // bound on |F(j) - F(0)| over a pixel's jitter footprint, film units, kept axes (host, double)
    float two_over_w, two_over_h;
    float4 *stage;               // [band_items][2^bundle_log2] finished samples {r,g,b,-} of the pixels in flight
    uint32_t *cnt;               // [band_items] samples staged for the row's current chunk
    const float2 *fold_w;        // [n_samples] {1/N, (N-1)/N} of main.cpp:484-487 for the launch's samples (IEEE, host)
    float neg_one, one;          // -1.0f / +1.0f, deliberately opaque to the compiler (srt_packed.cuh)
};

struct Hit {
    float t;
    uint32_t prim;
    bool hit, inside;
};

// v3x::Dot / LengthSquared of the intersection loop: profile I x64_math.h:299-309,
// profile F x64_math.h:409-420.
template <bool FMA>
__device__ __forceinline__ float lane_dot(float ax, float ay, float az, float bx, float by, float bz) {
    if (FMA) return __fmaf_rn(az, bz, __fmaf_rn(ay, by, __fmul_rn(ax, bx)));
    return __fadd_rn(__fadd_rn(__fmul_rn(ax, bx), __fmul_rn(ay, by)), __fmul_rn(az, bz));
}

// A3 + A4 with SIMD-lane semantics (main.cpp:392-455): sphere s is lane s % W of group
// s / W.  Every lane keeps its own MinT (strict <), group and sticky inside bit; the winner
// is the first lane holding the horizontal minimum.  The always-executed part is 15 FP32
// issue slots (profile F) or 19 (profile I) + 1 compare per test; the rest sits behind the
// rare `d2 < r2` branch.
template <int W, bool FMA>
__device__ __forceinline__ Hit intersect_lanes(const float4 *__restrict__ sph, uint32_t n_pad, f3 o, f3 d) {
    float min_t[W];
    uint32_t group[W];
    uint32_t inside_bits = 0;
#pragma unroll
    for (int l = 0; l < W; ++l) { min_t[l] = kMax; group[l] = 0; }

#pragma unroll 1
    for (uint32_t base = 0; base < n_pad; base += 8) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float4 s = sph[base + j];                     // warp-uniform LDS.128 broadcast
            const float scx = __fsub_rn(s.x, o.x), scy = __fsub_rn(s.y, o.y), scz = __fsub_rn(s.z, o.z);   // :401
            const float t = lane_dot<FMA>(scx, scy, scz, d.x, d.y, d.z);                                  // :402
            const float vx = __fsub_rn(scx, __fmul_rn(d.x, t));                                          // :403,407
            const float vy = __fsub_rn(scy, __fmul_rn(d.y, t));
            const float vz = __fsub_rn(scz, __fmul_rn(d.z, t));
            const float d2 = lane_dot<FMA>(vx, vy, vz, vx, vy, vz);
            if (d2 < s.w) {                                                                               // :409
                constexpr int kLaneMask = W - 1;
                const int l = j & kLaneMask;                    // static after unrolling
                const float x = __fsqrt_rn(__fsub_rn(s.w, d2));                                           // :413
                float it = __fsub_rn(t, x);                                                               // :415
                const bool test = it < kEps;                                                              // :416
                if (test) it = __fadd_rn(t, x);                                                           // :417
                if (it < min_t[l] && it > kEps) {                                                         // :419-421
                    min_t[l] = it;                                                                        // :427
                    group[l] = (base + j) / W;                                                            // :426
                    if (test) inside_bits |= (1u << l);                                                   // :425 sticky
                }
            }
        }
    }
    float mv = min_t[0];
#pragma unroll
    for (int l = 1; l < W; ++l) mv = (min_t[l] < mv) ? min_t[l] : mv;                                     // HorizontalMin :432
    Hit h;
    h.t = mv;
    h.hit = !(mv == kMax);                                                                                // :433
    int idx = W - 1;
#pragma unroll
    for (int l = W - 2; l >= 0; --l) idx = (min_t[l] == mv) ? l : idx;                                    // FindFirstIndex :441
    uint32_t g = group[0];
#pragma unroll
    for (int l = 1; l < W; ++l) g = (idx == l) ? group[l] : g;
    h.prim = h.hit ? g * W + (uint32_t)idx : 0xFFFFFFFFu;                                                 // :443
    h.inside = (inside_bits >> idx) & 1u;                                                                 // :455
    return h;
}

// A3 + A4 with RenderTileScalar semantics (main.cpp:541-588): d2 == r2 is a hit, a later
// sphere wins ties, t == 1e-4 is accepted, inside is overwritten.
__device__ __forceinline__ Hit intersect_scalar(const float4 *__restrict__ sph, uint32_t n_pad, uint32_t n_real, f3 o, f3 d) {
    float min_t = kMax;
    uint32_t prim = 0;
    bool inside = false;
#pragma unroll 1
    for (uint32_t base = 0; base < n_pad; base += 8) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float4 s = sph[base + j];
            const float scx = __fsub_rn(s.x, o.x), scy = __fsub_rn(s.y, o.y), scz = __fsub_rn(s.z, o.z);
            const float t = lane_dot<false>(scx, scy, scz, d.x, d.y, d.z);
            const float vx = __fsub_rn(scx, __fmul_rn(d.x, t));
            const float vy = __fsub_rn(scy, __fmul_rn(d.y, t));
            const float vz = __fsub_rn(scz, __fmul_rn(d.z, t));
            const float d2 = lane_dot<false>(vx, vy, vz, vx, vy, vz);
            // `d2 > r2 -> continue` lets NaN through, so padding is excluded by index.
            if (!(d2 > s.w) && (base + j) < n_real) {                                                                            // :557
                const float x = __fsqrt_rn(__fsub_rn(s.w, d2));
                float it = __fsub_rn(t, x);
                const bool test = it < kEps;
                if (test) it = __fadd_rn(t, x);
                if (!(it > min_t) && !(it < kEps)) {                                                      // :568-569
                    inside = test; prim = base + j; min_t = it;
                }
            }
        }
    }
    Hit h;
    h.t = min_t;
    h.hit = !(min_t == kMax);
    h.prim = h.hit ? prim : 0xFFFFFFFFu;
    h.inside = inside;
    return h;
}

} // namespace srt
