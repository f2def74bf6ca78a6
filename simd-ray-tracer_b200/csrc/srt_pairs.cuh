// srt_pairs.cuh -- brute-force intersection for a PAIR OF RAYS per thread (sm_100a):
// a conservative fused-arithmetic FILTER over all spheres, then the reference's exact
// arithmetic on the few survivors.
//
// Why.  ncu on the sphere-pair kernel (profiles/r1b_ncu_full_render_paths_v1.txt) plus the
// microbenchmarks in scripts/microbench show that on B200 a packed FFMA2/FMUL2/FADD2 holds the
// scheduler's issue port for 2 cycles: the 15 separately rounded instructions of
// main.cpp:401-409 cost 30 issue cycles per sphere per warp however they are scheduled, and
// every other instruction (LDS, compare, hit bookkeeping) adds to that.  The only way below
// that floor is to NOT execute the reference's instruction sequence for spheres the ray
// misses -- while still testing every sphere against every ray and still producing the
// reference's bits.
//
// Scan (all spheres, 7 packed FMAs + 2 compares + 1 mask update per sphere per ray pair).
//   With w = C - O and tau = w.D the reference decides a hit on d2 = |w - D tau|^2
//   = |w|^2 - tau^2 (2 - |D|^2).  The scan evaluates the same quantity from per-ray and
//   per-sphere constants,
//       z = tt^2 + (2 C.O - kappa_i),   tt = C.D' - O.D',   D' = D sqrt(2 - |D|^2),
//       kappa_i = |C_i|^2 - R_i,        candidate  <=>  z > oo',
//   where R_i = r2_i (1 + 128u) + 128u |C_i|^2 is the sphere's squared radius inflated on the
//   host (in double) and oo' = |O|^2 (1 - 160u) is deflated per ray, u = 2^-24.  DESIGN.md
//   section 5.3 bounds every rounding of both evaluation orders and shows that the
//   reference's `d2 < r2` (SIMD) or `!(d2 > r2)` (scalar) implies z > oo' with a 1.8x safety
//   factor: the scan never loses a hit, it only admits a few near misses.
//   `srt_filter_audit` (tests/test_gpu_filter.py) checks that superset property exhaustively
//   on the GPU for every (ray, sphere) pair of whole renders.
// Select (candidates only).  The reference's own instruction sequence, packed over the two
//   rays, in increasing sphere index: t and d2 have the reference's bits, so hit/miss,
//   closest-hit choice, tie-break and inside flag are bit-identical.
//
// Lane semantics.  Sphere s is lane s % W of group s / W; strict `<` per lane; HorizontalMin,
// first lane holding it (main.cpp:419-455).  That reduces to: smallest t wins; equal t ->
// lower lane; equal lane -> earlier group.  The sticky InsideSphere bit (main.cpp:425) of the
// winning lane is the OR of `test` over that lane's accepted updates; it can differ from the
// winner's own `test` only if an earlier candidate of the same lane had test == true, which
// is detected (`test_lanes`) and resolved by the literal per-lane emulation (replay_inside).
#pragma once
#include "srt_device.cuh"
#include "srt_packed.cuh"

namespace srt {

constexpr int kPairsMaxSpheres = 8192;                  // resident-scene limit of the pair kernel
constexpr int kPairsMaxWords = kPairsMaxSpheres / 32;   // candidate bitmap words per thread
constexpr float kIdleOrigin = 3.0e18f;                  // parked slot: no sphere is ever a candidate
constexpr float kFilterU = 5.9604644775390625e-08f;     // u = 2^-24

// Running closest-hit state of one ray while candidates are replayed.
struct Best {
    float t;               // MinT after HorizontalMin
    uint32_t idx;          // sphere index of the winner
    uint32_t test;         // winner's own `t - x < eps` flag (main.cpp:416)
    uint32_t test_lanes;   // lanes in which a valid test == true candidate was seen
};
__device__ __forceinline__ Best best_init() { Best b; b.t = kMax; b.idx = 0; b.test = 0; b.test_lanes = 0; return b; }

// main.cpp:413-427 for one candidate of one ray (the hit test was done by the caller).
template <int W>
__device__ __forceinline__ void best_update(Best &B, float t, float d2, float r2, uint32_t idx) {
    const float x = __fsqrt_rn(__fsub_rn(r2, d2));                       // :413
    float it = __fsub_rn(t, x);                                          // :415
    const bool test = it < kEps;                                         // :416
    if (test) it = __fadd_rn(t, x);                                      // :417
    if (W == 1) {                                                        // RenderTileScalar :568-576
        if (!(it > B.t) && !(it < kEps)) { B.t = it; B.idx = idx; B.test = test; }
    } else if (it > kEps) {                                              // :419-421
        const uint32_t lane = idx & (uint32_t)(W - 1);
        if (test) B.test_lanes |= 1u << lane;
        if (it < B.t || (it == B.t && lane < (B.idx & (uint32_t)(W - 1)))) { B.t = it; B.idx = idx; B.test = test; }
    }
}

// Literal per-lane emulation for the ambiguous sticky-inside case; out of line on purpose
// (rare, and its MinT[W]/group[W] registers must not count against the hot kernel).
template <int W, bool FMA>
__device__ __noinline__ uint32_t replay_inside(const float4 *sph, uint32_t n_pad, float ox, float oy, float oz, float dx, float dy, float dz) {
    const Hit h = intersect_lanes<(W == 1 ? 4 : W), FMA>(sph, n_pad, mk3(ox, oy, oz), mk3(dx, dy, dz));
    return h.inside ? 1u : 0u;
}

// Per-ray constants of the filter (one ray; a few dozen instructions per pass).
struct FilterRay {
    float p2ox, p2oy, p2oz;   // 2 O (exact)
    float dx, dy, dz;         // D' = D sqrt(2 - |D|^2)
    float od;                 // -(O . D')
    float oo;                 // |O|^2 deflated by 160u; -inf if D is not (nearly) unit or zero
};
__device__ __forceinline__ FilterRay filter_ray(float ox, float oy, float oz, float dx, float dy, float dz) {
    FilterRay f;
    f.p2ox = __fmul_rn(2.0f, ox); f.p2oy = __fmul_rn(2.0f, oy); f.p2oz = __fmul_rn(2.0f, oz);
    const float dd = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
    const float g = __fsqrt_rn(__fsub_rn(2.0f, dd));
    f.dx = __fmul_rn(dx, g); f.dy = __fmul_rn(dy, g); f.dz = __fmul_rn(dz, g);
    f.od = -__fmaf_rn(oz, f.dz, __fmaf_rn(oy, f.dy, __fmul_rn(ox, f.dx)));
    const float oo = __fmaf_rn(oz, oz, __fmaf_rn(oy, oy, __fmul_rn(ox, ox)));
    f.oo = __fmul_rn(oo, 1.0f - 160.0f * kFilterU);
    if (!(dd < 1.5f)) {   // exotic direction (never produced by the path tracer): every sphere is a candidate
        f.dx = 0.0f; f.dy = 0.0f; f.dz = 0.0f; f.od = 0.0f;
        f.oo = -__int_as_float(0x7f800000);
    }
    return f;
}

__device__ __noinline__ FilterRay filter_ray_call(float ox, float oy, float oz, float dx, float dy, float dz) {
    return filter_ray(ox, oy, oz, dx, dy, dz);
}

// mask |= bit  iff  za > ooa || zb > oob : FSETP, FSETP.OR, predicated LOP3.
__device__ __forceinline__ void filter_mark(uint32_t &mask, float za, float ooa, float zb, float oob, uint32_t bit) {
    asm("{\n\t.reg .pred p;\n\tsetp.gt.f32 p, %1, %2;\n\tsetp.gt.or.f32 p, %3, %4, p;\n\t@p or.b32 %0, %0, %5;\n\t}"
        : "+r"(mask) : "f"(za), "f"(ooa), "f"(zb), "f"(oob), "r"(bit));
}

// One sphere of the scan: {cx, cy, cz, -kappa} against the packed ray pair.
__device__ __forceinline__ void filter_sphere(uint32_t &mask, uint32_t bit, const float4 s, pk2 p2ox, pk2 p2oy, pk2 p2oz,
                                              pk2 dx, pk2 dy, pk2 dz, pk2 od, float ooa, float oob) {
    const pk2 cx = pk_make(s.x, s.x), cy = pk_make(s.y, s.y), cz = pk_make(s.z, s.z);
    const pk2 nq = pk_fma(cx, p2ox, pk_fma(cy, p2oy, pk_fma(cz, p2oz, pk_make(s.w, s.w))));   // 2 C.O - kappa
    const pk2 tt = pk_fma(cx, dx, pk_fma(cy, dy, pk_fma(cz, dz, od)));                        // (C - O).D'
    const pk2 z = pk_fma(tt, tt, nq);
    float za, zb;
    pk_split(z, za, zb);
    filter_mark(mask, za, ooa, zb, oob, bit);
}

typedef uint2 CandQueue[kPairsMaxWords];   // {candidate mask, first sphere} of the 32-sphere chunks with candidates

// ---- scan: every sphere, both rays of the thread, fused arithmetic, branch-free -----------------
// A table (or a tile of it) is scanned in chunks of 32 spheres; each chunk yields a 32-bit
// candidate mask (bit 31 - j = sphere j of the chunk) and the non-empty ones are appended to a
// per-thread queue.  The queue lives in LOCAL memory on purpose: ptxas cannot prove that a
// shared-memory store does not alias the sphere tables and would serialise the next LDS.128
// behind it.  One predicated STL.64 per chunk.  `count` is a multiple of 8 and only the last
// tile of a table may have count % 32 != 0; `first` is the table index of tile[0] (multiple of 32).
// SUB = spheres per unrolled loop body (16 or 32): the loop bodies share the SM's 6 KB L0 / 32 KB
// L1.5 instruction caches with the select and shade code.

// Per-pass constants of a general ray pair.  A parked slot must hold O = kIdleOrigin, D = 0.
struct ScanRays {
    pk2 p2ox, p2oy, p2oz, dx, dy, dz, od;
    float ooa, oob;
};
__device__ __forceinline__ ScanRays make_scan_rays(pk2 ox, pk2 oy, pk2 oz, pk2 dx, pk2 dy, pk2 dz) {
    float oxa, oxb, oya, oyb, oza, ozb, dxa, dxb, dya, dyb, dza, dzb;
    pk_split(ox, oxa, oxb); pk_split(oy, oya, oyb); pk_split(oz, oza, ozb);
    pk_split(dx, dxa, dxb); pk_split(dy, dya, dyb); pk_split(dz, dza, dzb);
    const FilterRay fa = filter_ray_call(oxa, oya, oza, dxa, dya, dza), fb = filter_ray_call(oxb, oyb, ozb, dxb, dyb, dzb);
    ScanRays r;
    r.p2ox = pk_make(fa.p2ox, fb.p2ox); r.p2oy = pk_make(fa.p2oy, fb.p2oy); r.p2oz = pk_make(fa.p2oz, fb.p2oz);
    r.dx = pk_make(fa.dx, fb.dx); r.dy = pk_make(fa.dy, fb.dy); r.dz = pk_make(fa.dz, fb.dz);
    r.od = pk_make(fa.od, fb.od);
    r.ooa = fa.oo; r.oob = fb.oo;
    return r;
}
// ---- level 1: slab scan -- 2 packed FMAs + 2 compares + 1 mask update per sphere per ray pair -----
// A sphere the ray hits lies within its radius of the ray's LINE, hence within its radius of any
// plane that contains the line.  Take the plane spanned by the ray and one fixed coordinate axis
// (the scene's `slab axis`, chosen at upload as the axis along which the centres spread least):
// its unit normal U = (-D_q, D_p) / |(D_p, D_q)| has only the two other components (p, q), so
//       a = C.U - O.U = fma(C_p, U_p, fma(C_q, U_q, -O.U)),   candidate  <=>  |a| <= r'
// costs 2 FFMA2 per sphere per ray pair (FSETP takes |a| as an operand modifier).  On the
// 1,024-sphere scene 1.8 % of the spheres pass per ray; the survivors go through the 7-FMA filter
// above (level 2, refine_queue) and what is left (~2 per ray pair) through the reference's
// arithmetic (select).  Every sphere is still tested against every ray.
// Soundness: a reference hit has |w - D tau| < r(1+2.7u) + 6.3u|w|, and
// (w - D tau).U* = w.U* for the real unit normal U*; U, O.U and a are evaluated with errors
// <= 7.4u(|C|+|O|) in total.  r' = r(1+32u) + 32u(|C| + Omax) + 1e-30 covers both as long as
// |O| <= Omax (the launch's bound on ray origins); a ray with a larger origin, a near-axis
// direction (|(D_p, D_q)|^2 < 1e-16) or a NaN takes U = 0, O.U = 0: every sphere is a candidate.
// Parked halves (O = kIdleOrigin) take -O.U = -inf: no candidates.  Padding has r' = -1e30 (kSlabPadRadius): the
// level-0 bound r' + wl * eps of srt_bundle.cuh stays negative for it whatever the camera.
constexpr float kSlabPadRadius = -1.0e30f;
struct SlabRay { float up, uq, nou; };
__device__ __forceinline__ SlabRay slab_ray(int axis, float ox, float oy, float oz, float dx, float dy, float dz, float omax2) {
    const float op = axis == 0 ? oy : (axis == 1 ? oz : ox), oq = axis == 0 ? oz : (axis == 1 ? ox : oy);
    const float dp = axis == 0 ? dy : (axis == 1 ? dz : dx), dq = axis == 0 ? dz : (axis == 1 ? dx : dy);
    const float h2 = __fmaf_rn(dq, dq, __fmul_rn(dp, dp));
    const float oo = __fmaf_rn(oz, oz, __fmaf_rn(oy, oy, __fmul_rn(ox, ox)));
    SlabRay r;
    r.up = 0.0f; r.uq = 0.0f; r.nou = 0.0f;
    if (ox == kIdleOrigin) {
        r.nou = -__int_as_float(0x7f800000);
    } else if (h2 >= 1.0e-16f && oo <= omax2) {
        const float h = __fsqrt_rn(h2);
        r.up = __fdiv_rn(-dq, h);
        r.uq = __fdiv_rn(dp, h);
        r.nou = -__fmaf_rn(oq, r.uq, __fmul_rn(op, r.up));
    }
    return r;
}
// r' of one sphere for a launch whose ray origins satisfy |O| <= omax.
__device__ __forceinline__ float slab_radius(float r2, float cx, float cy, float cz, float omax) {
    const float c = __fsqrt_rn(__fmaf_rn(cz, cz, __fmaf_rn(cy, cy, __fmul_rn(cx, cx))));
    return __fmaf_rn(32.0f * kFilterU, __fadd_rn(c, omax), __fmaf_rn(__fsqrt_rn(r2), 1.0f + 32.0f * kFilterU, 1.0e-30f));
}
__device__ __forceinline__ bool slab_test(float cp, float cq, float rr, const SlabRay &r) {
    return fabsf(__fmaf_rn(cp, r.up, __fmaf_rn(cq, r.uq, r.nou))) <= rr;
}

// Out-of-line copies for the kernels that build the constants of two rays per pass: one copy of the IEEE
// square root / divisions instead of two (instruction-cache footprint).
__device__ __noinline__ SlabRay slab_ray_call(int axis, float ox, float oy, float oz, float dx, float dy, float dz, float omax2) {
    return slab_ray(axis, ox, oy, oz, dx, dy, dz, omax2);
}
struct SlabRays { pk2 up, uq, nou; };
__device__ __forceinline__ SlabRays make_slab_rays(int axis, pk2 ox, pk2 oy, pk2 oz, pk2 dx, pk2 dy, pk2 dz, float omax2) {
    float oxa, oxb, oya, oyb, oza, ozb, dxa, dxb, dya, dyb, dza, dzb;
    pk_split(ox, oxa, oxb); pk_split(oy, oya, oyb); pk_split(oz, oza, ozb);
    pk_split(dx, dxa, dxb); pk_split(dy, dya, dyb); pk_split(dz, dza, dzb);
    const SlabRay a = slab_ray_call(axis, oxa, oya, oza, dxa, dya, dza, omax2), b = slab_ray_call(axis, oxb, oyb, ozb, dxb, dyb, dzb, omax2);
    SlabRays r;
    r.up = pk_make(a.up, b.up); r.uq = pk_make(a.uq, b.uq); r.nou = pk_make(a.nou, b.nou);
    return r;
}
// mask |= bit  iff  |a_a| <= rr || |a_b| <= rr
__device__ __forceinline__ void slab_sphere(uint32_t &mask, uint32_t bit, float cp, float cq, float rr, const SlabRays &r) {
    const pk2 a = pk_fma(pk_make(cp, cp), r.up, pk_fma(pk_make(cq, cq), r.uq, r.nou));
    float aa, ab;
    pk_split(a, aa, ab);
    asm("{\n\t.reg .pred p;\n\t.reg .f32 t, v;\n\tabs.f32 t, %1;\n\tabs.f32 v, %2;\n\tsetp.le.f32 p, t, %3;\n\tsetp.le.or.f32 p, v, %3, p;\n\t@p or.b32 %0, %0, %4;\n\t}"
        : "+r"(mask) : "f"(aa), "f"(ab), "f"(rr), "r"(bit));
}
// Slab table: per 4 spheres three float4 {p0..p3}, {q0..q3}, {r'0..r'3} (p, q = the two kept axes).
// tile = the float4 of sphere group first/4; count % 8 == 0; only the last tile may have count % 32 != 0.
constexpr uint32_t kSlabFloat4Per32 = 24;
__device__ __forceinline__ void slab_group(uint32_t &m, uint32_t bit0, const float4 *g, const SlabRays &r) {
    const float4 P = g[0], Q = g[1], R = g[2];
    slab_sphere(m, bit0, P.x, Q.x, R.x, r);
    slab_sphere(m, bit0 >> 1, P.y, Q.y, R.y, r);
    slab_sphere(m, bit0 >> 2, P.z, Q.z, R.z, r);
    slab_sphere(m, bit0 >> 3, P.w, Q.w, R.w, r);
}
__device__ __forceinline__ void scan_slab_tile(const float4 *__restrict__ tile, uint32_t count, uint32_t first,
                                               const SlabRays &r, CandQueue &queue, uint32_t &qn) {
    const uint32_t full = count >> 5;
#pragma unroll 1
    for (uint32_t c = 0; c < full; ++c) {
        const float4 *blk = tile + c * kSlabFloat4Per32;
        uint32_t m = 0;
#pragma unroll
        for (int g = 0; g < 8; ++g) slab_group(m, 0x80000000u >> (4 * g), blk + 3 * g, r);
        SRT_CHECK(24, qn < kPairsMaxWords);
        if (m) { queue[qn] = make_uint2(m, first + (c << 5)); qn += 1; }
    }
    if (count & 31u) {                       // tail: 2, 4 or 6 groups of 4
        const float4 *blk = tile + full * kSlabFloat4Per32;
        uint32_t m = 0;
#pragma unroll 1
        for (uint32_t g = 0; g < ((count & 31u) >> 2); ++g) slab_group(m, 0x80000000u >> (4 * g), blk + 3 * g, r);
        if (m) { queue[qn] = make_uint2(m, first + (full << 5)); qn += 1; }
    }
}

// ---- level 2: the 7-FMA filter on the slab survivors, per lane, in place -------------------------
// Every lane walks its own queue (flattened like select_update: the warp iterates
// max-over-lanes(survivors) times), drops the bits the filter rejects and compacts the queue
// (select_update relies on non-zero masks).
template <class Fetch>
__device__ __forceinline__ void refine_queue(const Fetch &fetch_scan, CandQueue &queue, uint32_t &qn, const ScanRays &r) {
    // The queue lives in local memory and mostly misses the small L1 left beside 200 KB of shared memory (ncu: the
    // __clz behind the entry load held 13 % of all stall samples), so entries are fetched two ahead of their use.
    uint32_t qi = 0, qo = 0, m = 0, keep = 0, first = 0;
    uint2 n0 = make_uint2(0u, 0u), n1 = n0;
    if (qn > 0u) n0 = queue[0];
    if (qn > 1u) n1 = queue[1];
    for (;;) {
        if (m == 0) {
            if (keep) { queue[qo] = make_uint2(keep, first); qo += 1; }      // compacted in place: qo <= qi - 1 < prefetched entries
            if (qi == qn) break;
            m = n0.x; first = n0.y; keep = 0;                                // masks in the queue are never 0
            n0 = n1; qi += 1;
            if (qi + 1u < qn) n1 = queue[qi + 1u];
        }
        const uint32_t j = (uint32_t)__clz((int)m);
        const uint32_t bit = 0x80000000u >> j;
        m &= ~bit;
        SRT_CHECK(25, first + j < 65536u * 4u && qi <= qn && qn <= kPairsMaxWords);
        filter_sphere(keep, bit, fetch_scan(first + j), r.p2ox, r.p2oy, r.p2oz, r.dx, r.dy, r.dz, r.od, r.ooa, r.oob);
    }
    qn = qo;
}
struct FetchScanResident { const float4 *scan; __device__ __forceinline__ float4 operator()(uint32_t idx) const { return scan[idx]; } };
struct FetchScanGlobal { const float4 *scan; __device__ __forceinline__ float4 operator()(uint32_t idx) const { return __ldg(scan + idx); } };

// ---- select: the reference's arithmetic on the queued candidates, in index order ----------------
// Sphere fetchers: {cx, cy, cz, r2} of sphere idx with the reference's operands.
struct FetchResident {              // centres from the resident filter table, r2 = R*R (main.cpp:406) beside it
    const float4 *scan;
    const float *r2tab;
    __device__ __forceinline__ float4 operator()(uint32_t idx) const { float4 s = scan[idx]; s.w = r2tab[idx]; return s; }
};
struct FetchGlobal {                // streamed scenes: straight from the {cx,cy,cz,r2} table in global memory (L2)
    const float4 *sph;
    __device__ __forceinline__ float4 operator()(uint32_t idx) const { return __ldg(sph + idx); }
};

// Folds the queue into the running closest-hit state of the two rays.  Flattened over the
// thread's own candidates: the warp iterates max-over-lanes(candidates) times, not once per
// chunk that has a candidate in any lane.
template <int W, bool FMA, class Fetch>
__device__ __forceinline__ void select_update(const Fetch &fetch, uint32_t n_real, const CandQueue &queue, uint32_t qn,
                                              pk2 ox, pk2 oy, pk2 oz, pk2 dx, pk2 dy, pk2 dz,
                                              float neg_one_f, float one_f, Best &b0, Best &b1) {
    const pk2 neg_one = pk_make(neg_one_f, neg_one_f), one = pk_make(one_f, one_f);
    uint32_t qi = 0, m = 0, first = 0;
    uint2 n0 = make_uint2(0u, 0u);
    if (qn > 0u) n0 = queue[0];                                              // entries are fetched one ahead (see refine_queue)
    for (;;) {
        if (m == 0) {
            if (qi == qn) break;
            m = n0.x; first = n0.y; qi += 1;
            if (qi < qn) n0 = queue[qi];
        }
        const uint32_t j = (uint32_t)__clz((int)m);
        m &= ~(0x80000000u >> j);
        const uint32_t idx = first + j;
        const float4 s = fetch(idx);
        const pk2 scx = pk_sub(pk_make(s.x, s.x), ox);                         // main.cpp:401
        const pk2 scy = pk_sub(pk_make(s.y, s.y), oy);
        const pk2 scz = pk_sub(pk_make(s.z, s.z), oz);
        const pk2 t2 = pk_dot<FMA>(scx, scy, scz, dx, dy, dz, one);            // :402
        const pk2 vx = pk_fma(pk_mul(dx, t2), neg_one, scx);                   // :403,407
        const pk2 vy = pk_fma(pk_mul(dy, t2), neg_one, scy);
        const pk2 vz = pk_fma(pk_mul(dz, t2), neg_one, scz);
        const pk2 dd = pk_dot<FMA>(vx, vy, vz, vx, vy, vz, one);
        float d2a, d2b, ta, tb;
        pk_split(dd, d2a, d2b);
        pk_split(t2, ta, tb);
        if (W == 1) {
            // `d2 > r2 -> continue` (:557) lets NaN through, so padding is excluded by index.
            if (!(d2a > s.w) && idx < n_real) best_update<W>(b0, ta, d2a, s.w, idx);
            if (!(d2b > s.w) && idx < n_real) best_update<W>(b1, tb, d2b, s.w, idx);
        } else {
            if (d2a < s.w) best_update<W>(b0, ta, d2a, s.w, idx);              // :409
            if (d2b < s.w) best_update<W>(b1, tb, d2b, s.w, idx);
        }
    }
}

// Closest-hit records from the final state; resolves the sticky-inside ambiguity.
//   sph_global : global memory, n_pad x {cx, cy, cz, r2}, only read by the rare replay_inside
template <int W, bool FMA>
__device__ __forceinline__ void select_finish(const float4 *__restrict__ sph_global, uint32_t n_pad, const Best &b0, const Best &b1,
                                              pk2 ox, pk2 oy, pk2 oz, pk2 dx, pk2 dy, pk2 dz, Hit &h0, Hit &h1) {
    h0.t = b0.t; h0.hit = !(b0.t == kMax); h0.prim = h0.hit ? b0.idx : 0xFFFFFFFFu; h0.inside = b0.test != 0;
    h1.t = b1.t; h1.hit = !(b1.t == kMax); h1.prim = h1.hit ? b1.idx : 0xFFFFFFFFu; h1.inside = b1.test != 0;
    if (W != 1) {
        const bool amb0 = h0.hit && !b0.test && ((b0.test_lanes >> (b0.idx & (uint32_t)(W - 1))) & 1u);
        const bool amb1 = h1.hit && !b1.test && ((b1.test_lanes >> (b1.idx & (uint32_t)(W - 1))) & 1u);
        if (amb0 || amb1) {
            float oxa, oxb, oya, oyb, oza, ozb, dxa, dxb, dya, dyb, dza, dzb;
            pk_split(ox, oxa, oxb); pk_split(oy, oya, oyb); pk_split(oz, oza, ozb);
            pk_split(dx, dxa, dxb); pk_split(dy, dya, dyb); pk_split(dz, dza, dzb);
            if (amb0) h0.inside = replay_inside<W, FMA>(sph_global, n_pad, oxa, oya, oza, dxa, dya, dza) != 0;
            if (amb1) h1.inside = replay_inside<W, FMA>(sph_global, n_pad, oxb, oyb, ozb, dxb, dyb, dzb) != 0;
        }
    }
}

} // namespace srt
