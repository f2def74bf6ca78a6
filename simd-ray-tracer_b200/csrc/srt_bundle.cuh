// srt_bundle.cuh -- level 0 of the sphere filter for PRIMARY rays: one slab test per (pixel, sphere)
// for ALL samples of the pixel (sm_100a, pool kernel).
//
// Why.  The level-1 slab scan (srt_pairs.cuh) is 2/3 of the pool kernel's issue slots and sits on its
// floor (2 FFMA2 + 2 FSETP + 1 IADD + 0.75 LDS per sphere per ray pair).  45 % of the rays of the
// 4K / 256-spp job are primary rays, and the primary rays of one pixel differ only by the sub-pixel
// jitter of main.cpp:378-381: they leave the same origin inside a cone of half-angle ~1e-4.  When the
// 64 rays of a primary warp pass are 64 SAMPLES OF ONE PIXEL, the slab test can be done once for the
// whole bundle with a radius inflated by the bundle's width -- and the 32 lanes can share the table:
// each lane tests 1/32 of the spheres.  Every sphere is still tested against every ray (there is no
// spatial structure over the scene); what is shared is the arithmetic of the miss decision.
//
// Test.  Let F(j) = film point(j) - O be the un-normalised direction of the sample with jitter j and
// F(0) the one through the pixel centre; p, q the two kept axes (slab axis dropped).  With U_c the unit
// normal of the plane spanned by F(0) and the slab axis,
//       a = C.U_c - O.U_c,     candidate  <=>  |a| <= r' + (|C_p - O_p| + |C_q - O_q|) * eps,
//       eps = 1.01 rho / |F(0)_pq| + 1e-5   (the 1e-5 also absorbs the 2^-21 error of the fast normalisation),
// where r' is the level-1 radius (slab_radius) and rho >= |F(j)_pq - F(0)_pq| for every jitter in
// [-0.5, 0.5]^2 (host, double, plus the fp32 rounding of the film-point evaluation).
// Soundness: a reference hit of sample j has |w.U*(j)| < r(1+2.7u) + 6.3u|w| (5.3, (3)),
// U*(j) being the unit normal for the fp32 direction D(j).  The projections of D(j) and F(0) on the pq
// plane enclose an angle theta with sin(theta) <= rho / |F(0)_pq| (+ u for the normalisation of D), so
// |U*(j) - U_c| <= theta <= 1.0005 rho/|F(0)_pq| + u for rho/|F(0)_pq| <= 0.05, and
// |w.U_c| <= |w.U*(j)| + |w_pq| |U*(j) - U_c|.  The evaluation errors of a and U_c are the ones r' was
// inflated for.  Pixels with rho/|F(0)_pq| > 0.05 (rays nearly parallel to the slab axis), a degenerate
// film vector or a non-finite constant are not bundled: their warp pass runs the per-ray slab scan.
// srt_filter_audit checks the superset property for every (primary ray, sphere) pair on the device.
#pragma once
#include "srt_pairs.cuh"
#include "srt_path.cuh"

namespace srt {

struct BundlePixel { float up, uq, nou, eps; };   // eps < 0: this pixel's samples must take the per-ray scan

// Nothing here needs to be exact: (up, uq) only has to be a unit vector to 1e-6 (r' carries a (1+32u) factor
// and eps a 1e-5 slack for it) and rho an upper bound, so the film point uses a multiplication by 2/W and the
// normalisation one MUFU.RSQ (<= 2 ulp) -- no IEEE division or square root (1.5 KB of SASS less in the hot code).
__device__ __forceinline__ BundlePixel bundle_pixel(const KernelParams &P, uint32_t px, uint32_t py) {
    // film point through the pixel centre: main.cpp:378-384 with zero jitter
    const float ax = __fmul_rn(__fmul_rn(__fmaf_rn((float)px, P.two_over_w, -1.0f), P.cam.film_w), 0.5f);
    const float ay = __fmul_rn(__fmul_rn(__fmaf_rn((float)py, P.two_over_h, -1.0f), P.cam.film_h), 0.5f);
    const int axis = P.slab_axis;
    const int ip = axis == 0 ? 1 : (axis == 1 ? 2 : 0), iq = axis == 0 ? 2 : (axis == 1 ? 0 : 1);
    const float fx = __fsub_rn(__fadd_rn(__fadd_rn(P.cam.film_center[0], __fmul_rn(ax, P.cam.x[0])), __fmul_rn(ay, P.cam.y[0])), P.cam.pos[0]);
    const float fy = __fsub_rn(__fadd_rn(__fadd_rn(P.cam.film_center[1], __fmul_rn(ax, P.cam.x[1])), __fmul_rn(ay, P.cam.y[1])), P.cam.pos[1]);
    const float fz = __fsub_rn(__fadd_rn(__fadd_rn(P.cam.film_center[2], __fmul_rn(ax, P.cam.x[2])), __fmul_rn(ay, P.cam.y[2])), P.cam.pos[2]);
    const float fp = ip == 0 ? fx : (ip == 1 ? fy : fz), fq = iq == 0 ? fx : (iq == 1 ? fy : fz);
    const float h2 = __fmaf_rn(fq, fq, __fmul_rn(fp, fp));
    const float ff = __fmaf_rn(fz, fz, __fmaf_rn(fy, fy, __fmul_rn(fx, fx)));
    const float rs = rsqrtf(h2);
    const float x = __fmul_rn(P.bundle_rho, rs);
    BundlePixel b;
    b.up = __fmul_rn(-fq, rs);
    b.uq = __fmul_rn(fp, rs);
    b.nou = -__fmaf_rn(P.cam.pos[iq], b.uq, __fmul_rn(P.cam.pos[ip], b.up));
    b.eps = __fmaf_rn(1.01f, x, 1.0e-5f);
    if (!(h2 >= 1.0e-12f && ff >= 0.01f && ff < 1.0e12f && x <= 0.05f && P.bundle_rho > 0.0f)) b.eps = -1.0f;   // NaN lands here too
    return b;
}

// One sphere of the slab table against a pixel bundle; (op, oq) = camera position on the kept axes.
__device__ __forceinline__ bool bundle_test(float cp, float cq, float rr, float op, float oq, const BundlePixel &b) {
    const float a = __fmaf_rn(cp, b.up, __fmaf_rn(cq, b.uq, b.nou));
    const float wl = __fadd_rn(fabsf(__fsub_rn(cp, op)), fabsf(__fsub_rn(cq, oq)));
    return fabsf(a) <= __fmaf_rn(wl, b.eps, rr);
}

// Warp-cooperative level-0 scan of spheres [first, first + count) (first % 128 == 0, count % 8 == 0) of a
// slab table (3 float4 per 4 spheres; `tab` = float4 of sphere 0) for the pixel bundles held by the lanes
// in `heads0` (constants b0) and `heads1` (constants b1).  Lane l tests sphere group l of every 128-sphere
// round; the 4-bit results of 8 neighbouring lanes are OR-reduced into the 32-sphere chunk masks of the
// candidate queue, and EVERY lane appends the same entries to its own queue (refine / select then run on
// warp-uniform sphere indices: broadcast LDS, no divergence).
template <bool GLOBAL>
__device__ __forceinline__ void bundle_scan(const float4 *__restrict__ tab, uint32_t first, uint32_t count, float op, float oq,
                                            unsigned heads0, unsigned heads1, const BundlePixel &b0, const BundlePixel &b1,
                                            unsigned lane, CandQueue &queue, uint32_t &qn) {
    const uint32_t g_end = (first + count) >> 2;
#pragma unroll 1
    for (uint32_t g0 = first >> 2; g0 < g_end; g0 += 32u) {
        const uint32_t grp = g0 + lane;
        const bool in = grp < g_end;
        float4 Pp = make_float4(0.f, 0.f, 0.f, 0.f), Qq = Pp, Rr = make_float4(-1.0e30f, -1.0e30f, -1.0e30f, -1.0e30f);
        if (in) {
            const float4 *g = tab + 3u * grp;
            if (GLOBAL) { Pp = __ldg(g); Qq = __ldg(g + 1); Rr = __ldg(g + 2); }
            else { Pp = g[0]; Qq = g[1]; Rr = g[2]; }
        }
        uint32_t nib = 0;
#pragma unroll 1
        for (int half = 0; half < 2; ++half) {
            unsigned hm = half ? heads1 : heads0;
            while (hm) {
                const int src = __ffs(hm) - 1;
                hm &= hm - 1u;
                BundlePixel b;
                b.up = __shfl_sync(0xFFFFFFFFu, half ? b1.up : b0.up, src);
                b.uq = __shfl_sync(0xFFFFFFFFu, half ? b1.uq : b0.uq, src);
                b.nou = __shfl_sync(0xFFFFFFFFu, half ? b1.nou : b0.nou, src);
                b.eps = __shfl_sync(0xFFFFFFFFu, half ? b1.eps : b0.eps, src);
                if (bundle_test(Pp.x, Qq.x, Rr.x, op, oq, b)) nib |= 8u;
                if (bundle_test(Pp.y, Qq.y, Rr.y, op, oq, b)) nib |= 4u;
                if (bundle_test(Pp.z, Qq.z, Rr.z, op, oq, b)) nib |= 2u;
                if (bundle_test(Pp.w, Qq.w, Rr.w, op, oq, b)) nib |= 1u;
            }
        }
        // Lanes past the end of the table (n_pad % 128 != 0) ran the test on {0, 0, -1}: |nou| <= wl * eps - 1 CAN hold for a
        // far camera and a coarse image (eps ~ 0.05, |O| ~ 1e2), and a set bit would name a sphere index >= n_pad.
        if (!in) nib = 0u;
        uint32_t word = nib << (28u - 4u * (lane & 7u));          // bit 31 - j = sphere j of the 32-sphere chunk
        word |= __shfl_xor_sync(0xFFFFFFFFu, word, 1);
        word |= __shfl_xor_sync(0xFFFFFFFFu, word, 2);
        word |= __shfl_xor_sync(0xFFFFFFFFu, word, 4);
        unsigned nz = __ballot_sync(0xFFFFFFFFu, word != 0u) & 0x01010101u;   // lanes 0, 8, 16, 24: chunks in index order
        while (nz) {
            const int src = __ffs(nz) - 1;
            nz &= nz - 1u;
            const uint32_t m = __shfl_sync(0xFFFFFFFFu, word, src);
            SRT_CHECK(26, qn < kPairsMaxWords);
            queue[qn] = make_uint2(m, (g0 + (uint32_t)src) << 2);
            qn += 1;
        }
    }
}

// ---- work items of bundle mode --------------------------------------------------------------------
// item -> (j, pixel chunk): j = item % S is the sample within the chunk, S = 2^bundle_log2.  Pixel chunks are
// ordered band by band (a band = a few whole rows of 8x4 pixel blocks, srt_path.cuh), inside a band
// chunk-major: (band b, chunk c, pixel item).  All chunks of a band are handed out before the next band
// starts, so the accumulator, the staging rows and the chunk counters of the pixels in flight stay in L2,
// and the chunks of one pixel are band_items * S work items apart (far more than the paths in flight).
struct BundleItem { uint32_t row, n, px, py; };   // staging row (= pixel item within its band), sample ordinal, pixel
__device__ __forceinline__ bool bundle_item(const KernelParams &P, uint32_t item, BundleItem &it) {
    const uint32_t j = item & ((1u << P.bundle_log2) - 1u), pc = item >> P.bundle_log2;      // pc < 2^30
    const uint32_t b = magic_div(pc, P.div_full), r = pc - b * (P.band_items * P.n_chunks);
    const bool last = b + 1u == P.n_bands;
    const uint32_t size = last ? P.band_last : P.band_items;
    MagicDiv dv = P.div_band;
    if (last) dv = P.div_last;
    const uint32_t c = magic_div(r, dv);
    it.row = r - c * size;
    it.n = (c << P.bundle_log2) + j;
    return it.n < P.n_samples && item_to_pixel(P, b * P.band_items + it.row, it.px, it.py);
}

} // namespace srt
