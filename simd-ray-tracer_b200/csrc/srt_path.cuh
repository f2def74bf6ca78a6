// srt_path.cuh -- the per-path steps shared by every kernel of the backend: primary ray (A1),
// miss / sky (A5), shade + scatter (A6), running mean (A8).  Citations: reference file:line.
#pragma once
#include "../../include/srt_backend.h"
#include "srt_device.cuh"

namespace srt {

constexpr uint32_t kInvalidPixel = 0xFFFFFFFFu;

// Work item i -> pixel: 8x4 pixel blocks, row-major over blocks, so that a fresh warp
// covers a compact screen patch (coherent primary rays: the rare hit branch and the
// material branch diverge less).  Returns false for items outside the image.
// Image partition (multi-device, srt_set_partition): a handle renders every n_parts-th row of blocks -- the
// reference hands 32x32 tiles to its worker threads (main.cpp:9, 362-368, 824-827); here the unit is a row of
// 8x4 blocks, interleaved so that sky rows and ground rows spread evenly over the devices.
__device__ __forceinline__ bool item_to_pixel(const KernelParams &P, uint32_t item, uint32_t &px, uint32_t &py) {
    const uint32_t blk = item >> 5, in = item & 31u;
    const uint32_t by = magic_div(blk, P.div_bx), bx = blk - by * P.blocks_x;
    px = bx * 8u + (in & 7u);
    py = (by * P.n_parts + P.part) * 4u + (in >> 3);
    return px < P.width && py < P.height;
}

// Launch epilogue of every render kernel: the warp's ray count (main.cpp:390) joins the launch total; the LAST CTA
// to leave publishes the total to pinned host memory and re-arms the launch counters, so a frame costs the host one
// kernel launch -- no memset, no copy (the kick replaces WorkQueueStart, main.cpp:852).  All threads must call it.
__device__ __forceinline__ void finish_launch(const KernelParams &P, unsigned long long rays) {
    for (int off = 16; off > 0; off >>= 1) rays += __shfl_down_sync(0xFFFFFFFFu, rays, off);
    if ((threadIdx.x & 31u) == 0 && rays) atomicAdd(P.rays, rays);
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(P.ticket, 1u) == gridDim.x - 1u) {
            __threadfence();
            const unsigned long long total = atomicExch(P.rays, 0ull);
            *P.work_counter = 0u;
            *P.ticket = 0u;
            *(volatile unsigned long long *)P.host_rays = total;
            __threadfence_system();
        }
    }
}

// A1 primary ray, main.cpp:375-385.  Stream contract: the path (sample k,
// pixel) owns the RNG stream scramble(stream_base + k*W*H + pix); jitter x is drawn first.
template <bool OUTLINE = false>
__device__ __forceinline__ void primary_ray(const KernelParams &P, uint32_t px, uint32_t py, uint32_t pix, uint32_t k,
                                            uint64_t &rng, f3 &o, f3 &d) {
    rng = stream_seed(P.stream_base + (uint64_t)k * ((uint64_t)P.width * P.height) + pix);
    const float jx = rand_float(rng, kInvU32Range1, -0.5f);
    const float jy = rand_float(rng, kInvU32Range1, -0.5f);
    const float film_x = __fadd_rn(-1.0f, __fdiv_rn(__fmul_rn(__fadd_rn((float)px, jx), 2.0f), (float)P.width));
    const float film_y = __fadd_rn(-1.0f, __fdiv_rn(__fmul_rn(__fadd_rn((float)py, jy), 2.0f), (float)P.height));
    const float ax = __fmul_rn(__fmul_rn(film_x, P.cam.film_w), 0.5f);
    const float ay = __fmul_rn(__fmul_rn(film_y, P.cam.film_h), 0.5f);
    const f3 cam_x = mk3(P.cam.x[0], P.cam.x[1], P.cam.x[2]);
    const f3 cam_y = mk3(P.cam.y[0], P.cam.y[1], P.cam.y[2]);
    const f3 fc = mk3(P.cam.film_center[0], P.cam.film_center[1], P.cam.film_center[2]);
    const f3 film_p = add3(add3(fc, scale3(ax, cam_x)), scale3(ay, cam_y));
    o = mk3(P.cam.pos[0], P.cam.pos[1], P.cam.pos[2]);
    d = normalize3_sel<OUTLINE>(sub3(film_p, o));
}

// A5 + A6 for one path segment whose closest hit is `h` (`s` = winner's {cx,cy,cz,r2}, read
// by the caller from whatever layout its kernel keeps; ignored on a miss).  Updates the ray,
// attenuation, radiance, RNG and bounce count; returns true when the path ends.
template <bool OUTLINE = false>
__device__ __forceinline__ bool shade_scatter(const KernelParams &P, const Hit &h, const float4 s, f3 &o, f3 &d,
                                              f3 &atten, f3 &out, uint64_t &rng, uint32_t &bounce) {
    if (!h.hit) {
        if (P.use_sky) {                                              // A5 main.cpp:433-440
            const float a = __fmul_rn(__fadd_rn(d.y, 1.0f), 0.5f);
            const float ia = __fsub_rn(1.0f, a);
            const f3 sky = add3(scale3(ia, mk3(1.0f, 1.0f, 1.0f)), scale3(a, mk3(0.5f, 0.7f, 1.0f)));
            out = add3(out, mul3(sky, atten));
        }
        return true;
    }
    // ---- A6 shade + scatter, main.cpp:443-481 -------------------------------------------------
    const DevMaterial *M = &P.materials[h.prim];
    const float4 cs = __ldg(&M->color_spec);
    const float4 ei = __ldg(&M->emis_ior);
    const float4 dv = __ldg(&M->derived);
    out = add3(out, mul3(mk3(ei.x, ei.y, ei.z), atten));
    atten = mul3(atten, mk3(cs.x, cs.y, cs.z));
    // winner's HitNormal / NextRayOrigin, recomputed with the ops of main.cpp:423-429
    const f3 sc = sub3(mk3(s.x, s.y, s.z), o);
    const f3 ip = scale3(h.t, d);
    o = add3(o, ip);
    f3 n = normalize3_sel<OUTLINE>(sub3(ip, sc));
    const f3 pure = sub3(d, scale3(__fmul_rn(2.0f, dot3(d, n)), n));   // :453
    if (h.inside) n = neg3(n);
    if (ei.w == 0.0f) {
        const float rx = rand_float(rng, kInvU32Range2, -1.0f);
        const float ry = rand_float(rng, kInvU32Range2, -1.0f);
        const float rz = rand_float(rng, kInvU32Range2, -1.0f);
        const f3 r = normalize3_sel<OUTLINE>(mk3(rx, ry, rz));                      // :461
        const f3 rb = add3(n, r);
        d = normalize3_sel<OUTLINE>(add3(scale3(dv.x, rb), scale3(cs.w, pure)));    // :463-464
    } else {
        const float eta = h.inside ? ei.w : dv.y;                      // :466
        const float cd = dot3(neg3(d), n);
        const float c = (cd < 1.0f) ? cd : 1.0f;                       // _mm_min_ss :468
        const float cc = __fmul_rn(c, c);
        const float sn = __fsqrt_rn(__double2float_rn(__dsub_rn(1.0, (double)cc)));   // :469
        const bool cant = __fmul_rn(eta, sn) > 1.0f;                   // :471
        const f3 perp = scale3(eta, add3(d, scale3(c, n)));            // :472
        const float kk = -__fsqrt_rn(fabsf(__fsub_rn(1.0f, dot3(perp, perp))));   // :473
        const f3 refr = normalize3_sel<OUTLINE>(add3(perp, scale3(kk, n)));         // :474
        bool reflect = cant;                                            // :476 (|| short-circuit)
        if (!cant) reflect = reflectance(c, eta) > rand_float(rng, kInvU32Range1, 0.0f);
        d = (reflect && !h.inside) ? pure : refr;
    }
    bounce += 1;
    return bounce == kMaxBounce;
}

// A8: fold the finished sample k into the pixel's accumulator (main.cpp:484-489 with
// k == PreviousRayCount), or add it to the plain sum of the multi-GPU path.
__device__ __forceinline__ f3 accumulate_sample(uint32_t accum_mode, f3 mean, f3 out, uint32_t k) {
    if (accum_mode == SRT_ACCUM_MEAN) {
        const float total = (float)(k + 1u);
        const float wa = __fdiv_rn(1.0f, total);
        const float wb = __fdiv_rn((float)k, total);
        return add3(scale3(wa, out), scale3(wb, mean));
    }
    return add3(mean, out);
}

} // namespace srt
