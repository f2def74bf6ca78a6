// srt_backend.cu -- CUDA kernels (sm_100a) and the C ABI of include/srt_backend.h.
//
// Replaces the reference's per-tile callbacks RenderTile / RenderTileScalar
// (main.cpp:348-640) and the work-queue kick that runs them (main.cpp:852).
// No CPU fallback: every entry point fails with SRT_ERR_NO_DEVICE / SRT_ERR_CUDA
// when there is no usable device.
#include "../../include/srt_backend.h"
#include "srt_device.cuh"
#include "srt_packed.cuh"
#include "srt_pairs.cuh"
#include "srt_path.cuh"
#include "srt_pool.cuh"
#include "srt_bundle.cuh"
#include "srt_stream.cuh"
#include "srt_comm.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

namespace srt {

constexpr int kBlockThreads = 256;

// ------------------------------------------------------------------------------------------
// render_paths: the whole per-pixel pipeline A1..A9, ONE path per
// thread.  Used for small scenes (shading-bound) and as the diagnostic baseline.
//
// Persistent CTAs, PATH REGENERATION: a thread whose path ended (miss or 5th bounce)
// immediately starts the next sample of its pixel, or pulls the next pixel from a global
// counter (warp-aggregated atomic), so all 32 lanes of a warp enter the brute-force
// intersection loop together no matter how many bounces each path takes.  A thread owns a
// pixel for all n_samples of the launch: the running mean lives in registers and the
// framebuffers are touched once per pixel per launch (16 B read when first_sample > 0,
// 16 B + 4 B written).  The scene SoA sits in shared memory and is read with warp-uniform
// LDS.128 broadcasts.
// ------------------------------------------------------------------------------------------
template <int W, bool FMA, bool PACKED, int PAIRS = 4, int MINB = 1>
__global__ void __launch_bounds__(kBlockThreads, MINB)
render_paths(const __grid_constant__ KernelParams P) {
    extern __shared__ float4 s_spheres[];
    for (uint32_t i = threadIdx.x; i < P.n_pad; i += blockDim.x) s_spheres[i] = P.spheres[i];
    __syncthreads();

    const unsigned lane = threadIdx.x & 31u;
    const uint32_t k_end = P.first_sample + P.n_samples * P.sample_stride;

    uint32_t pix = kInvalidPixel, px = 0, py = 0;
    uint32_t k = k_end;            // current sample index; k == k_end: pixel finished
    bool active = false, done = false;
    uint32_t bounce = 0;
    uint64_t rng = 0;
    unsigned long long rays = 0;
    f3 o = mk3(0, 0, 0), d = mk3(0, 0, 0), atten = mk3(1, 1, 1), out = mk3(0, 0, 0);
    f3 mean = mk3(0, 0, 0);

    for (;;) {
        // ---- regenerate: finish pixels, fetch new ones, start new paths -------------------
        // Starting a path costs ~150 issue slots for the whole warp however few lanes need
        // it, so idle lanes wait until `regen_threshold` of them have gathered (1 for large
        // scenes, where an idle lane in the intersection loop is the bigger waste).
        const unsigned idle_mask = __ballot_sync(0xFFFFFFFFu, !active && !done);
        const unsigned active_mask = __ballot_sync(0xFFFFFFFFu, active);
        if (idle_mask != 0u && ((uint32_t)__popc(idle_mask) >= P.regen_threshold || active_mask == 0u)) {
            bool need = !active && !done && (k == k_end);
            if (need && pix != kInvalidPixel) {
                // A8 result + A9 tonemap/pack, main.cpp:484-492
                P.accum[pix] = make_float4(mean.x, mean.y, mean.z, 1.0f);
                if (P.accum_mode == SRT_ACCUM_MEAN) P.rgba[pix] = pack_rgba8(mean.x, mean.y, mean.z);
                pix = kInvalidPixel;
            }
            unsigned m = __ballot_sync(0xFFFFFFFFu, need);
            while (m) {
                const unsigned leader = __ffs(m) - 1;
                unsigned base = 0;
                if (lane == leader) base = atomicAdd(P.work_counter, __popc(m));
                base = __shfl_sync(0xFFFFFFFFu, base, leader);
                if (need) {
                    const uint32_t item = base + __popc(m & ((1u << lane) - 1u));
                    if (item >= P.total_items) { done = true; need = false; }
                    else if (item_to_pixel(P, item, px, py)) {
                        pix = py * P.width + px;
                        k = P.first_sample;
                        need = false;
                        if (P.first_sample > 0 || P.accum_mode == SRT_ACCUM_SUM) {
                            const float4 prev = P.accum[pix];
                            mean = mk3(prev.x, prev.y, prev.z);
                        } else {
                            mean = mk3(0.0f, 0.0f, 0.0f);   // PreviousColor * (0/1) with a zeroed arena
                        }
                    }
                }
                m = __ballot_sync(0xFFFFFFFFu, need);
            }
            if (!active && !done) {
                primary_ray(P, px, py, pix, k, rng, o, d);
                atten = mk3(1.0f, 1.0f, 1.0f);
                out = mk3(0.0f, 0.0f, 0.0f);
                bounce = 0;
                active = true;
            }
        }
        if (__all_sync(0xFFFFFFFFu, done)) break;

        if (active) {
            // ---- A3/A4 brute-force intersection over ALL spheres, main.cpp:399-455 --------
            rays += 1;                                                    // main.cpp:390
            Hit h;
            if (W == 1) h = intersect_scalar(s_spheres, P.n_pad, P.n_spheres, o, d);
            else if (PACKED) h = intersect_lanes_packed<(W == 1 ? 4 : W), FMA, PAIRS>(s_spheres, P.n_pad, o, d, P.neg_one, P.one);
            else h = intersect_lanes<(W == 1 ? 4 : W), FMA>(s_spheres, P.n_pad, o, d);
            if (bounce == 0 && P.tap_t != nullptr && k == P.first_sample) {
                P.tap_t[pix] = h.t;
                P.tap_prim[pix] = h.prim;
            }
            float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
            if (h.hit) s = PACKED ? paired_sphere(s_spheres, h.prim) : s_spheres[h.prim];
            if (shade_scatter(P, h, s, o, d, atten, out, rng, bounce)) {
                mean = accumulate_sample(P.accum_mode, mean, out, k);
                k += P.sample_stride;
                active = false;
            }
        }
    }

    finish_launch(P, rays);
}

// Path state of one slot of the pool kernel, parked in shared memory (field-major: conflict-free).
enum : int { S_OX, S_OY, S_OZ, S_DX, S_DY, S_DZ, S_AX, S_AY, S_AZ, S_LX, S_LY, S_LZ,
             S_RNG_LO, S_RNG_HI, S_PIX, S_K, S_PXY, S_FLAGS, S_KEND, S_MX, S_MY, S_MZ, S_COUNT };
constexpr int kBundleFields = S_COUNT - 3;   // bundle mode keeps no running mean in the slot (S_MX..S_MZ): 23 KB more L1 for the queues
#ifndef SRT_POOL_SPARE
#define SRT_POOL_SPARE 2
#endif
constexpr uint32_t kPoolSpareBatches = SRT_POOL_SPARE;   // path slots per CTA = 64 x (warps + spare)
constexpr uint32_t kFlagBounceMask = 7u, kFlagPending = 32u;

// ------------------------------------------------------------------------------------------
// render_pool: TWO paths per thread in the halves of the f32x2 instructions (srt_pairs.cuh), path slots
// owned by the CTA (srt_pool.cuh).  A warp pass is either 64 slots that need a NEW PRIMARY ray (fetch work
// items, regenerate; in bundle mode all samples of one pixel, srt_bundle.cuh) or 64 slots with a SECONDARY
// ray ready; both run level 1 (slab test of every sphere) -> level 2 (7-FMA filter) -> exact select -> shade.
// ------------------------------------------------------------------------------------------
// Slab table entry of sphere i (srt_pairs.cuh): kept-axis coordinates and the inflated radius.
__device__ __forceinline__ void slab_entry(float *table, uint32_t i, float4 c, float r2, bool real, int axis, float omax) {
    const float cp = axis == 0 ? c.y : (axis == 1 ? c.z : c.x), cq = axis == 0 ? c.z : (axis == 1 ? c.x : c.y);
    float *g = table + (i >> 2) * 12u + (i & 3u);
    g[0] = cp; g[4] = cq; g[8] = real ? slab_radius(r2, c.x, c.y, c.z, omax) : kSlabPadRadius;
}
// Streamed scenes: the slab table is built once per launch in global memory.
__global__ void build_slab_kernel(const float4 *scan, const float4 *spheres, float *table, uint32_t n_pad, uint32_t n_real, int axis, float omax) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pad) return;
    slab_entry(table, i, scan[i], spheres[i].w, i < n_real, axis, omax);
}

// One pass of a warp over the streamed slab table (srt_stream.cuh): TMA bulk copies into the
// warp's ring of stages, level-1 scan of each chunk; refine (level 2) + select after every
// kPairsMaxSpheres spheres against the {c,-kappa} and {c,r2} tables in global memory (L2).
template <int W, bool FMA>
__device__ __forceinline__ void stream_pass(WarpStream &ws, const float4 *__restrict__ slab, const float4 *__restrict__ scan,
                                            const float4 *__restrict__ spheres, uint32_t n_pad, uint32_t n_real,
                                            const SlabRays &sr, const ScanRays &rays, unsigned lane,
                                            pk2 ox, pk2 oy, pk2 oz, pk2 dx, pk2 dy, pk2 dz, float neg_one, float one, Best &b0, Best &b1) {
    constexpr uint32_t kF4 = kStreamStageF4;                  // float4 per chunk of the slab table
    const uint32_t n_chunks = (n_pad + kStreamChunk - 1) / kStreamChunk;
    if (lane == 0) {
        for (uint32_t c = 0; c < (uint32_t)kStreamStages && c < n_chunks; ++c) {
            const uint32_t cnt = min((uint32_t)kStreamChunk, n_pad - c * kStreamChunk);
            stream_issue(ws.stage + c * kF4, slab + c * kF4, cnt * 12u, ws.bar + c);
        }
    }
    CandQueue queue;
    uint32_t qn = 0;
    const FetchGlobal fetch = {spheres};
    const FetchScanGlobal fetch_scan = {scan};
#pragma unroll 1
    for (uint32_t c = 0; c < n_chunks; ++c) {
        const uint32_t st = c % kStreamStages;
        mbar_wait(ws.bar + st, (ws.parity >> st) & 1u);
        ws.parity ^= 1u << st;
        const uint32_t cnt = min((uint32_t)kStreamChunk, n_pad - c * kStreamChunk);
        scan_slab_tile(ws.stage + st * kF4, cnt, c * kStreamChunk, sr, queue, qn);
        __syncwarp();
        const uint32_t nc = c + kStreamStages;
        if (lane == 0 && nc < n_chunks) {
            const uint32_t ncnt = min((uint32_t)kStreamChunk, n_pad - nc * kStreamChunk);
            stream_issue(ws.stage + st * kF4, slab + nc * kF4, ncnt * 12u, ws.bar + st);
        }
        if (((c + 1) * kStreamChunk) % kPairsMaxSpheres == 0 || c + 1 == n_chunks) {
            refine_queue(fetch_scan, queue, qn, rays);
            select_update<W, FMA>(fetch, n_real, queue, qn, ox, oy, oz, dx, dy, dz, neg_one, one, b0, b1);
            qn = 0;
        }
    }
}

// Bundle mode, streamed scenes: the level-0 scan reads the launch's slab table straight from global memory
// (L2; every lane a different 48-byte group, coalesced), refine + select per kPairsMaxSpheres segment.
template <int W, bool FMA>
__device__ __forceinline__ void stream_bundle_pass(const float4 *__restrict__ slab, const float4 *__restrict__ scan, const float4 *__restrict__ spheres,
                                                   uint32_t n_pad, uint32_t n_real, float op, float oq, unsigned heads0, unsigned heads1,
                                                   const BundlePixel &bp0, const BundlePixel &bp1, const ScanRays &rays, unsigned lane,
                                                   pk2 ox, pk2 oy, pk2 oz, pk2 dx, pk2 dy, pk2 dz, float neg_one, float one, Best &b0, Best &b1) {
    CandQueue queue;
    const FetchGlobal fetch = {spheres};
    const FetchScanGlobal fetch_scan = {scan};
#pragma unroll 1
    for (uint32_t first = 0; first < n_pad; first += (uint32_t)kPairsMaxSpheres) {
        uint32_t qn = 0;
        bundle_scan<true>(slab, first, min((uint32_t)kPairsMaxSpheres, n_pad - first), op, oq, heads0, heads1, bp0, bp1, lane, queue, qn);
        refine_queue(fetch_scan, queue, qn, rays);
        select_update<W, FMA>(fetch, n_real, queue, qn, ox, oy, oz, dx, dy, dz, neg_one, one, b0, b1);
    }
}

// Bundle mode, A8 (+ A9): all samples of chunk c of pixel `pix` are staged in row `row`; fold them into the
// running mean in sample order (main.cpp:484-489 is sequential in k) and publish the chunk count.  One chunk per
// LANE: the slots that delivered a chunk's last sample queue up in ring F and a warp folds 32-64 chunks at a time
// (folding each chunk by a whole warp as soon as it completed cost 8 % of the kernel's instructions).
__device__ __forceinline__ void fold_chunk(const KernelParams &P, uint32_t pix, uint32_t row, uint32_t c) {
    const uint32_t n0 = c << P.bundle_log2;
    const uint32_t nvalid = min(1u << P.bundle_log2, P.n_samples - n0);
    const float4 *st = P.stage + ((size_t)row << P.bundle_log2);
    const bool mean_mode = P.accum_mode == SRT_ACCUM_MEAN;
    float4 m = make_float4(0.0f, 0.0f, 0.0f, 0.0f);                       // PreviousColor * (0/1), zeroed arena
    if (c != 0u || P.first_sample > 0u || !mean_mode) m = __ldcg(&P.accum[pix]);
#pragma unroll 4
    for (uint32_t j = 0; j < nvalid; ++j) {
        const float4 v = __ldcg(st + j);
        if (mean_mode) {                                                  // Out * (1/N) + Prev * (PrevCount/N), N = k + 1 (host table, IEEE)
            const float2 w = __ldg(P.fold_w + n0 + j);
            m.x = __fadd_rn(__fmul_rn(w.x, v.x), __fmul_rn(w.y, m.x));
            m.y = __fadd_rn(__fmul_rn(w.x, v.y), __fmul_rn(w.y, m.y));
            m.z = __fadd_rn(__fmul_rn(w.x, v.z), __fmul_rn(w.y, m.z));
        } else {
            m.x = __fadd_rn(m.x, v.x); m.y = __fadd_rn(m.y, v.y); m.z = __fadd_rn(m.z, v.z);
        }
    }
    P.accum[pix] = make_float4(m.x, m.y, m.z, 1.0f);
    if (c + 1u == P.n_chunks && mean_mode) P.rgba[pix] = pack_rgba8(m.x, m.y, m.z);
    P.cnt[row] = 0u;
    __threadfence();                                                      // accum, cnt before the chunk count
    *(volatile uint32_t *)&P.done[pix] = P.done_tag | (c + 1u);
}

#ifdef SRT_SPIN_LIMIT
// Debug build: who holds each path slot (0 = in a ring, 1 + warp = held by that warp, 0xFFFF = retired).  A slot id taken
// twice (site 30) or pushed / retired by a warp that does not hold it (sites 31-33) is a broken ring protocol; sites 16-26, 34
// check the indices read from slot state and queues before they are used.  The report (site, CTA, thread, one value) comes
// back through srt_wait.  History: the first run of scripts/soak.py on this build found the missing __syncwarp() after the
// ring pops.
__device__ unsigned int g_dbg_owner[1024 * 2048];
#define SRT_OWN_SET(site, s, expect, val) do { const unsigned int o_ = atomicExch(&g_dbg_owner[(blockIdx.x & 1023u) * 2048u + (s)], (val)); \
        SRT_CHECKV(site, o_ == (expect), ((unsigned long long)o_ << 32) | ((s) << 8) | kind); } while (0)
#else
#define SRT_OWN_SET(site, s, expect, val) do { } while (0)
#endif
template <int W, bool FMA, int THREADS, int MINB, bool STREAM, bool BUNDLE = false>
__global__ void __launch_bounds__(THREADS, MINB)
render_pool(const __grid_constant__ KernelParams P) {
    constexpr uint32_t NS = 64u * (THREADS / 32 + kPoolSpareBatches);   // slots per CTA
    constexpr uint32_t RCAP = NS <= 1024u ? 1024u : 2048u;     // ring capacity (power of two >= NS)
    extern __shared__ float4 s_scan[];                         // resident: {cx,cy,cz,-kappa}
    const uint32_t n_res = STREAM ? 0u : P.n_pad;              // spheres resident in shared memory
    float4 *const s_slab = s_scan + n_res;                     // resident: slab table, 3 float4 per 4 spheres (srt_pairs.cuh)
    float *const s_r2 = reinterpret_cast<float *>(s_slab + (n_res >> 2) * 3u);
    float *const st = s_r2 + n_res;
    uint32_t *const su = reinterpret_cast<uint32_t *>(st);
    ring_t *const ring0 = reinterpret_cast<ring_t *>(su + (BUNDLE ? kBundleFields : S_COUNT) * NS);
    ring_t *const ring1 = ring0 + RCAP;
    ring_t *const ring2 = ring1 + RCAP;                       // bundle mode only
    PoolCtl *const ctl = reinterpret_cast<PoolCtl *>(ring2 + (BUNDLE ? RCAP : 0u));
    const uint32_t tid = threadIdx.x;
    const unsigned lane = tid & 31u;
    // streamed: per-warp TMA stages + barriers behind the control block (16-byte aligned)
    WarpStream ws;
    {
        unsigned char *base = reinterpret_cast<unsigned char *>(ctl) + ((sizeof(PoolCtl) + 15u) & ~15u);
        ws.stage = reinterpret_cast<float4 *>(base) + (tid >> 5) * (kStreamStages * kStreamStageF4);
        ws.bar = reinterpret_cast<uint64_t *>(base + (size_t)(THREADS / 32) * kStreamStages * kStreamStageF4 * sizeof(float4)) + (tid >> 5) * kStreamStages;
        ws.parity = 0u;
    }
#define SRT_ST(f, id) st[(f) * NS + (id)]
#define SRT_SU(f, id) su[(f) * NS + (id)]
    const uint32_t k_end = P.first_sample + P.n_samples * P.sample_stride;
    const f3 cam_o = mk3(P.cam.pos[0], P.cam.pos[1], P.cam.pos[2]);
    {
        if (!STREAM) {
            for (uint32_t i = tid; i < P.n_pad; i += THREADS) {
                const float4 c = P.scan_table[i];
                const float r2 = P.spheres[i].w;
                s_scan[i] = c;
                s_r2[i] = r2;
                slab_entry(reinterpret_cast<float *>(s_slab), i, c, r2, i < P.n_spheres, P.slab_axis, P.omax);
            }
        } else if (lane == 0) {
            for (int i = 0; i < kStreamStages; ++i) mbar_init(ws.bar + i, 1u);
            mbar_fence_init();
        }
        for (uint32_t i = tid; i < NS; i += THREADS) {
            SRT_SU(S_PIX, i) = kInvalidPixel;
            SRT_SU(S_K, i) = k_end;
            SRT_SU(S_KEND, i) = k_end;
            SRT_SU(S_FLAGS, i) = 0u;
        }
#ifdef SRT_SPIN_LIMIT
        SRT_CHECK(15, gridDim.x <= kDbgCtas);
        for (uint32_t i = tid; i < 2048u; i += THREADS) g_dbg_owner[(blockIdx.x & 1023u) * 2048u + i] = 0u;
        for (uint32_t i = tid; i < 3u * 2048u; i += THREADS) { g_dbg_busy[dbg_ring_index(0u, 0u) + i] = 0u; g_dbg_wpos[dbg_ring_index(0u, 0u) + i] = i < NS ? i : 0xFFFFFFFFu; }
#endif
        for (uint32_t i = tid; i < RCAP; i += THREADS) { ring0[i] = i < NS ? (ring_t)i : kRingSlotEmpty; ring1[i] = kRingSlotEmpty; if (BUNDLE) ring2[i] = kRingSlotEmpty; }
        if (tid == 0) {
            ctl->head[0] = 0; ctl->tail[0] = NS; ctl->avail[0] = NS;
            ctl->head[1] = 0; ctl->tail[1] = 0; ctl->avail[1] = 0;
            ctl->head[2] = 0; ctl->tail[2] = 0; ctl->avail[2] = 0;
            ctl->inflight = 0; ctl->retired = 0; ctl->spins_steady = 0; ctl->spins_tail = 0;
            ctl->scanning[0] = ctl->scanning[1] = ctl->scanning[2] = ctl->scanning[3] = 0;
        }
    }
    __syncthreads();
    unsigned long long rays = 0;
    bool hungry = false;

    for (;;) {
        uint32_t kind = 0, n = 0, pos = 0;
        if (lane == 0) pool_acquire<BUNDLE>(ctl, NS, hungry, kind, n, pos);
        kind = __shfl_sync(0xFFFFFFFFu, kind, 0);
        n = __shfl_sync(0xFFFFFFFFu, n, 0);
        pos = __shfl_sync(0xFFFFFFFFu, pos, 0);
        if (n == 0) break;
        ring_t *const ring = kind == 2u ? ring2 : (kind ? ring1 : ring0);
        uint32_t id0 = lane < n ? pool_pop(ring, RCAP - 1u, pos + lane, kind) : kRingEmpty;
        uint32_t id1 = lane + 32u < n ? pool_pop(ring, RCAP - 1u, pos + 32u + lane, kind) : kRingEmpty;
        // Reconverge HERE.  A lane whose entry is still being written spins in pool_pop while the lanes that already hold
        // their ids could run ahead -- there is no other full-mask operation before the end of the scan, so under independent
        // thread scheduling the early lanes may finish a whole scan / refine / select before the late ones even read their
        // entries.  Those entries then stay in the ring for a pass; the producers of the next lap queue up behind them, two
        // producers end up waiting on one entry, both write when it is cleared, and a slot id is lost or taken twice.  Where
        // the compiler places the convergence barrier after the spin loop is its choice: the release build happened to
        // reconverge right after the loop, the -DSRT_SPIN_LIMIT build did not (found by scripts/soak.py on that build).
        __syncwarp();
        __threadfence_block();
        SRT_CHECK(16, (id0 == kRingEmpty || id0 < NS) && (id1 == kRingEmpty || id1 < NS));
        if (id0 != kRingEmpty) SRT_OWN_SET(30, id0, 0u, 1u + (tid >> 5));
        if (id1 != kRingEmpty) SRT_OWN_SET(30, id1, 0u, 1u + (tid >> 5));

        if (BUNDLE && kind == 2u) {
            // ---- up to 64 slots that delivered the last sample of a pixel chunk: fold the chunks, one per lane (A8, A9) ----
            __threadfence();                                              // the other samples' staging stores
#pragma unroll 1
            for (int r = 0; r < 2; ++r) {
                const uint32_t s = r ? id1 : id0;
                if (s != kRingEmpty) {
                    SRT_CHECKV(17, SRT_SU(S_PIX, s) < P.width * P.height, ((unsigned long long)SRT_SU(S_PIX, s) << 32) | (s << 16) | (SRT_SU(S_FLAGS, s) & 0xFFFFu));
                    SRT_CHECK(18, SRT_SU(S_PXY, s) < P.band_items);
                    SRT_CHECK(19, (SRT_SU(S_K, s) >> P.bundle_log2) < P.n_chunks);
                    fold_chunk(P, SRT_SU(S_PIX, s), SRT_SU(S_PXY, s), SRT_SU(S_K, s) >> P.bundle_log2);
                }
                __syncwarp();
                if (s != kRingEmpty) SRT_OWN_SET(31, s, 1u + (tid >> 5), 0u);
                pool_push(ctl, ring0, RCAP - 1u, 0, s != kRingEmpty, s, lane);
            }
            hungry = false;
            if (lane == 0) atomicSub(&ctl->inflight, 1u);
            continue;
        }
        f3 o0, o1, d0, d1;
        Hit h0, h1;
        Best b0 = best_init(), b1 = best_init();
        pk2 ox, oy, oz, dx, dy, dz;
        CandQueue queue;                 // resident tables: one queue for either kind of pass
        uint32_t qn = 0;
        uint32_t pend = 0;               // bit r: slot of half r waits for the preceding sample chunk of its pixel
        bool any_real = true;
        BundlePixel bp0 = {0.0f, 0.0f, 0.0f, -1.0f}, bp1 = bp0;
        unsigned heads0 = 0u, heads1 = 0u;
        bool use_bundle = false;
        if (BUNDLE && kind == 0) {
            // ---- bundle mode: 64 free slots take 64 consecutive work items = consecutive samples of one pixel --
            bool need0 = id0 != kRingEmpty && !(SRT_SU(S_FLAGS, id0) & kFlagPending);
            bool need1 = id1 != kRingEmpty && !(SRT_SU(S_FLAGS, id1) & kFlagPending);
            unsigned m0 = __ballot_sync(0xFFFFFFFFu, need0), m1 = __ballot_sync(0xFFFFFFFFu, need1);
            unsigned n_retired = 0;
            while (m0 | m1) {
                unsigned base = 0;
                if (lane == 0) base = atomicAdd(P.work_counter, (unsigned)(__popc(m0) + __popc(m1)));
                base = __shfl_sync(0xFFFFFFFFu, base, 0);
                const unsigned lt = (1u << lane) - 1u;
#pragma unroll 1
                for (int r = 0; r < 2; ++r) {
                    if (r ? need1 : need0) {
                        const uint32_t s = r ? id1 : id0;
                        const uint32_t item = base + (r ? (uint32_t)(__popc(m0) + __popc(m1 & lt)) : (uint32_t)__popc(m0 & lt));
                        bool need = true;
                        if (item >= P.total_items) {                     // out of work: the slot retires
                            SRT_OWN_SET(32, s, 1u + (tid >> 5), 0xFFFFu);
                            SRT_SU(S_PIX, s) = kInvalidPixel;
                            if (r) id1 = kRingEmpty; else id0 = kRingEmpty;
                            n_retired += 1u;
                            need = false;
                        } else {
                            BundleItem it;
                            if (bundle_item(P, item, it)) {
                                SRT_SU(S_PIX, s) = it.py * P.width + it.px; SRT_SU(S_K, s) = it.n; SRT_SU(S_PXY, s) = it.row;
                                SRT_SU(S_KEND, s) = it.px | (it.py << 16); SRT_SU(S_FLAGS, s) = kFlagPending;
                                need = false;
                            }
                        }
                        if (r) need1 = need; else need0 = need;
                    }
                }
                m0 = __ballot_sync(0xFFFFFFFFu, need0); m1 = __ballot_sync(0xFFFFFFFFu, need1);
            }
            for (int off = 16; off > 0; off >>= 1) n_retired += __shfl_down_sync(0xFFFFFFFFu, n_retired, off);
            if (lane == 0 && n_retired) atomicAdd(&ctl->retired, n_retired);
            bool real = false, odd = false;
            uint32_t key0 = kInvalidPixel, key1 = kInvalidPixel;
#pragma unroll 1
            for (int r = 0; r < 2; ++r) {
                const uint32_t s = r ? id1 : id0;
                f3 o = mk3(kIdleOrigin, kIdleOrigin, kIdleOrigin), d = mk3(0.0f, 0.0f, 0.0f);   // parked unless a ray is generated
                BundlePixel bp = {0.0f, 0.0f, 0.0f, -1.0f};
                uint32_t key = kInvalidPixel;
                if (s != kRingEmpty) {
                    const uint32_t pix = SRT_SU(S_PIX, s), n = SRT_SU(S_K, s), pxy = SRT_SU(S_KEND, s);
                    const uint32_t c = n >> P.bundle_log2;
                    // the staging row of this pixel chunk must have been released by its previous user: chunk c-1 of
                    // the same pixel, or the last chunk of the pixel at the same position one band earlier (bands are
                    // whole block rows: same column, inside the image as well)
                    bool row_free = true;
                    SRT_CHECKV(23, pix < P.width * P.height, ((unsigned long long)pix << 32) | (s << 16) | (SRT_SU(S_FLAGS, s) & 0xFFFFu));
                    if (c != 0u) row_free = *(volatile uint32_t *)&P.done[pix] == (P.done_tag | c);
                    else if (pix >= P.band_pix) row_free = *(volatile uint32_t *)&P.done[pix - P.band_pix] == (P.done_tag | P.n_chunks);
                    if (row_free) {
                        uint64_t rng;
                        primary_ray<true>(P, pxy & 0xFFFFu, pxy >> 16, pix, P.first_sample + n * P.sample_stride, rng, o, d);
                        SRT_ST(S_AX, s) = 1.0f; SRT_ST(S_AY, s) = 1.0f; SRT_ST(S_AZ, s) = 1.0f;
                        SRT_ST(S_LX, s) = 0.0f; SRT_ST(S_LY, s) = 0.0f; SRT_ST(S_LZ, s) = 0.0f;
                        SRT_SU(S_RNG_LO, s) = (uint32_t)rng; SRT_SU(S_RNG_HI, s) = (uint32_t)(rng >> 32);
                        SRT_SU(S_FLAGS, s) = 0u;                                      // bounce 0
                        bp = bundle_pixel(P, pxy & 0xFFFFu, pxy >> 16);
                        key = pix;
                        real = true;
                        odd = odd || bp.eps < 0.0f;
                    } else pend |= 1u << r;
                }
                if (r) { o1 = o; d1 = d; bp1 = bp; key1 = key; } else { o0 = o; d0 = d; bp0 = bp; key0 = key; }
            }
            any_real = __any_sync(0xFFFFFFFFu, real);
            use_bundle = any_real && !__any_sync(0xFFFFFFFFu, odd);
            // one level-0 scan per run of equal pixels (work items are handed out in pixel order)
            const uint32_t up0 = __shfl_up_sync(0xFFFFFFFFu, key0, 1), up1 = __shfl_up_sync(0xFFFFFFFFu, key1, 1);
            const uint32_t last0 = __shfl_sync(0xFFFFFFFFu, key0, 31);
            heads0 = __ballot_sync(0xFFFFFFFFu, key0 != kInvalidPixel && (lane == 0 || up0 != key0));
            heads1 = __ballot_sync(0xFFFFFFFFu, key1 != kInvalidPixel && ((lane == 0 ? last0 : up1) != key1));
            if (P.debug && lane == 0) {
                atomicAdd(&P.debug[use_bundle ? 2 : 3], 1ull);
                atomicAdd(&P.debug[4], (unsigned long long)(__popc(heads0) + __popc(heads1)));
                atomicAdd(&P.debug[5], (unsigned long long)(pend != 0u));
            }
        } else if (kind == 0) {
            // ---- 64 slots that need a new primary ray: finish items, fetch new ones (A1) -------
            bool real = false;
#pragma unroll 1
            for (int r = 0; r < 2; ++r) {
                const uint32_t s = r ? id1 : id0;
                const bool valid = s != kRingEmpty;
                uint32_t pix = kInvalidPixel, k = k_end, kend = k_end, pxy = 0;
                bool pending = false;
                if (valid) {
                    pix = SRT_SU(S_PIX, s); k = SRT_SU(S_K, s); kend = SRT_SU(S_KEND, s); pxy = SRT_SU(S_PXY, s);
                    pending = (SRT_SU(S_FLAGS, s) & kFlagPending) != 0u;
                }
                bool need = valid && !pending && (k == kend);
                if (need && pix != kInvalidPixel) {
                    // the item's samples are folded: A8 result (+ A9 tonemap/pack after the pixel's last chunk), main.cpp:484-492
                    const float mx = SRT_ST(S_MX, s), my = SRT_ST(S_MY, s), mz = SRT_ST(S_MZ, s);
                    P.accum[pix] = make_float4(mx, my, mz, 1.0f);
                    if (kend == k_end) {
                        if (P.accum_mode == SRT_ACCUM_MEAN) P.rgba[pix] = pack_rgba8(mx, my, mz);
                    } else {
                        __threadfence();                                   // accum before the chunk count
                        *(volatile uint32_t *)&P.done[pix] = P.done_tag | ((kend - P.first_sample) / P.chunk_span);
                    }
                    pix = kInvalidPixel;
                }
                bool retire = false;
                unsigned m = __ballot_sync(0xFFFFFFFFu, need);
                while (m) {
                    const unsigned leader = __ffs(m) - 1;
                    unsigned base = 0;
                    if (lane == leader) base = atomicAdd(P.work_counter, __popc(m));
                    base = __shfl_sync(0xFFFFFFFFu, base, leader);
                    if (need) {
                        const uint32_t item = base + __popc(m & ((1u << lane) - 1u));
                        uint32_t px, py;
                        if (item >= P.total_items) { retire = true; need = false; }
                        else {
                            const uint32_t c = item / P.items_per_gen;      // chunk-major: every pixel's chunk c before any chunk c+1
                            if (item_to_pixel(P, item - c * P.items_per_gen, px, py)) {
                                pix = py * P.width + px;
                                pxy = px | (py << 16);
                                k = P.first_sample + c * P.chunk_span;
                                kend = (k_end - k <= P.chunk_span) ? k_end : k + P.chunk_span;
                                need = false;
                                pending = true;
                            }
                        }
                    }
                    m = __ballot_sync(0xFFFFFFFFu, need);
                }
                if (pending) {
                    // the running mean is sequential in k (main.cpp:484-489): chunk c starts from the mean chunk c-1 published
                    const uint32_t c = (k - P.first_sample) / P.chunk_span;
                    if (c == 0u || *(volatile uint32_t *)&P.done[pix] == (P.done_tag | c)) {
                        float4 prev = make_float4(0.0f, 0.0f, 0.0f, 0.0f);   // PreviousColor * (0/1), zeroed arena
                        if (c != 0u) { __threadfence(); prev = __ldcg(&P.accum[pix]); }
                        else if (P.first_sample > 0 || P.accum_mode == SRT_ACCUM_SUM) prev = __ldcg(&P.accum[pix]);
                        SRT_ST(S_MX, s) = prev.x; SRT_ST(S_MY, s) = prev.y; SRT_ST(S_MZ, s) = prev.z;
                        pending = false;
                    }
                }
                const unsigned rm = __ballot_sync(0xFFFFFFFFu, retire);
                if (rm && lane == (unsigned)(__ffs(rm) - 1)) atomicAdd(&ctl->retired, (unsigned)__popc(rm));
                f3 o = mk3(kIdleOrigin, kIdleOrigin, kIdleOrigin), d = mk3(0.0f, 0.0f, 0.0f);   // parked unless a ray is generated
                if (retire) { SRT_SU(S_PIX, s) = kInvalidPixel; if (r) id1 = kRingEmpty; else id0 = kRingEmpty; }
                else if (valid && pending) {
                    SRT_SU(S_PIX, s) = pix; SRT_SU(S_K, s) = k; SRT_SU(S_KEND, s) = kend; SRT_SU(S_PXY, s) = pxy; SRT_SU(S_FLAGS, s) = kFlagPending;
                    pend |= 1u << r;
                } else if (valid) {
                    uint64_t rng;
                    primary_ray<true>(P, pxy & 0xFFFFu, pxy >> 16, pix, k, rng, o, d);
                    SRT_ST(S_AX, s) = 1.0f; SRT_ST(S_AY, s) = 1.0f; SRT_ST(S_AZ, s) = 1.0f;
                    SRT_ST(S_LX, s) = 0.0f; SRT_ST(S_LY, s) = 0.0f; SRT_ST(S_LZ, s) = 0.0f;
                    SRT_SU(S_RNG_LO, s) = (uint32_t)rng; SRT_SU(S_RNG_HI, s) = (uint32_t)(rng >> 32);
                    SRT_SU(S_PIX, s) = pix; SRT_SU(S_K, s) = k; SRT_SU(S_KEND, s) = kend; SRT_SU(S_PXY, s) = pxy; SRT_SU(S_FLAGS, s) = 0u;   // bounce 0
                    real = true;
                }
                if (r) { o1 = o; d1 = d; } else { o0 = o; d0 = d; }
            }
            any_real = __any_sync(0xFFFFFFFFu, real);
        } else {
            // ---- 64 slots with a secondary ray ready ------------------------------------------
            o0 = o1 = mk3(kIdleOrigin, kIdleOrigin, kIdleOrigin);
            d0 = d1 = mk3(0.0f, 0.0f, 0.0f);
            if (id0 != kRingEmpty) {
                o0 = mk3(SRT_ST(S_OX, id0), SRT_ST(S_OY, id0), SRT_ST(S_OZ, id0));
                d0 = mk3(SRT_ST(S_DX, id0), SRT_ST(S_DY, id0), SRT_ST(S_DZ, id0));
            }
            if (id1 != kRingEmpty) {
                o1 = mk3(SRT_ST(S_OX, id1), SRT_ST(S_OY, id1), SRT_ST(S_OZ, id1));
                d1 = mk3(SRT_ST(S_DX, id1), SRT_ST(S_DY, id1), SRT_ST(S_DZ, id1));
            }
        }
        ox = pk_make(o0.x, o1.x); oy = pk_make(o0.y, o1.y); oz = pk_make(o0.z, o1.z);
        dx = pk_make(d0.x, d1.x); dy = pk_make(d0.y, d1.y); dz = pk_make(d0.z, d1.z);
        if (any_real) {
            // level 1 (slab, every sphere; level 0 = once per pixel bundle for the primary passes of bundle mode)
            // -> level 2 (7-FMA filter on the survivors) -> exact select
            const int ip = P.slab_axis == 0 ? 1 : (P.slab_axis == 1 ? 2 : 0), iq = P.slab_axis == 0 ? 2 : (P.slab_axis == 1 ? 0 : 1);
            if (STREAM) {
                const ScanRays rays = make_scan_rays(ox, oy, oz, dx, dy, dz);
                if (BUNDLE && use_bundle) {
                    stream_bundle_pass<W, FMA>(P.ptab, P.scan_table, P.spheres, P.n_pad, P.n_spheres, P.cam.pos[ip], P.cam.pos[iq], heads0, heads1, bp0, bp1,
                                               rays, lane, ox, oy, oz, dx, dy, dz, P.neg_one, P.one, b0, b1);
                } else {
                    const SlabRays sr = make_slab_rays(P.slab_axis, ox, oy, oz, dx, dy, dz, P.omax2);
                    if (BUNDLE && kScanTarget && lane == 0) atomicAdd(&ctl->scanning[(tid >> 5) & 3u], 1u);
                    stream_pass<W, FMA>(ws, P.ptab, P.scan_table, P.spheres, P.n_pad, P.n_spheres, sr, rays, lane, ox, oy, oz, dx, dy, dz, P.neg_one, P.one, b0, b1);
                    if (BUNDLE && kScanTarget && lane == 0) atomicSub(&ctl->scanning[(tid >> 5) & 3u], 1u);
                }
            } else {
                if (BUNDLE && use_bundle) {
                    bundle_scan<false>(s_slab, 0u, P.n_pad, P.cam.pos[ip], P.cam.pos[iq], heads0, heads1, bp0, bp1, lane, queue, qn);
                } else {
                    const SlabRays sr = make_slab_rays(P.slab_axis, ox, oy, oz, dx, dy, dz, P.omax2);
                    if (BUNDLE && kScanTarget && lane == 0) atomicAdd(&ctl->scanning[(tid >> 5) & 3u], 1u);
                    scan_slab_tile(s_slab, P.n_pad, 0u, sr, queue, qn);
                    if (BUNDLE && kScanTarget && lane == 0) atomicSub(&ctl->scanning[(tid >> 5) & 3u], 1u);
                }
                const ScanRays rays = make_scan_rays(ox, oy, oz, dx, dy, dz);
                const FetchScanResident fetch_scan = {s_scan};
                refine_queue(fetch_scan, queue, qn, rays);
                const FetchResident fetch = {s_scan, s_r2};
                select_update<W, FMA>(fetch, P.n_spheres, queue, qn, ox, oy, oz, dx, dy, dz, P.neg_one, P.one, b0, b1);
            }
            select_finish<W, FMA>(P.spheres, P.n_pad, b0, b1, ox, oy, oz, dx, dy, dz, h0, h1);
        }
        hungry = !any_real && __any_sync(0xFFFFFFFFu, pend != 0u);

        // ---- shade / scatter, one half after the other; push every slot to its next ring ------
#pragma unroll 1
        for (int r = 0; r < 2; ++r) {
            const uint32_t s = r ? id1 : id0;
            const bool waits = ((pend >> r) & 1u) != 0u;
            const bool valid = s != kRingEmpty && !waits;
            bool ended = false, folds = false;
            if (valid) {
                const Hit h = r ? h1 : h0;
                f3 o = r ? o1 : o0, d = r ? d1 : d0;
                rays += 1;                                                    // main.cpp:390
                const uint32_t k = SRT_SU(S_K, s);
                uint32_t bounce = SRT_SU(S_FLAGS, s) & kFlagBounceMask;
                f3 atten = mk3(SRT_ST(S_AX, s), SRT_ST(S_AY, s), SRT_ST(S_AZ, s));
                f3 out = mk3(SRT_ST(S_LX, s), SRT_ST(S_LY, s), SRT_ST(S_LZ, s));
                uint64_t rng = (uint64_t)SRT_SU(S_RNG_LO, s) | ((uint64_t)SRT_SU(S_RNG_HI, s) << 32);
                if (bounce == 0 && P.tap_t != nullptr && k == (BUNDLE ? 0u : P.first_sample)) {
                    const uint32_t pix = SRT_SU(S_PIX, s);
                    P.tap_t[pix] = h.t;
                    P.tap_prim[pix] = h.prim;
                }
                float4 sp = make_float4(0.f, 0.f, 0.f, 0.f);
                SRT_CHECK(22, !h.hit || h.prim < P.n_pad);
                if (h.hit) sp = STREAM ? __ldg(P.spheres + h.prim) : s_scan[h.prim];   // centre only; .w is not used by shading
                ended = shade_scatter<true>(P, h, sp, o, d, atten, out, rng, bounce);
                if (BUNDLE && ended) {
                    // stage the sample (S_K holds its ordinal); the slot that delivers the chunk's last one folds the chunk
                    const uint32_t row = SRT_SU(S_PXY, s), c = k >> P.bundle_log2;
                    SRT_CHECK(20, row < P.band_items);
                    SRT_CHECK(21, k < P.n_samples);
                    P.stage[((size_t)row << P.bundle_log2) + (k & ((1u << P.bundle_log2) - 1u))] = make_float4(out.x, out.y, out.z, 0.0f);
                    __threadfence();
                    folds = atomicAdd(&P.cnt[row], 1u) + 1u == min(1u << P.bundle_log2, P.n_samples - (c << P.bundle_log2));
                } else if (ended) {
                    const f3 mean = accumulate_sample(P.accum_mode, mk3(SRT_ST(S_MX, s), SRT_ST(S_MY, s), SRT_ST(S_MZ, s)), out, k);
                    SRT_ST(S_MX, s) = mean.x; SRT_ST(S_MY, s) = mean.y; SRT_ST(S_MZ, s) = mean.z;
                    SRT_SU(S_K, s) = k + P.sample_stride;
                } else {
                    SRT_SU(S_RNG_LO, s) = (uint32_t)rng; SRT_SU(S_RNG_HI, s) = (uint32_t)(rng >> 32);
                    SRT_ST(S_OX, s) = o.x; SRT_ST(S_OY, s) = o.y; SRT_ST(S_OZ, s) = o.z;
                    SRT_ST(S_DX, s) = d.x; SRT_ST(S_DY, s) = d.y; SRT_ST(S_DZ, s) = d.z;
                    SRT_ST(S_AX, s) = atten.x; SRT_ST(S_AY, s) = atten.y; SRT_ST(S_AZ, s) = atten.z;
                    SRT_ST(S_LX, s) = out.x; SRT_ST(S_LY, s) = out.y; SRT_ST(S_LZ, s) = out.z;
                    SRT_SU(S_FLAGS, s) = bounce;
                }
            }
            {   // ONE copy of the push code for the three rings (they are contiguous in shared memory): the hot code sits right at
                // the 32 KB of the L1.5 instruction cache; three inlined copies cost 1.6 % of the 4K / 256-spp launch
                const int dest = folds ? 2 : (((valid && ended) || waits) ? 0 : ((valid && !ended) ? 1 : 3));
                if (dest != 3) { SRT_OWN_SET(33, s, 1u + (tid >> 5), 0u); SRT_CHECKV(34, SRT_SU(S_PIX, s) < P.width * P.height, (s << 8) | (unsigned)dest); }
#pragma unroll 1
                for (int q = BUNDLE ? 2 : 1; q >= 0; --q) pool_push(ctl, ring0 + (uint32_t)q * RCAP, RCAP - 1u, q, dest == q, s, lane);
            }
        }
        if (lane == 0) atomicSub(&ctl->inflight, 1u);
    }
#undef SRT_ST
#undef SRT_SU

    __syncthreads();
    if (tid == 0 && P.debug) { atomicAdd(&P.debug[0], (unsigned long long)ctl->spins_steady); atomicAdd(&P.debug[1], (unsigned long long)ctl->spins_tail); }
    finish_launch(P, rays);
}

// ------------------------------------------------------------------------------------------
// audit_filter: diagnostic kernel behind srt_filter_audit().  One path per thread (pixel),
// advanced with the literal per-lane reference arithmetic; for EVERY (ray, sphere) pair of
// every path segment it evaluates both the reference's hit predicate and the ray-pair
// kernel's filter predicate (the very filter_ray / filter_sphere code) and counts
//   tests, exact hits, filter candidates, and VIOLATIONS (exact hit that the filter rejects).
// The filter is sound iff violations == 0.
// ------------------------------------------------------------------------------------------
template <int W, bool FMA>
__global__ void __launch_bounds__(kBlockThreads)
audit_filter(const __grid_constant__ KernelParams P, unsigned long long *counters) {
    unsigned long long n_tests = 0, n_hits = 0, n_cand = 0, n_viol = 0, n_slab = 0, n_btests = 0, n_bcand = 0, n_bviol = 0;
    const uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix < P.width * P.height) {
        const uint32_t px = pix % P.width, py = pix / P.width;
        for (uint32_t k = P.first_sample; k < P.first_sample + P.n_samples; ++k) {
            uint64_t rng;
            f3 o, d, atten = mk3(1.0f, 1.0f, 1.0f), out = mk3(0.0f, 0.0f, 0.0f);
            uint32_t bounce = 0;
            primary_ray(P, px, py, pix, k, rng, o, d);
            const BundlePixel bp = bundle_pixel(P, px, py);            // level 0 of bundle mode: primary rays only
            const int ip = P.slab_axis == 0 ? 1 : (P.slab_axis == 1 ? 2 : 0), iq = P.slab_axis == 0 ? 2 : (P.slab_axis == 1 ? 0 : 1);
            for (;;) {
                const FilterRay f = filter_ray(o.x, o.y, o.z, d.x, d.y, d.z);
                const pk2 p2ox = pk_make(f.p2ox, f.p2ox), p2oy = pk_make(f.p2oy, f.p2oy), p2oz = pk_make(f.p2oz, f.p2oz);
                const pk2 fdx = pk_make(f.dx, f.dx), fdy = pk_make(f.dy, f.dy), fdz = pk_make(f.dz, f.dz), od = pk_make(f.od, f.od);
                const SlabRay sl = slab_ray(P.slab_axis, o.x, o.y, o.z, d.x, d.y, d.z, P.omax2);
                for (uint32_t i = 0; i < P.n_pad; ++i) {
                    const float4 s = P.spheres[i];
                    const float cp = P.slab_axis == 0 ? s.y : (P.slab_axis == 1 ? s.z : s.x), cq = P.slab_axis == 0 ? s.z : (P.slab_axis == 1 ? s.x : s.y);
                    const bool slab = slab_test(cp, cq, i < P.n_spheres ? slab_radius(s.w, s.x, s.y, s.z, P.omax) : kSlabPadRadius, sl);
                    const float scx = __fsub_rn(s.x, o.x), scy = __fsub_rn(s.y, o.y), scz = __fsub_rn(s.z, o.z);
                    const float t = lane_dot<(W != 1) && FMA>(scx, scy, scz, d.x, d.y, d.z);
                    const float vx = __fsub_rn(scx, __fmul_rn(d.x, t));
                    const float vy = __fsub_rn(scy, __fmul_rn(d.y, t));
                    const float vz = __fsub_rn(scz, __fmul_rn(d.z, t));
                    const float d2 = lane_dot<(W != 1) && FMA>(vx, vy, vz, vx, vy, vz);
                    const bool exact = (W == 1) ? (!(d2 > s.w) && i < P.n_spheres) : (d2 < s.w);
                    uint32_t mask = 0;
                    filter_sphere(mask, 1u, P.scan_table[i], p2ox, p2oy, p2oz, fdx, fdy, fdz, od, f.oo, f.oo);
                    n_tests += 1;
                    n_hits += exact ? 1 : 0;
                    n_slab += slab ? 1 : 0;
                    n_cand += (slab && mask) ? 1 : 0;
                    n_viol += (exact && !(slab && mask)) ? 1 : 0;
                    if (bounce == 0 && !(bp.eps < 0.0f)) {             // a pixel with eps < 0 takes the per-ray scan
                        const bool bund = bundle_test(cp, cq, i < P.n_spheres ? slab_radius(s.w, s.x, s.y, s.z, P.omax) : kSlabPadRadius, P.cam.pos[ip], P.cam.pos[iq], bp);
                        n_btests += 1;
                        n_bcand += bund ? 1 : 0;
                        n_bviol += (exact && !bund) ? 1 : 0;
                    }
                }
                Hit h;
                if (W == 1) h = intersect_scalar(P.spheres, P.n_pad, P.n_spheres, o, d);
                else h = intersect_lanes<(W == 1 ? 4 : W), FMA>(P.spheres, P.n_pad, o, d);
                float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
                if (h.hit) s = P.spheres[h.prim];
                if (shade_scatter(P, h, s, o, d, atten, out, rng, bounce)) break;
            }
        }
    }
    for (int off = 16; off > 0; off >>= 1) {
        n_tests += __shfl_down_sync(0xFFFFFFFFu, n_tests, off);
        n_hits += __shfl_down_sync(0xFFFFFFFFu, n_hits, off);
        n_cand += __shfl_down_sync(0xFFFFFFFFu, n_cand, off);
        n_viol += __shfl_down_sync(0xFFFFFFFFu, n_viol, off);
        n_slab += __shfl_down_sync(0xFFFFFFFFu, n_slab, off);
        n_btests += __shfl_down_sync(0xFFFFFFFFu, n_btests, off);
        n_bcand += __shfl_down_sync(0xFFFFFFFFu, n_bcand, off);
        n_bviol += __shfl_down_sync(0xFFFFFFFFu, n_bviol, off);
    }
    if ((threadIdx.x & 31u) == 0) {
        atomicAdd(&counters[0], n_tests);
        atomicAdd(&counters[1], n_hits);
        atomicAdd(&counters[2], n_cand);
        atomicAdd(&counters[3], n_viol);
        atomicAdd(&counters[4], n_slab);
        atomicAdd(&counters[5], n_btests);
        atomicAdd(&counters[6], n_bcand);
        atomicAdd(&counters[7], n_bviol);
    }
}

// ------------------------------------------------------------------------------------------
// kat_kernel: diagnostic behind srt_device_kat().  Runs the device functions the render kernels are built
// from -- pcg_next / rand_float (base.h:954-989), stream_seed (main.cpp:667-676), normalize3
// (x64_math.h:234-245), reflectance (main.cpp:292-300), linear_to_srgb / pack_rgba8 (main.cpp:312-346) --
// on caller-provided inputs, one record per thread, so that the reference's known answers
// (tests/golden/ref_kat.json), edge values included, are checked on the GPU and not only through images.
// ------------------------------------------------------------------------------------------
__global__ void kat_kernel(uint32_t op, const void *in, uint32_t n, void *out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    switch (op) {
        case SRT_KAT_PCG: {                       // in: u64 seed -> out: 16 x u32
            uint64_t s = ((const uint64_t *)in)[i];
            for (int j = 0; j < 16; ++j) ((uint32_t *)out)[i * 16u + j] = pcg_next(s);
        } break;
        case SRT_KAT_RANDOM_FLOAT: {              // in: {u64 seed, f32 inverse_max_int, f32 min} -> out: 16 x f32
            const uint64_t *rec = (const uint64_t *)in + 2u * i;
            uint64_t s = rec[0];
            const float inv = __uint_as_float((uint32_t)rec[1]), mn = __uint_as_float((uint32_t)(rec[1] >> 32));
            for (int j = 0; j < 16; ++j) ((float *)out)[i * 16u + j] = rand_float(s, inv, mn);
        } break;
        case SRT_KAT_STREAM_SEED:                 // in: u64 index -> out: u64
            ((uint64_t *)out)[i] = stream_seed(((const uint64_t *)in)[i]);
            break;
        case SRT_KAT_NORMALIZE: {                 // in: 3 x f32 -> out: 3 x f32
            const float *v = (const float *)in + 3u * i;
            const f3 r = normalize3(mk3(v[0], v[1], v[2]));
            float *o = (float *)out + 3u * i;
            o[0] = r.x; o[1] = r.y; o[2] = r.z;
        } break;
        case SRT_KAT_REFLECTANCE: {               // in: {cos, eta} -> out: f32
            const float *v = (const float *)in + 2u * i;
            ((float *)out)[i] = reflectance(v[0], v[1]);
        } break;
        case SRT_KAT_LINEAR_TO_SRGB:              // in: f32 -> out: f32
            ((float *)out)[i] = linear_to_srgb(((const float *)in)[i]);
            break;
        case SRT_KAT_PACK: {                      // in: 4 x f32 {r, g, b, -} -> out: u32 RGBA8
            const float *v = (const float *)in + 4u * i;
            ((uint32_t *)out)[i] = pack_rgba8(v[0], v[1], v[2]);
        } break;
        default: break;
    }
}

// srt_write_accum on a partitioned handle: keep the rows this device owns (block row by = py / 4 with by % n_parts == part).
__global__ void clear_foreign_rows(float4 *accum, uint32_t width, uint32_t n_pixels, uint32_t part, uint32_t n_parts) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_pixels && ((i / width) >> 2) % n_parts != part) accum[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
}

// mean = sum * (1/N); tonemap + pack.  HBM-bound: 16 B read + 16 B + 4 B written per pixel.
// Two adjacent pixels per thread: both 16-byte loads are issued before anything depends on them
// (a warp reads 1 KiB contiguous), the two RGBA8 words go out as one 8-byte store.
__global__ void __launch_bounds__(256) resolve_sum_kernel(float4 *__restrict__ accum, uint32_t *__restrict__ rgba, uint32_t n_pixels, float inv_n) {
    const uint32_t i = 2u * (blockIdx.x * blockDim.x + threadIdx.x);
    if (i + 1u < n_pixels) {
        const float4 s0 = __ldcs(accum + i), s1 = __ldcs(accum + i + 1u);
        const float r0 = __fmul_rn(s0.x, inv_n), g0 = __fmul_rn(s0.y, inv_n), b0 = __fmul_rn(s0.z, inv_n);
        const float r1 = __fmul_rn(s1.x, inv_n), g1 = __fmul_rn(s1.y, inv_n), b1 = __fmul_rn(s1.z, inv_n);
        accum[i] = make_float4(r0, g0, b0, 1.0f);
        accum[i + 1u] = make_float4(r1, g1, b1, 1.0f);
        *reinterpret_cast<uint2 *>(rgba + i) = make_uint2(pack_rgba8(r0, g0, b0), pack_rgba8(r1, g1, b1));
    } else if (i < n_pixels) {
        const float4 s = accum[i];
        const float r = __fmul_rn(s.x, inv_n), g = __fmul_rn(s.y, inv_n), b = __fmul_rn(s.z, inv_n);
        accum[i] = make_float4(r, g, b, 1.0f);
        rgba[i] = pack_rgba8(r, g, b);
    }
}
// one pixel per thread (SRT_RESOLVE_VARIANT=1; kept for A/B measurements)
__global__ void __launch_bounds__(256) resolve_sum_kernel_1px(float4 *accum, uint32_t *rgba, uint32_t n_pixels, float inv_n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pixels) return;
    const float4 s = accum[i];
    const float r = __fmul_rn(s.x, inv_n), g = __fmul_rn(s.y, inv_n), b = __fmul_rn(s.z, inv_n);
    accum[i] = make_float4(r, g, b, 1.0f);
    rgba[i] = pack_rgba8(r, g, b);
}

} // namespace srt

// =============================================================================================
// Host side: the C ABI
// =============================================================================================
using namespace srt;

// Diagnostic knobs (environment), read ONCE per srt_set_scene -- never on the per-frame path.
struct Knobs {
    int variant = 0;                     // SRT_KERNEL_VARIANT: 0 auto (pool kernel from 64 spheres, else one path per thread);
                                         // 1 scalar-issue loop; 7 force one path per thread; 9 force the pool kernel; 12 force
                                         // the STREAMED pool kernel
    int slab_axis = -1;                  // SRT_SLAB_AXIS: force the slab axis (0/1/2)
    uint32_t pool_chunk = 0;             // SRT_POOL_CHUNK: samples per work item of the non-bundle pool kernel, 0 = kPoolChunkSamples
    int bundle = 1;                      // SRT_BUNDLE: 1 auto (>= kBundleMinSamples samples, >= kBundleMinSpheres spheres), 0 off,
                                         // k >= 2 forced with chunks of <= 2^k samples
    int regen_threshold = 0;             // SRT_REGEN_THRESHOLD: 0 = automatic
    int pool_threads = 0;                // SRT_POOL_THREADS: force the resident pool configuration (896 / 512 / 256)
    bool pool_debug = false;             // SRT_POOL_DEBUG: print ring statistics after every launch (synchronous)
    uint32_t band_rows = 0;              // SRT_BUNDLE_BAND_ROWS: block rows per staging band (tests: tiny bands)
    size_t stage_limit = ~(size_t)0;     // SRT_BUNDLE_STAGE_LIMIT: pretend larger staging allocations fail (tests)
    int resolve_variant = 0;             // SRT_RESOLVE_VARIANT: 1 = one pixel per thread
};
static Knobs read_knobs() {
    Knobs k;
    if (const char *v = getenv("SRT_KERNEL_VARIANT")) k.variant = atoi(v);
    if (const char *v = getenv("SRT_SLAB_AXIS")) k.slab_axis = std::min(2, std::max(0, atoi(v)));
    if (const char *v = getenv("SRT_POOL_CHUNK")) k.pool_chunk = (uint32_t)std::max(0, atoi(v));
    if (const char *v = getenv("SRT_BUNDLE")) k.bundle = atoi(v);
    if (const char *v = getenv("SRT_REGEN_THRESHOLD")) k.regen_threshold = std::min(32, std::max(1, atoi(v)));
    if (const char *v = getenv("SRT_POOL_THREADS")) k.pool_threads = atoi(v);
    k.pool_debug = getenv("SRT_POOL_DEBUG") != nullptr;
    if (const char *v = getenv("SRT_BUNDLE_BAND_ROWS")) k.band_rows = (uint32_t)std::max(1, atoi(v));
    if (const char *v = getenv("SRT_BUNDLE_STAGE_LIMIT")) k.stage_limit = (size_t)atoll(v);
    if (const char *v = getenv("SRT_RESOLVE_VARIANT")) k.resolve_variant = atoi(v);
    return k;
}

typedef void (*render_kernel_t)(const KernelParams);
// Launch configuration of one (kernel, block size, shared memory) triple: the attribute calls and the occupancy query
// are made once, not per frame (the per-frame kick replaces WorkQueueStart, main.cpp:852 -- it has to be cheap).
struct LaunchCfg { render_kernel_t kernel; int threads; size_t smem; int per_sm; };

#ifdef SRT_SPIN_LIMIT
static unsigned long long *g_spin_report_host = nullptr;   // host address of the word g_spin_report (device) points at
#endif
struct srt_backend {
    int device = -1;
    int sm_count = 0, clock_khz = 0;
    size_t max_smem_optin = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_start = nullptr, ev_stop = nullptr;
    bool events_recorded = false;
    // scene
    float4 *d_spheres = nullptr;         // {cx,cy,cz,r2} per sphere (scalar-issue kernels)
    float4 *d_spheres2 = nullptr;        // sphere-pair layout (packed f32x2 kernels)
    float4 *d_scan = nullptr;            // filter table of the pool kernel {cx,cy,cz,-kappa}
    float4 *d_ptab = nullptr;            // slab table of streamed scenes (rebuilt per launch)
    DevMaterial *d_materials = nullptr;
    Knobs knobs;
    uint32_t regen_threshold = 1;
    uint32_t n_spheres = 0, n_pad = 0;
    srt_scene_options opt = {0, 8, 1, 0};
    // images
    uint32_t width = 0, height = 0;
    float4 *d_accum_own = nullptr, *d_accum = nullptr;
    uint32_t *d_rgba_own = nullptr, *d_rgba = nullptr;
    uint32_t *d_tap_prim = nullptr;
    float *d_tap_t = nullptr;
    bool tap_valid = false;
    // image partition over devices (srt_set_partition): this handle renders block rows by = part + k * n_parts
    uint32_t part = 0, n_parts = 1;
    // counters: [0..7] rays (u64), [8..11] work counter, [12..15] exit ticket, [16..63] diagnostics.  The last CTA of a
    // launch publishes the ray count to pinned host memory and re-arms the counters (finish_launch): no memset / copy per frame.
    unsigned char *d_counters = nullptr;
    bool counters_dirty = true;          // first launch, or a failed one: memset before the next launch
    uint32_t *d_done = nullptr;          // pool kernel: per-pixel chunk counters (KernelParams::done)
    uint32_t *d_static_done = nullptr;   // device address of the file-scope scratch of srt_render_static (resolve_static_scratch)
    uint32_t pool_seq = 0;               // pool launches issued: done-tag sequence (20 bits, 0xFFFFF reserved for the 0xFF fill)
    int slab_axis = 1;                   // axis contained in the slab planes: least spread of the centres
    double omax_scene = 0.0;             // >= |C| + r of every sphere (bound on secondary-ray origins)
    float4 *d_stage = nullptr;           // bundle mode: staging rows of the pixels in flight (KernelParams::stage)
    uint32_t *d_cnt = nullptr;           // bundle mode: samples staged per row
    size_t stage_bytes = 0, cnt_rows = 0;
    float2 *d_fold_w = nullptr;          // bundle mode: running-mean weights of the launch's samples
    uint32_t fold_w_cap = 0;
    uint32_t fold_w_first = 0, fold_w_n = 0, fold_w_stride = 0;   // what d_fold_w currently holds
    std::vector<float2> h_fold_w;
    unsigned long long *h_rays = nullptr;   // pinned + mapped: written by the last CTA of a launch
    bool in_flight = false;
    uint64_t launches = 0;
    std::string last_kernel;             // the render kernel of the last srt_render / srt_render_static, as ncu prints it
    // multi-device (srt_comm_*): NCCL communicator of this handle, gather buffers on the root
    void *comm = nullptr;
    int comm_rank = 0, comm_size = 1;
    uint32_t *d_gather_rgba = nullptr;
    float4 *d_gather_accum = nullptr;
    bool gathered_rgba = false, gathered_accum = false;
    std::string err;
};

static std::string g_create_error;

static int fail(srt_backend *b, int code, const std::string &msg) {
    if (b) b->err = msg; else g_create_error = msg;
    return code;
}
#define SRT_CUDA(b, call)                                                                       \
    do {                                                                                        \
        cudaError_t e_ = (call);                                                                \
        if (e_ != cudaSuccess)                                                                  \
            return fail((b), SRT_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

// One path per thread (render_paths): tiny scenes are shading- and latency-bound, more resident warps beat the
// 104-register configuration.  Demo scene, 1080p: scalar-issue loop at 4 CTAs / SM 32.7 Grays/s at 64 spp and 0.139 ms at
// 1 spp; packed loop at 3 CTAs / SM 31.4 and 0.147.  (A packed body must cover whole SIMD groups: 2 * PAIRS % W == 0.)
static render_kernel_t pick_kernel(const srt_scene_options &o, bool packed) {
#ifdef SRT_FAST_BUILD                       // development / debug builds: only the lane_width 8 + FMA profile is compiled
    if (o.lane_width != 8 || !o.fma_dot) return nullptr;
    return packed ? render_paths<8, true, true, 4, 3> : render_paths<8, true, false, 4, 4>;
#else
    if (o.lane_width == 1) return render_paths<1, false, false, 4, 4>;
    if (packed) {
        if (o.lane_width == 4) return o.fma_dot ? render_paths<4, true, true, 2, 4> : render_paths<4, false, true, 2, 4>;
        return o.fma_dot ? render_paths<8, true, true, 4, 3> : render_paths<8, false, true, 4, 3>;
    }
    if (o.lane_width == 4) return o.fma_dot ? render_paths<4, true, false, 4, 4> : render_paths<4, false, false, 4, 4>;
    return o.fma_dot ? render_paths<8, true, false, 4, 4> : render_paths<8, false, false, 4, 4>;
#endif
}

// Pool kernel (srt_pool.cuh): CTA-owned path slots, primary / secondary warp passes; used from kPoolMinSpheres
// spheres, where the intersection loop dominates.  Resident tables: one large CTA per SM -- the refine / select /
// shade phases of a pass are latency-bound, the slab scan is issue-bound, and only warps in the scan keep the issue
// port busy: 28 warps (72 registers) beat 16 warps (2 x 256 threads) by 24 % on the 1,024-sphere scene.  Larger
// scenes trade warps for table space; beyond that the tables are streamed (2 CTAs of 256 threads).
constexpr uint32_t kPoolMinSpheres = 64;
constexpr int kPoolThreads = 256;           // streamed kernel: 2 CTAs of 256 threads
constexpr int kPoolMinBlocks = 2;
constexpr uint32_t kPoolChunkSamples = 1;   // samples per pool work item (measured best: 1; SRT_POOL_CHUNK overrides)
constexpr int kPoolResidentThreads[3] = {896, 512, 256};
template <int THREADS, int MINB, bool STREAM, bool BUNDLE>
static render_kernel_t pick_pool_kernel_cfg(const srt_scene_options &o) {
#ifdef SRT_FAST_BUILD
    if (o.lane_width != 8 || !o.fma_dot) return nullptr;
    return render_pool<8, true, THREADS, MINB, STREAM, BUNDLE>;
#else
    if (o.lane_width == 1) return render_pool<1, false, THREADS, MINB, STREAM, BUNDLE>;
    if (o.lane_width == 4) return o.fma_dot ? render_pool<4, true, THREADS, MINB, STREAM, BUNDLE> : render_pool<4, false, THREADS, MINB, STREAM, BUNDLE>;
    return o.fma_dot ? render_pool<8, true, THREADS, MINB, STREAM, BUNDLE> : render_pool<8, false, THREADS, MINB, STREAM, BUNDLE>;
#endif
}
template <bool BUNDLE>
static render_kernel_t pick_pool_kernel_b(const srt_scene_options &o, bool stream, int threads) {
    if (stream) return pick_pool_kernel_cfg<kPoolThreads, kPoolMinBlocks, true, BUNDLE>(o);
    if (threads == 896) return pick_pool_kernel_cfg<896, 1, false, BUNDLE>(o);
    if (threads == 512) return pick_pool_kernel_cfg<512, 1, false, BUNDLE>(o);
    return pick_pool_kernel_cfg<256, 1, false, BUNDLE>(o);
}
static render_kernel_t pick_pool_kernel(const srt_scene_options &o, bool stream, int threads, bool bundle) {
    return bundle ? pick_pool_kernel_b<true>(o, stream, threads) : pick_pool_kernel_b<false>(o, stream, threads);
}
static size_t pool_smem_bytes(uint32_t n_pad, int threads, bool stream, bool bundle = false) {
    const size_t ns = 64u * (threads / 32 + kPoolSpareBatches), rcap = ns <= 1024 ? 1024 : 2048, warps = threads / 32;
    size_t bytes = (size_t)(bundle ? kBundleFields : S_COUNT) * ns * sizeof(float) + (bundle ? 3 : 2) * rcap * sizeof(ring_t) + ((sizeof(PoolCtl) + 15u) & ~(size_t)15u);
    if (stream) bytes += warps * kStreamStages * (kStreamStageF4 * sizeof(float4) + sizeof(uint64_t));
    else bytes += (size_t)n_pad * (sizeof(float4) + 12 + sizeof(float));   // {c,-kappa}, slab table, r2
    return bytes;
}

// Bundle mode (srt_bundle.cuh).
constexpr uint32_t kBundleMinSamples = 8;        // launches with fewer samples keep (chunk, pixel) work items
constexpr uint64_t kBundleStageBytes = 1ull << 30;   // staging rows of one band (16 B x samples per chunk x pixels of the band)
constexpr uint32_t kBundleMinSpheres = 384;      // measured (1080p, 32 spp): 256 spheres +1 %, 482 +9 %, 1,024 +18 %, 65,536 +57 %
static MagicDiv make_magic(uint32_t d) {          // srt_device.cuh: x / d for x < 2^30
    uint32_t l = 0;
    while ((1ull << l) < d) ++l;
    MagicDiv m;
    m.m = (uint32_t)((((uint64_t)1 << (30 + l)) + d - 1) / d);
    m.sh = 30 + l;
    return m;
}    // work items per band and chunk: 2 Mi >> paths in flight (148 x 1,920)
// rho >= |F(j)_pq - F(0)_pq| for every jitter j in [-0.5, 0.5]^2 of any pixel (film units, kept axes p, q), plus the
// fp32 rounding of both film-point evaluations (main.cpp:378-384: <= u (2 |fc_i| + 12) per component).  0 = do not bundle.
static double bundle_footprint(const srt_camera *cam, uint32_t width, uint32_t height, int axis) {
    const int ip = axis == 0 ? 1 : (axis == 1 ? 2 : 0), iq = axis == 0 ? 2 : (axis == 1 ? 0 : 1);
    const double fw = cam->film_w, fh = cam->film_h;
    double nx = 0, ny = 0, mag = 0;
    for (int i = 0; i < 3; ++i) {
        nx += (double)cam->camera_x[i] * cam->camera_x[i];
        ny += (double)cam->camera_y[i] * cam->camera_y[i];
        mag = std::max({mag, fabs((double)cam->film_center[i]), fabs((double)cam->camera_position[i])});
    }
    if (!(fw > 0 && fw <= 4 && fh > 0 && fh <= 4 && nx <= 4 && ny <= 4 && mag < 1e6)) return 0.0;   // NaN fails too
    const double dax = fw / (2.0 * width), day = fh / (2.0 * height);
    const double xp = cam->camera_x[ip], xq = cam->camera_x[iq], yp = cam->camera_y[ip], yq = cam->camera_y[iq];
    const double a2 = dax * dax * (xp * xp + xq * xq), b2 = day * day * (yp * yp + yq * yq), cross = 2.0 * dax * day * fabs(xp * yp + xq * yq);
    const double u = 5.9604644775390625e-08;
    const double fp_err = 8.0 * u * (fabs((double)cam->film_center[ip]) + fabs((double)cam->film_center[iq]) +
                                     fabs((double)cam->camera_position[ip]) + fabs((double)cam->camera_position[iq]) + 16.0);
    return sqrt(a2 + b2 + cross) * 1.001 + fp_err;
}

static void comm_release(srt_backend *b) {
    if (b->comm) {
        NcclApi &N = nccl_api();
        if (N.CommDestroy) N.CommDestroy((ncclComm_t)b->comm);
        b->comm = nullptr;
    }
    b->comm_rank = 0;
    b->comm_size = 1;
}
#define SRT_NCCL(b, call)                                                                                         \
    do {                                                                                                          \
        ncclResult_t r_ = (call);                                                                                 \
        if (r_ != ncclSuccess)                                                                                    \
            return fail((b), SRT_ERR_COMM, std::string(#call) + ": " + (nccl_api().GetErrorString ? nccl_api().GetErrorString(r_) : "NCCL error")); \
    } while (0)

extern "C" {

const char *srt_last_error(const srt_backend *b) { return b ? b->err.c_str() : g_create_error.c_str(); }

// ---- srt_render_static: scratch without cudaMalloc ----------------------------------------------------------
// The pool kernel's per-pixel chunk counters (KernelParams::done) are scratch memory that srt_render allocates on its
// first pool launch.  Callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
// that forbid it) use srt_render_static, whose scratch is this file-scope array: kStaticDonePixels x 4 B = 8 MiB, held by
// every GPU that loads the library until the process exits.  In HOST code `g_static_done` is only a stand-in for the
// symbol, not a device address: the address is taken once per device with cudaGetSymbolAddress (that device current) and
// that pointer is what the kernel gets.
constexpr uint32_t kStaticDonePixels = 2u << 20;             // 2,097,152 pixels: up to 1920 x 1080 (and 2048 x 1024)
static __device__ __align__(256) uint32_t g_static_done[kStaticDonePixels];
constexpr int kStaticMaxDevices = 64;
static std::mutex g_static_mu;
static uint32_t *g_static_done_ptr[kStaticMaxDevices];       // device address of g_static_done on each device, once resolved
static int resolve_static_scratch(srt_backend *b) {          // the caller has made b->device current
    if (b->device < 0 || b->device >= kStaticMaxDevices) return SRT_OK;     // srt_render_static will refuse; everything else works
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_done_ptr[b->device]) {
        uint32_t *ptr = nullptr;
        SRT_CUDA(b, cudaGetSymbolAddress((void **)&ptr, g_static_done));
        g_static_done_ptr[b->device] = ptr;
    }
    b->d_static_done = g_static_done_ptr[b->device];
    return SRT_OK;
}

static int create_resources(srt_backend *b) {
    SRT_CUDA(b, cudaSetDevice(b->device));
    SRT_CUDA(b, cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking));
    SRT_CUDA(b, cudaEventCreate(&b->ev_start));
    SRT_CUDA(b, cudaEventCreate(&b->ev_stop));
    SRT_CUDA(b, cudaMalloc(&b->d_counters, 64));
    SRT_CUDA(b, cudaMemset(b->d_counters, 0, 64));
    SRT_CUDA(b, cudaHostAlloc(&b->h_rays, 4 * sizeof(unsigned long long), cudaHostAllocMapped | cudaHostAllocPortable));
    b->h_rays[0] = 0;                    // [0] rays of the last launch, [1] spin-limit report of the debug build
    b->h_rays[1] = 0;
    b->h_rays[2] = 0;                    // [2] value attached to an index-check report
    {
        const int rc = resolve_static_scratch(b);            // set-up of srt_render_static: once per device, this device current
        if (rc != SRT_OK) return rc;
    }
#ifdef SRT_SPIN_LIMIT
    {
        unsigned long long *report = b->h_rays + 1;
        SRT_CUDA(b, cudaMemcpyToSymbol(g_spin_report, &report, sizeof(report)));
        g_spin_report_host = report;         // the device symbol is per device, not per handle: the newest handle's word takes every report
    }
#endif
    return SRT_OK;
}

int srt_create(srt_backend **out, int device_id) {
    if (!out) return fail(nullptr, SRT_ERR_INVALID, "srt_create: out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(nullptr, SRT_ERR_NO_DEVICE,
                    std::string("srt_create: no CUDA device (") + cudaGetErrorString(e) + "); this backend has no CPU fallback");
    if (device_id < 0 || device_id >= count) return fail(nullptr, SRT_ERR_INVALID, "srt_create: bad device ordinal");
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device_id) != cudaSuccess) return fail(nullptr, SRT_ERR_CUDA, "cudaGetDeviceProperties failed");
    if (prop.major < 10)
        return fail(nullptr, SRT_ERR_NO_DEVICE, std::string("srt_create: device '") + prop.name + "' is not sm_100 class; kernels are built for sm_100a only");
    srt_backend *b = new srt_backend();
    b->device = device_id;
    b->sm_count = prop.multiProcessorCount;
    b->clock_khz = prop.clockRate;
    b->max_smem_optin = prop.sharedMemPerBlockOptin;
    b->knobs = read_knobs();
    const int rc = create_resources(b);
    if (rc != SRT_OK) {                  // no half-built handle leaves this function: *out stays NULL, the reason is srt_last_error(NULL)
        g_create_error = "srt_create: " + b->err;
        srt_destroy(b);
        return rc;
    }
    *out = b;
    return SRT_OK;
}

void srt_destroy(srt_backend *b) {
    if (!b) return;
    cudaSetDevice(b->device);
    if (b->stream) cudaStreamSynchronize(b->stream);
    comm_release(b);
    cudaFree(b->d_spheres); cudaFree(b->d_spheres2); cudaFree(b->d_scan); cudaFree(b->d_ptab); cudaFree(b->d_materials);
    cudaFree(b->d_accum_own); cudaFree(b->d_rgba_own);
    cudaFree(b->d_tap_prim); cudaFree(b->d_tap_t);
    cudaFree(b->d_counters);
    cudaFree(b->d_done); cudaFree(b->d_stage); cudaFree(b->d_cnt); cudaFree(b->d_fold_w);
    cudaFree(b->d_gather_rgba); cudaFree(b->d_gather_accum);
    if (b->h_rays) cudaFreeHost(b->h_rays);
    if (b->ev_start) cudaEventDestroy(b->ev_start);
    if (b->ev_stop) cudaEventDestroy(b->ev_stop);
    if (b->stream) cudaStreamDestroy(b->stream);
    delete b;
}

static int drain(srt_backend *b) {
    if (b->in_flight) {
        SRT_CUDA(b, cudaStreamSynchronize(b->stream));
        b->in_flight = false;
    }
    return SRT_OK;
}

int srt_set_scene(srt_backend *b, const srt_sphere *spheres, uint32_t n_spheres, const srt_scene_options *options) {
    if (!b) return SRT_ERR_INVALID;
    if (!spheres || n_spheres == 0 || !options) return fail(b, SRT_ERR_INVALID, "srt_set_scene: null/empty scene");
    if (options->lane_width != 1 && options->lane_width != 4 && options->lane_width != 8)
        return fail(b, SRT_ERR_INVALID, "srt_set_scene: lane_width must be 1, 4 or 8");
    SRT_CUDA(b, cudaSetDevice(b->device));
    int rc = drain(b);
    if (rc) return rc;
    const uint32_t n_pad = (n_spheres + 7u) & ~7u;
    if (n_spheres > (1u << 24)) return fail(b, SRT_ERR_TOO_LARGE, "srt_set_scene: more than 2^24 spheres");
    std::vector<float4> sph(n_pad, make_float4(0.f, 0.f, 0.f, 0.f));   // zero spheres: d2 < 0 never holds
    std::vector<DevMaterial> mats(n_pad);
    memset(mats.data(), 0, mats.size() * sizeof(DevMaterial));
    for (uint32_t i = 0; i < n_spheres; ++i) {
        const srt_sphere &s = spheres[i];
        const volatile float r = s.radius;
        const volatile float r2 = r * r;                                  // main.cpp:406, one IEEE f32 multiply
        sph[i] = make_float4(s.position[0], s.position[1], s.position[2], r2);
        const srt_material &m = s.material;
        mats[i].color_spec = make_float4(m.color[0], m.color[1], m.color[2], m.specular);
        mats[i].emis_ior = make_float4(m.emissive[0], m.emissive[1], m.emissive[2], m.index_of_refraction);
        const float oms = (float)(1.0 - (double)m.specular);              // main.cpp:463
        const volatile float one = 1.0f;
        const volatile float inv_ior = one / m.index_of_refraction;       // main.cpp:466 (inf when IOR == 0: unused)
        mats[i].derived = make_float4(oms, inv_ior, 0.f, 0.f);
    }
    std::vector<float4> sph2(n_pad);                                     // pair layout, srt_packed.cuh
    for (uint32_t p = 0; p < n_pad / 2; ++p) {
        const float4 &e = sph[2 * p], &o = sph[2 * p + 1];
        sph2[2 * p] = make_float4(e.x, o.x, e.y, o.y);
        sph2[2 * p + 1] = make_float4(e.z, o.z, e.w, o.w);
    }
    // Filter table of the ray-pair kernel (srt_pairs.cuh): -kappa_i = R_i - |C_i|^2
    // with R_i = r2 (1 + 128u) + 128u |C|^2, evaluated in double.  Padding can never be a candidate.
    std::vector<float4> scan(n_pad, make_float4(0.f, 0.f, 0.f, -1.0e30f));
    for (uint32_t i = 0; i < n_spheres; ++i) {
        const double u = 5.9604644775390625e-08;
        const double cx = sph[i].x, cy = sph[i].y, cz = sph[i].z, r2 = sph[i].w;
        const double c2 = cx * cx + cy * cy + cz * cz;
        const double R = r2 * (1.0 + 128.0 * u) + 128.0 * u * c2 + 1.0e-30;
        scan[i] = make_float4(sph[i].x, sph[i].y, sph[i].z, (float)(R - c2));
    }
    // Slab filter (srt_pairs.cuh): the slab planes contain the coordinate axis along which the centres
    // spread least (ignoring spheres much larger than the median one, e.g. a ground sphere); Omax bounds
    // the origins of secondary rays, which lie on sphere surfaces.
    {
        std::vector<double> radii(n_spheres);
        for (uint32_t i = 0; i < n_spheres; ++i) radii[i] = sqrt((double)sph[i].w);
        std::vector<double> sorted = radii;
        std::nth_element(sorted.begin(), sorted.begin() + n_spheres / 2, sorted.end());
        const double cutoff = 10.0 * sorted[n_spheres / 2];
        double sum[3] = {0, 0, 0}, sq[3] = {0, 0, 0}, cnt = 0, omax = 0;
        for (uint32_t i = 0; i < n_spheres; ++i) {
            const double c[3] = {sph[i].x, sph[i].y, sph[i].z};
            omax = std::max(omax, sqrt(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]) + radii[i]);
            if (radii[i] > cutoff) continue;
            for (int a = 0; a < 3; ++a) { sum[a] += c[a]; sq[a] += c[a] * c[a]; }
            cnt += 1;
        }
        int axis = 1;
        if (cnt >= 2) {
            double best = 0;
            for (int a = 0; a < 3; ++a) {
                const double var = sq[a] / cnt - (sum[a] / cnt) * (sum[a] / cnt);
                if (a == 0 || var < best) { best = var; axis = a; }
            }
        }
        b->knobs = read_knobs();                 // the only place the SRT_* diagnostics are read after srt_create
        if (b->knobs.variant == 0) b->knobs.variant = options->_reserved;
        b->slab_axis = b->knobs.slab_axis >= 0 ? b->knobs.slab_axis : axis;
        b->omax_scene = omax * 1.01 + 1e-3;
    }
    cudaFree(b->d_spheres); b->d_spheres = nullptr;
    cudaFree(b->d_scan); b->d_scan = nullptr;
    cudaFree(b->d_ptab); b->d_ptab = nullptr;
    cudaFree(b->d_spheres2); b->d_spheres2 = nullptr;
    cudaFree(b->d_materials); b->d_materials = nullptr;
    SRT_CUDA(b, cudaMalloc(&b->d_spheres, sph.size() * sizeof(float4)));
    SRT_CUDA(b, cudaMalloc(&b->d_scan, scan.size() * sizeof(float4)));
    SRT_CUDA(b, cudaMalloc(&b->d_ptab, scan.size() * sizeof(float4)));
    SRT_CUDA(b, cudaMemcpy(b->d_scan, scan.data(), scan.size() * sizeof(float4), cudaMemcpyHostToDevice));
    SRT_CUDA(b, cudaMalloc(&b->d_spheres2, sph2.size() * sizeof(float4)));
    SRT_CUDA(b, cudaMemcpy(b->d_spheres2, sph2.data(), sph2.size() * sizeof(float4), cudaMemcpyHostToDevice));
    SRT_CUDA(b, cudaMalloc(&b->d_materials, mats.size() * sizeof(DevMaterial)));
    SRT_CUDA(b, cudaMemcpy(b->d_spheres, sph.data(), sph.size() * sizeof(float4), cudaMemcpyHostToDevice));
    SRT_CUDA(b, cudaMemcpy(b->d_materials, mats.data(), mats.size() * sizeof(DevMaterial), cudaMemcpyHostToDevice));
    b->n_spheres = n_spheres;
    b->n_pad = n_pad;
    b->opt = *options;
    // Path start-up costs the whole warp ~200 issue slots however few lanes need it.  For
    // tiny scenes (intersection pass cheaper than that) a warp therefore refills only once
    // all its lanes are idle, which also keeps its primary rays coherent; measured on B200:
    // scene 0 (5 spheres) 2.11 ms with 32 vs 2.43 ms with 1; 256+ spheres: 1 is best.
    b->regen_threshold = n_pad <= 16 ? 32u : (n_pad <= 32 ? 8u : 1u);
    if (b->knobs.regen_threshold) b->regen_threshold = (uint32_t)b->knobs.regen_threshold;
    return SRT_OK;
}

int srt_reset_accumulation(srt_backend *b) {
    if (!b) return SRT_ERR_INVALID;
    if (!b->d_accum) return fail(b, SRT_ERR_INVALID, "srt_reset_accumulation: no image (call srt_resize)");
    SRT_CUDA(b, cudaSetDevice(b->device));
    const size_t n = (size_t)b->width * b->height;
    SRT_CUDA(b, cudaMemsetAsync(b->d_accum, 0, n * sizeof(float4), b->stream));
    SRT_CUDA(b, cudaMemsetAsync(b->d_rgba, 0, n * sizeof(uint32_t), b->stream));
    b->tap_valid = false;
    b->gathered_rgba = b->gathered_accum = false;
    return SRT_OK;
}

int srt_resize(srt_backend *b, uint32_t width, uint32_t height) {
    if (!b) return SRT_ERR_INVALID;
    if (width == 0 || height == 0) return fail(b, SRT_ERR_INVALID, "srt_resize: zero size");
    // pixel coordinates travel as px | py << 16 in the path slots, pixel indices as 32-bit words
    if (width > 65535u || height > 65535u || (uint64_t)width * height > 0x7FFFFFFFull)
        return fail(b, SRT_ERR_TOO_LARGE, "srt_resize: image too large (width, height <= 65535 and width * height < 2^31)");
    SRT_CUDA(b, cudaSetDevice(b->device));
    int rc = drain(b);
    if (rc) return rc;
    const size_t n = (size_t)width * height;
    if (width != b->width || height != b->height || !b->d_accum_own) {
        cudaFree(b->d_accum_own); cudaFree(b->d_rgba_own); cudaFree(b->d_tap_prim); cudaFree(b->d_tap_t); cudaFree(b->d_done);
        cudaFree(b->d_gather_rgba); cudaFree(b->d_gather_accum);
        b->d_gather_rgba = nullptr; b->d_gather_accum = nullptr;
        b->d_done = nullptr;
        b->d_accum_own = nullptr; b->d_rgba_own = nullptr; b->d_tap_prim = nullptr; b->d_tap_t = nullptr;
        SRT_CUDA(b, cudaMalloc(&b->d_accum_own, n * sizeof(float4)));
        SRT_CUDA(b, cudaMalloc(&b->d_rgba_own, n * sizeof(uint32_t)));
        b->width = width;
        b->height = height;
    }
    b->d_accum = b->d_accum_own;
    b->d_rgba = b->d_rgba_own;
    return srt_reset_accumulation(b);
}

int srt_bind_device_buffers(srt_backend *b, void *accum_dev, void *rgba_dev) {
    if (!b) return SRT_ERR_INVALID;
    if (!b->d_accum_own) return fail(b, SRT_ERR_INVALID, "srt_bind_device_buffers: call srt_resize first");
    int rc = drain(b);
    if (rc) return rc;
    b->d_accum = accum_dev ? (float4 *)accum_dev : b->d_accum_own;
    b->d_rgba = rgba_dev ? (uint32_t *)rgba_dev : b->d_rgba_own;
    return SRT_OK;
}

// Argument checks + the fields every kernel shares.
static int make_params(srt_backend *b, const srt_camera *cam, const srt_render_args *args, const char *who, KernelParams &P) {
    const std::string w(who);
    if (!cam || !args) return fail(b, SRT_ERR_INVALID, w + ": null argument");
    if (!b->d_spheres) return fail(b, SRT_ERR_INVALID, w + ": no scene (call srt_set_scene)");
    if (!b->d_accum) return fail(b, SRT_ERR_INVALID, w + ": no image (call srt_resize)");
    if (args->n_samples == 0 || args->sample_stride == 0) return fail(b, SRT_ERR_INVALID, w + ": n_samples and sample_stride must be >= 1");
    if (args->accum_mode != SRT_ACCUM_MEAN && args->accum_mode != SRT_ACCUM_SUM) return fail(b, SRT_ERR_INVALID, w + ": bad accum_mode");
    if (args->accum_mode == SRT_ACCUM_MEAN && args->sample_stride != 1)
        return fail(b, SRT_ERR_INVALID, w + ": the exact running mean needs contiguous samples (sample_stride == 1)");
    if ((uint64_t)args->first_sample + (uint64_t)args->n_samples * args->sample_stride > 0xFFFFFFFFull)
        return fail(b, SRT_ERR_INVALID, w + ": sample index overflows 32 bits");
    memset(&P, 0, sizeof(P));
    P.spheres = b->d_spheres;
    P.scan_table = b->d_scan;
    P.regen_threshold = b->regen_threshold;
    if (b->n_pad <= 16u && args->n_samples <= 2u && !b->knobs.regen_threshold) P.regen_threshold = 16u;   // 1 spp / frame: 0.139 vs 0.158 ms
    P.neg_one = -1.0f;
    P.one = 1.0f;
    P.materials = b->d_materials;
    P.n_pad = b->n_pad;
    P.n_spheres = b->n_spheres;
    P.use_sky = b->opt.use_sky_color;
    for (int i = 0; i < 3; ++i) {
        P.cam.pos[i] = cam->camera_position[i];
        P.cam.x[i] = cam->camera_x[i];
        P.cam.y[i] = cam->camera_y[i];
        P.cam.film_center[i] = cam->film_center[i];
    }
    P.cam.film_w = cam->film_w;
    P.cam.film_h = cam->film_h;
    P.width = b->width;
    P.height = b->height;
    P.blocks_x = (b->width + 7u) / 8u;
    P.div_bx = make_magic(P.blocks_x);
    P.two_over_w = 2.0f / (float)b->width;
    P.two_over_h = 2.0f / (float)b->height;
    // this handle's share of the image: block rows by = part + k * n_parts (srt_set_partition; the whole image by default)
    const uint32_t blocks_y = (b->height + 3u) / 4u;
    P.part = b->part;
    P.n_parts = b->n_parts;
    P.total_items = P.blocks_x * (blocks_y > b->part ? (blocks_y - b->part + b->n_parts - 1u) / b->n_parts : 0u) * 32u;
    P.first_sample = args->first_sample;
    P.n_samples = args->n_samples;
    P.sample_stride = args->sample_stride;
    P.accum_mode = args->accum_mode;
    P.stream_base = args->stream_base;
    P.accum = b->d_accum;
    P.rgba = b->d_rgba;
    P.rays = (unsigned long long *)b->d_counters;
    P.work_counter = (unsigned int *)(b->d_counters + 8);
    P.ticket = (unsigned int *)(b->d_counters + 12);
    P.host_rays = b->h_rays;
    P.debug = (unsigned long long *)(b->d_counters + 16);
    P.items_per_gen = P.total_items;
    P.chunk_span = 0;
    // slab filter (srt_pairs.cuh): bound on the ray origins of this launch = the scene's bound or the camera
    const double cam_len = sqrt((double)cam->camera_position[0] * cam->camera_position[0] + (double)cam->camera_position[1] * cam->camera_position[1] +
                                (double)cam->camera_position[2] * cam->camera_position[2]);
    P.slab_axis = b->slab_axis;
    P.omax = (float)(std::max(b->omax_scene, cam_len * 1.001 + 1e-3) * 1.0001);
    P.omax2 = P.omax * P.omax;
    return SRT_OK;
}

// Name of a render kernel instantiation the way ncu / cuobjdump print it (bench.py matches its profile against it).
static std::string kernel_label(const srt_scene_options &o, bool pool, bool stream, int pool_threads, bool bundle, bool packed) {
    const int w = o.lane_width, fma = (w != 1 && o.fma_dot) ? 1 : 0;
    char buf[96];
    if (pool) snprintf(buf, sizeof(buf), "render_pool<%d, %d, %d, %d, %d, %d>", w, fma, pool_threads, stream ? kPoolMinBlocks : 1, stream ? 1 : 0, bundle ? 1 : 0);
    else if (w == 1) snprintf(buf, sizeof(buf), "render_paths<1, 0, 0, 4, 4>");
    else if (packed) snprintf(buf, sizeof(buf), "render_paths<%d, %d, 1, %d, %d>", w, fma, w == 4 ? 2 : 4, w == 4 ? 4 : 3);
    else snprintf(buf, sizeof(buf), "render_paths<%d, %d, 0, 4, 4>", w, fma);
    return buf;
}

// Attribute + occupancy calls of a (kernel, threads, shared memory) triple are made once per handle.
// Function attributes belong to the (device, kernel) pair, not to a handle: several handles of one device share them, and
// one handle renders scenes of different sizes with the same kernel.  The dynamic shared memory limit is therefore only
// ever RAISED (a launch may use less than the limit, never more), the carve-out follows the launch that is about to run,
// and all of it is cached process-wide per device so that the per-frame path makes no CUDA attribute call at all.
struct KernelCfg {
    int device; render_kernel_t kernel;
    size_t smem_limit = 0;               // cudaFuncAttributeMaxDynamicSharedMemorySize currently set
    int carve = -2;                      // cudaFuncAttributePreferredSharedMemoryCarveout currently set (-2: never)
    std::vector<LaunchCfg> occ;          // occupancy per (threads, smem)
};
static std::mutex g_cfg_mu;
static std::vector<KernelCfg> g_cfgs;

static int launch_cfg(srt_backend *b, render_kernel_t kernel, int threads, size_t smem, int carve, int *per_sm) {
    std::lock_guard<std::mutex> lock(g_cfg_mu);
    KernelCfg *k = nullptr;
    for (KernelCfg &c : g_cfgs) if (c.device == b->device && c.kernel == kernel) { k = &c; break; }
    if (!k) { g_cfgs.push_back(KernelCfg{b->device, kernel}); k = &g_cfgs.back(); }
    if (smem > k->smem_limit) {
        SRT_CUDA(b, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k->smem_limit = smem;
    }
    if (carve >= 0 && carve != k->carve) {
        SRT_CUDA(b, cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, carve));
        k->carve = carve;
    }
    for (const LaunchCfg &c : k->occ)
        if (c.threads == threads && c.smem == smem) { *per_sm = c.per_sm; return SRT_OK; }
    int n = 0;
    SRT_CUDA(b, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, threads, smem));
    if (n < 1) return fail(b, SRT_ERR_TOO_LARGE, "srt_render: kernel does not fit on an SM");
    k->occ.push_back({kernel, threads, smem, n});
    *per_sm = n;
    return SRT_OK;
}

// Bundle mode set-up of one launch (srt_bundle.cuh): staging bands, item decoding constants, fold weights.
// Returns false (with P untouched where it matters) when this launch has to run without bundle mode.
static int plan_bundle(srt_backend *b, const srt_camera *cam, const srt_render_args *args, KernelParams &P, bool *out_bundle) {
    *out_bundle = false;
    const Knobs &K = b->knobs;
    // auto: scenes of >= kBundleMinSpheres (below that the per-ray scan is too short for the saving to pay for the
    // staging and the larger code) and launches of >= kBundleMinSamples samples (8 spp +8 %, 4 spp -2 %);
    // SRT_BUNDLE=0 off, =k >= 2 forces it
    const bool want = K.bundle > 1 || (K.bundle == 1 && b->n_pad >= kBundleMinSpheres);
    if (!want || args->n_samples < (K.bundle > 1 ? 4u : kBundleMinSamples) || K.pool_chunk || P.items_per_gen == 0u) return SRT_OK;
    uint32_t blog = 0;
    while (blog < 6u && (2u << blog) <= args->n_samples) ++blog;                 // S = min(64, 2^floor(log2 n))
    if (K.bundle > 1) blog = std::min<uint32_t>(blog, (uint32_t)K.bundle);       // SRT_BUNDLE=k (k >= 2): S <= 2^k (tests)
    const uint64_t S = 1ull << blog, n_chunks = (args->n_samples + S - 1) / S;
    const double rho = bundle_footprint(cam, b->width, b->height, P.slab_axis);
    // needs a sane pinhole camera (the footprint bound) and 32-bit item indices
    if (!(rho > 0.0) || n_chunks > 4095 || (uint64_t)P.items_per_gen * n_chunks * S >= 0xF0000000ull) return SRT_OK;
    // bands of whole block rows, as large as kBundleStageBytes of staging rows allow (SRT_BUNDLE_BAND_ROWS: tests,
    // tiny bands).  Measured: larger bands are faster (4K / 256 spp: 2 block rows 6,353, 64 rows 6,525 Mrays/s;
    // 65,536 spheres: 17 rows 155, whole image 190) -- a chunk's fold waits in ring F for 63 others, and the next
    // chunk / band of the same staging row must not start before it; L2 residency of the rows does not matter.
    const uint32_t row_items = P.blocks_x * 32u, blocks_y = P.items_per_gen / row_items;   // local block rows of this handle
    uint32_t band_rows = std::max<uint32_t>(1u, (uint32_t)((kBundleStageBytes / (S * sizeof(float4))) / row_items));
    if (K.band_rows) band_rows = K.band_rows;
    band_rows = std::min(band_rows, blocks_y);
    size_t rows = 0;
    for (;;) {                                                // smaller bands if the device is short of memory
        rows = (size_t)band_rows * row_items;
        const size_t need = rows * S * sizeof(float4);
        if (need <= b->stage_bytes && rows <= b->cnt_rows) break;
        cudaFree(b->d_stage); cudaFree(b->d_cnt);
        b->d_stage = nullptr; b->d_cnt = nullptr; b->stage_bytes = 0; b->cnt_rows = 0;
        if (need <= K.stage_limit && cudaMalloc(&b->d_stage, need) == cudaSuccess && cudaMalloc(&b->d_cnt, rows * sizeof(uint32_t)) == cudaSuccess) {
            b->stage_bytes = need;
            b->cnt_rows = rows;
            break;
        }
        cudaGetLastError();
        cudaFree(b->d_stage); cudaFree(b->d_cnt);
        b->d_stage = nullptr; b->d_cnt = nullptr;
        if (band_rows == 1u) return SRT_OK;                   // not even one block row: this launch runs without bundle mode
        band_rows = (band_rows + 1u) / 2u;
    }
    P.bundle_log2 = blog;
    P.n_chunks = (uint32_t)n_chunks;
    P.n_bands = (blocks_y + band_rows - 1) / band_rows;
    P.band_items = band_rows * row_items;
    P.band_last = P.items_per_gen - (P.n_bands - 1) * P.band_items;
    P.band_pix = band_rows * 4u * b->width * b->n_parts;      // local block rows are n_parts block rows apart in the image
    P.div_full = make_magic(P.band_items * P.n_chunks);
    P.div_band = make_magic(P.band_items);
    P.div_last = make_magic(P.band_last);
    P.bundle_rho = (float)rho;
    P.total_items = (uint32_t)((uint64_t)P.items_per_gen * n_chunks * S);
    SRT_CUDA(b, cudaMemsetAsync(b->d_cnt, 0, rows * sizeof(uint32_t), b->stream));
    P.stage = b->d_stage;
    P.cnt = b->d_cnt;
    if (args->accum_mode == SRT_ACCUM_MEAN) {                 // A8 weights of the launch's samples (main.cpp:484-487)
        if (args->n_samples > b->fold_w_cap) {
            cudaFree(b->d_fold_w);
            b->d_fold_w = nullptr; b->fold_w_cap = 0; b->fold_w_n = 0;
            SRT_CUDA(b, cudaMalloc(&b->d_fold_w, (size_t)args->n_samples * sizeof(float2)));
            b->fold_w_cap = args->n_samples;
        }
        if (b->fold_w_n != args->n_samples || b->fold_w_first != args->first_sample || b->fold_w_stride != args->sample_stride) {
            b->h_fold_w.resize(args->n_samples);
            for (uint32_t i = 0; i < args->n_samples; ++i) {
                const uint32_t k = args->first_sample + i * args->sample_stride;
                const float total = (float)(k + 1u);
                b->h_fold_w[i] = make_float2(1.0f / total, (float)k / total);
            }
            // pageable source: the copy has left h_fold_w when the call returns
            SRT_CUDA(b, cudaMemcpyAsync(b->d_fold_w, b->h_fold_w.data(), (size_t)args->n_samples * sizeof(float2), cudaMemcpyHostToDevice, b->stream));
            b->fold_w_n = args->n_samples; b->fold_w_first = args->first_sample; b->fold_w_stride = args->sample_stride;
        }
        P.fold_w = b->d_fold_w;
    }
    *out_bundle = true;
    return SRT_OK;
}

int srt_render(srt_backend *b, const srt_camera *cam, const srt_render_args *args) {
    if (!b) return SRT_ERR_INVALID;
    KernelParams P;
    int rc = make_params(b, cam, args, "srt_render", P);
    if (rc) return rc;
    SRT_CUDA(b, cudaSetDevice(b->device));
    rc = drain(b);
    if (rc) return rc;
    b->gathered_rgba = b->gathered_accum = false;
    if (args->record_primary) {
        const size_t n = (size_t)b->width * b->height;
        if (!b->d_tap_prim) {
            SRT_CUDA(b, cudaMalloc(&b->d_tap_prim, n * sizeof(uint32_t)));
            SRT_CUDA(b, cudaMalloc(&b->d_tap_t, n * sizeof(float)));
        }
        if (b->n_parts > 1u) {            // image partition: the rows of the other devices read as "not rendered here" (all bits set)
            SRT_CUDA(b, cudaMemsetAsync(b->d_tap_prim, 0xFF, n * sizeof(uint32_t), b->stream));
            SRT_CUDA(b, cudaMemsetAsync(b->d_tap_t, 0xFF, n * sizeof(float), b->stream));
        }
        P.tap_prim = b->d_tap_prim;
        P.tap_t = b->d_tap_t;
        b->tap_valid = true;
    }
    if (P.total_items == 0u) {            // more devices than block rows: this handle owns no pixel
        *b->h_rays = 0;
        b->events_recorded = false;
        return SRT_OK;
    }

    // kernel choice: tiny scenes -> one path per thread; otherwise the pool kernel, with the sphere tables resident
    // in shared memory (the configuration with the most warps whose tables + path slots fit) or streamed through
    // per-warp TMA stages when they do not fit
    const Knobs &K = b->knobs;
    int pool_threads = 0;
    for (int t : kPoolResidentThreads) {
        if (K.pool_threads && K.pool_threads != t) continue;
        if (b->n_pad <= (uint32_t)kPairsMaxSpheres && pool_smem_bytes(b->n_pad, t, false) <= b->max_smem_optin) { pool_threads = t; break; }
    }
    const bool pool = K.variant == 9 || K.variant == 12 || (K.variant == 0 && b->n_pad >= kPoolMinSpheres);
    const bool stream = pool && (K.variant == 12 || pool_threads == 0);
    // <= 16 spheres (the demo scene): shading-bound, the scalar-issue loop at 4 CTAs per SM is the fastest valid configuration
    const bool packed = !pool && K.variant != 1 && b->opt.lane_width != 1 && b->n_pad > 16u;
    if (!pool && (size_t)b->n_pad * sizeof(float4) > b->max_smem_optin)
        return fail(b, SRT_ERR_TOO_LARGE, "srt_render: scene too large for the one-path kernel");
    if (packed) P.spheres = b->d_spheres2;

    bool use_bundle = false;
    if (!K.pool_debug) P.debug = nullptr;
    if (pool) {
        // work items = (sample chunk, pixel), chunk-major.  A pixel's chunks are folded in order through
        // done[pix]; short chunks keep the end of the launch short (few slots left with many samples to go).
        const size_t done_bytes = (size_t)b->width * b->height * sizeof(uint32_t);
        bool done_fresh = false;
        if (!b->d_done) {   // 0xFF..: tag 0xFFFFF is never issued (recycled device memory may hold another backend's counts)
            SRT_CUDA(b, cudaMalloc(&b->d_done, done_bytes));
            done_fresh = true;
        }
        rc = plan_bundle(b, cam, args, P, &use_bundle);
        if (rc) return rc;
        if (!use_bundle) {
            uint32_t chunk = K.pool_chunk ? K.pool_chunk : kPoolChunkSamples;
            const uint32_t max_chunks = std::min<uint32_t>(4095u, 0xF0000000u / P.items_per_gen);
            if ((args->n_samples + chunk - 1) / chunk > max_chunks) chunk = (args->n_samples + max_chunks - 1) / max_chunks;
            const uint32_t n_chunks = (args->n_samples + chunk - 1) / chunk;
            P.chunk_span = chunk * args->sample_stride;
            P.total_items = P.items_per_gen * n_chunks;
            P.n_chunks = n_chunks;
        }
        // done[pix] = tag << 12 | chunks folded (<= 4095).  The tag is this handle's pool-launch sequence number modulo
        // 0xFFFFF (a dedicated counter: other launches do not advance it); when it wraps, every count of the previous
        // cycle is forgotten, whatever this launch's chunk count.
        const uint32_t tag = b->pool_seq % 0xFFFFFu;
        b->pool_seq = tag + 1u;
        if (done_fresh || tag == 0u) SRT_CUDA(b, cudaMemsetAsync(b->d_done, 0xFF, done_bytes, b->stream));
        P.done = b->d_done;
        P.done_tag = tag << 12;
    }
    if (stream) pool_threads = kPoolThreads;
    render_kernel_t kernel = pool ? pick_pool_kernel(b->opt, stream, pool_threads, use_bundle) : pick_kernel(b->opt, packed);
    b->last_kernel = kernel_label(b->opt, pool, stream, pool_threads, use_bundle, packed);
    if (!kernel) return fail(b, SRT_ERR_INVALID, "srt_render: this build (SRT_FAST_BUILD) only carries the lane_width 8 + FMA kernels");
    const int threads = pool ? pool_threads : kBlockThreads;
    const size_t smem = pool ? pool_smem_bytes(b->n_pad, pool_threads, stream, use_bundle) : (size_t)b->n_pad * sizeof(float4);
    int carve = -1;
    if (pool) {   // leave the rest of the 256 KB to L1: the candidate queues live in local memory
        const size_t ctas = stream ? kPoolMinBlocks : 1;
        carve = (int)std::min<size_t>(100, (ctas * (smem + 1024) * 100 + 228 * 1024 - 1) / (228 * 1024));
    }
    int per_sm = 0;
    rc = launch_cfg(b, kernel, threads, smem, carve, &per_sm);
    if (rc) return rc;
    uint32_t grid = (uint32_t)(b->sm_count * per_sm);                 // persistent: one resident wave
    const uint32_t paths_per_block = pool ? 64u * (threads / 32 + kPoolSpareBatches) : (uint32_t)threads;
    const uint32_t max_useful = (P.total_items + paths_per_block - 1) / paths_per_block;
    if (grid > max_useful) grid = max_useful;

    if (b->counters_dirty || K.pool_debug) {   // normally the previous launch's last CTA has re-armed the counters
        SRT_CUDA(b, cudaMemsetAsync(b->d_counters, 0, 64, b->stream));
        b->counters_dirty = false;
    }
    SRT_CUDA(b, cudaEventRecord(b->ev_start, b->stream));
    if (stream) {
        P.ptab = b->d_ptab;
        build_slab_kernel<<<(b->n_pad + 255) / 256, 256, 0, b->stream>>>(b->d_scan, b->d_spheres, (float *)b->d_ptab, b->n_pad, b->n_spheres,
                                                                        P.slab_axis, P.omax);
        b->launches += 1;
    }
    kernel<<<grid, threads, smem, b->stream>>>(P);
    const cudaError_t le = cudaGetLastError();
    if (le != cudaSuccess) { b->counters_dirty = true; return fail(b, SRT_ERR_CUDA, std::string("srt_render: launch: ") + cudaGetErrorString(le)); }
    SRT_CUDA(b, cudaEventRecord(b->ev_stop, b->stream));
    b->events_recorded = true;
    b->launches += 1;
    b->in_flight = true;
    if (K.pool_debug) {
        unsigned long long dbg[6] = {0, 0, 0, 0, 0, 0};
        cudaStreamSynchronize(b->stream);
        cudaMemcpy(dbg, b->d_counters + 16, sizeof(dbg), cudaMemcpyDeviceToHost);
        fprintf(stderr, "[srt pool] acquire spins: steady %llu, tail %llu; primary passes: bundled %llu, per-ray %llu, pixel runs %llu, lane-0 waits %llu\n",
                dbg[0], dbg[1], dbg[2], dbg[3], dbg[4], dbg[5]);
    }
    return SRT_OK;
}

// srt_render without any allocation: the pool kernel's chunk counters live in the file-scope array g_static_done.
// Same kernels, same results as srt_render (bundle mode needs staging rows that only srt_render allocates, so launches
// that srt_render would run in bundle mode run the non-bundle pool kernel here -- bit-identical by construction, slower).
// The array is shared by every handle and every call on the device and keeps the last call's contents: the part this
// launch reads is reset on the stream first; calls that overlap on the GPU (two handles of one device) race.
int srt_render_static(srt_backend *b, const srt_camera *cam, const srt_render_args *args) {
    if (!b) return SRT_ERR_INVALID;
    KernelParams P;
    int rc = make_params(b, cam, args, "srt_render_static", P);
    if (rc) return rc;
    if (args->record_primary) return fail(b, SRT_ERR_INVALID, "srt_render_static: record_primary needs buffers that only srt_render allocates");
    const uint64_t n_pixels = (uint64_t)b->width * b->height;
    if (n_pixels > kStaticDonePixels)
        return fail(b, SRT_ERR_TOO_LARGE, "srt_render_static: image of " + std::to_string(n_pixels) + " pixels exceeds the static scratch of " +
                                          std::to_string(kStaticDonePixels) + " pixels (SRT_STATIC_MAX_PIXELS); use srt_render");
    if (!b->d_static_done) return fail(b, SRT_ERR_INVALID, "srt_render_static: static scratch was not resolved for this device");
    SRT_CUDA(b, cudaSetDevice(b->device));
    rc = drain(b);
    if (rc) return rc;
    b->gathered_rgba = b->gathered_accum = false;
    if (P.total_items == 0u) {
        *b->h_rays = 0;
        b->events_recorded = false;
        return SRT_OK;
    }
    const Knobs &K = b->knobs;
    int pool_threads = 0;
    for (int t : kPoolResidentThreads) {
        if (K.pool_threads && K.pool_threads != t) continue;
        if (b->n_pad <= (uint32_t)kPairsMaxSpheres && pool_smem_bytes(b->n_pad, t, false) <= b->max_smem_optin) { pool_threads = t; break; }
    }
    const bool pool = K.variant == 9 || K.variant == 12 || (K.variant == 0 && b->n_pad >= kPoolMinSpheres);
    const bool stream = pool && (K.variant == 12 || pool_threads == 0);
    const bool packed = !pool && K.variant != 1 && b->opt.lane_width != 1 && b->n_pad > 16u;
    if (!pool && (size_t)b->n_pad * sizeof(float4) > b->max_smem_optin)
        return fail(b, SRT_ERR_TOO_LARGE, "srt_render_static: scene too large for the one-path kernel");
    if (packed) P.spheres = b->d_spheres2;
    P.debug = nullptr;
    if (pool) {
        uint32_t chunk = K.pool_chunk ? K.pool_chunk : kPoolChunkSamples;
        const uint32_t max_chunks = std::min<uint32_t>(4095u, 0xF0000000u / P.items_per_gen);
        if ((args->n_samples + chunk - 1) / chunk > max_chunks) chunk = (args->n_samples + max_chunks - 1) / max_chunks;
        const uint32_t n_chunks = (args->n_samples + chunk - 1) / chunk;
        P.chunk_span = chunk * args->sample_stride;
        P.total_items = P.items_per_gen * n_chunks;
        P.n_chunks = n_chunks;
        // The counters of this image's pixels must not match this launch's tag: other handles and earlier calls wrote the
        // array with tags of their own, so it is reset (0xFF: the reserved tag) before every launch that reads it.
        if (n_chunks > 1u) SRT_CUDA(b, cudaMemsetAsync(b->d_static_done, 0xFF, (size_t)n_pixels * sizeof(uint32_t), b->stream));
        const uint32_t tag = b->pool_seq % 0xFFFFFu;
        b->pool_seq = tag + 1u;
        P.done = b->d_static_done;
        P.done_tag = tag << 12;
    }
    if (stream) pool_threads = kPoolThreads;
    render_kernel_t kernel = pool ? pick_pool_kernel(b->opt, stream, pool_threads, false) : pick_kernel(b->opt, packed);
    b->last_kernel = kernel_label(b->opt, pool, stream, pool_threads, false, packed);
    if (!kernel) return fail(b, SRT_ERR_INVALID, "srt_render_static: this build (SRT_FAST_BUILD) only carries the lane_width 8 + FMA kernels");
    const int threads = pool ? pool_threads : kBlockThreads;
    const size_t smem = pool ? pool_smem_bytes(b->n_pad, pool_threads, stream, false) : (size_t)b->n_pad * sizeof(float4);
    int carve = -1;
    if (pool) {
        const size_t ctas = stream ? kPoolMinBlocks : 1;
        carve = (int)std::min<size_t>(100, (ctas * (smem + 1024) * 100 + 228 * 1024 - 1) / (228 * 1024));
    }
    int per_sm = 0;
    rc = launch_cfg(b, kernel, threads, smem, carve, &per_sm);
    if (rc) return rc;
    uint32_t grid = (uint32_t)(b->sm_count * per_sm);
    const uint32_t paths_per_block = pool ? 64u * (threads / 32 + kPoolSpareBatches) : (uint32_t)threads;
    const uint32_t max_useful = (P.total_items + paths_per_block - 1) / paths_per_block;
    if (grid > max_useful) grid = max_useful;
    if (b->counters_dirty) {
        SRT_CUDA(b, cudaMemsetAsync(b->d_counters, 0, 64, b->stream));
        b->counters_dirty = false;
    }
    SRT_CUDA(b, cudaEventRecord(b->ev_start, b->stream));
    if (stream) {
        P.ptab = b->d_ptab;
        build_slab_kernel<<<(b->n_pad + 255) / 256, 256, 0, b->stream>>>(b->d_scan, b->d_spheres, (float *)b->d_ptab, b->n_pad, b->n_spheres,
                                                                        P.slab_axis, P.omax);
        b->launches += 1;
    }
    kernel<<<grid, threads, smem, b->stream>>>(P);
    const cudaError_t le = cudaGetLastError();
    if (le != cudaSuccess) { b->counters_dirty = true; return fail(b, SRT_ERR_CUDA, std::string("srt_render_static: launch: ") + cudaGetErrorString(le)); }
    SRT_CUDA(b, cudaEventRecord(b->ev_stop, b->stream));
    b->events_recorded = true;
    b->launches += 1;
    b->in_flight = true;
    return SRT_OK;
}

int srt_poll(srt_backend *b) {
    if (!b) return SRT_ERR_INVALID;
    if (!b->in_flight) return 1;
    cudaError_t e = cudaStreamQuery(b->stream);
    if (e == cudaSuccess) return 1;
    if (e == cudaErrorNotReady) return 0;
    return fail(b, SRT_ERR_CUDA, std::string("srt_poll: ") + cudaGetErrorString(e));
}

int srt_wait(srt_backend *b, double *out_ms, uint64_t *out_rays) {
    if (!b) return SRT_ERR_INVALID;
    SRT_CUDA(b, cudaSetDevice(b->device));
    const cudaError_t e = cudaStreamSynchronize(b->stream);
    b->in_flight = false;
    if (e != cudaSuccess) {
        b->counters_dirty = true;
        std::string msg = std::string("srt_wait: ") + cudaGetErrorString(e);
        unsigned long long rep = *(volatile unsigned long long *)(b->h_rays + 1);
#ifdef SRT_SPIN_LIMIT
        unsigned long long rep_value = *(volatile unsigned long long *)(b->h_rays + 2);
        if ((rep >> 48) != 0xDEADull && g_spin_report_host) { rep = *(volatile unsigned long long *)g_spin_report_host; rep_value = *(volatile unsigned long long *)(g_spin_report_host + 1); }
#endif
        if ((rep >> 48) == 0xDEADull) {      // -DSRT_SPIN_LIMIT build: a wait loop gave up (site: 1 ring pop, 2 ring push, 3 acquire, 4 TMA barrier) or an index check failed (site >= 16)
            char buf[160];
#ifdef SRT_SPIN_LIMIT
            snprintf(buf, sizeof(buf), " [value 0x%llx]", rep_value);
            msg += buf;
#endif
            snprintf(buf, sizeof(buf), "; spin limit exceeded or index check failed at site %u (CTA %u, thread %u)", (unsigned)((rep >> 32) & 0xFFFFu), (unsigned)((rep >> 12) & 0xFFFFFu), (unsigned)(rep & 0xFFFu));
            msg += buf;
        }
        return fail(b, SRT_ERR_CUDA, msg);
    }
    if (out_ms) {
        float ms = 0.f;
        if (b->events_recorded) SRT_CUDA(b, cudaEventElapsedTime(&ms, b->ev_start, b->ev_stop));
        *out_ms = ms;
    }
    if (out_rays) *out_rays = *(volatile unsigned long long *)b->h_rays;   // written by the last CTA of the last render
    return SRT_OK;
}

int srt_filter_audit(srt_backend *b, const srt_camera *cam, const srt_render_args *args, uint64_t out_counts[8]) {
    if (!b) return SRT_ERR_INVALID;
    if (!out_counts) return fail(b, SRT_ERR_INVALID, "srt_filter_audit: null output");
    KernelParams P;
    int rc = make_params(b, cam, args, "srt_filter_audit", P);
    if (rc) return rc;
    if (args->sample_stride != 1) return fail(b, SRT_ERR_INVALID, "srt_filter_audit: sample_stride must be 1");
    P.bundle_rho = (float)bundle_footprint(cam, b->width, b->height, P.slab_axis);
    SRT_CUDA(b, cudaSetDevice(b->device));
    rc = drain(b);
    if (rc) return rc;
    unsigned long long *d_cnt = nullptr;
    SRT_CUDA(b, cudaMalloc(&d_cnt, 8 * sizeof(unsigned long long)));
    SRT_CUDA(b, cudaMemsetAsync(d_cnt, 0, 8 * sizeof(unsigned long long), b->stream));
    const uint32_t n = b->width * b->height;
    const dim3 grid((n + kBlockThreads - 1) / kBlockThreads);
    const srt_scene_options &o = b->opt;
#ifdef SRT_FAST_BUILD
    if (o.lane_width != 8 || !o.fma_dot) { cudaFree(d_cnt); return fail(b, SRT_ERR_INVALID, "srt_filter_audit: SRT_FAST_BUILD only carries lane_width 8 + FMA"); }
    audit_filter<8, true><<<grid, kBlockThreads, 0, b->stream>>>(P, d_cnt);
#else
    if (o.lane_width == 1) audit_filter<1, false><<<grid, kBlockThreads, 0, b->stream>>>(P, d_cnt);
    else if (o.lane_width == 4 && o.fma_dot) audit_filter<4, true><<<grid, kBlockThreads, 0, b->stream>>>(P, d_cnt);
    else if (o.lane_width == 4) audit_filter<4, false><<<grid, kBlockThreads, 0, b->stream>>>(P, d_cnt);
    else if (o.fma_dot) audit_filter<8, true><<<grid, kBlockThreads, 0, b->stream>>>(P, d_cnt);
    else audit_filter<8, false><<<grid, kBlockThreads, 0, b->stream>>>(P, d_cnt);
#endif
    cudaError_t e = cudaGetLastError();
    unsigned long long h[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (e == cudaSuccess) e = cudaMemcpyAsync(h, d_cnt, sizeof(h), cudaMemcpyDeviceToHost, b->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(b->stream);
    cudaFree(d_cnt);
    if (e != cudaSuccess) return fail(b, SRT_ERR_CUDA, std::string("srt_filter_audit: ") + cudaGetErrorString(e));
    for (int i = 0; i < 8; ++i) out_counts[i] = h[i];
    b->launches += 1;
    return SRT_OK;
}

int srt_device_kat(srt_backend *b, uint32_t op, const void *in, uint32_t n, void *out) {
    if (!b) return SRT_ERR_INVALID;
    static const uint32_t in_bytes[SRT_KAT_COUNT] = {8, 16, 8, 12, 8, 4, 16}, out_bytes[SRT_KAT_COUNT] = {64, 64, 8, 12, 4, 4, 4};
    if (op >= SRT_KAT_COUNT || !in || !out || n == 0) return fail(b, SRT_ERR_INVALID, "srt_device_kat: bad arguments");
    SRT_CUDA(b, cudaSetDevice(b->device));
    void *d_in = nullptr, *d_out = nullptr;
    cudaError_t e = cudaMalloc(&d_in, (size_t)n * in_bytes[op]);
    if (e == cudaSuccess) e = cudaMalloc(&d_out, (size_t)n * out_bytes[op]);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_in, in, (size_t)n * in_bytes[op], cudaMemcpyHostToDevice, b->stream);
    if (e == cudaSuccess) {
        kat_kernel<<<(n + 127) / 128, 128, 0, b->stream>>>(op, d_in, n, d_out);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_out, (size_t)n * out_bytes[op], cudaMemcpyDeviceToHost, b->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(b->stream);
    cudaFree(d_in); cudaFree(d_out);
    if (e != cudaSuccess) return fail(b, SRT_ERR_CUDA, std::string("srt_device_kat: ") + cudaGetErrorString(e));
    b->launches += 1;
    return SRT_OK;
}

int srt_resolve_sum(srt_backend *b, uint32_t total_samples) {
    if (!b) return SRT_ERR_INVALID;
    if (!b->d_accum || total_samples == 0) return fail(b, SRT_ERR_INVALID, "srt_resolve_sum: no image or zero samples");
    SRT_CUDA(b, cudaSetDevice(b->device));
    const uint32_t n = b->width * b->height;
    const volatile float one = 1.0f;
    const float inv_n = one / (float)total_samples;
    // the 8-byte RGBA8 store of the two-pixel kernel needs an 8-byte aligned image (cudaMalloc / torch buffers are)
    if (b->knobs.resolve_variant == 1 || (reinterpret_cast<uintptr_t>(b->d_rgba) & 7u) != 0)
        resolve_sum_kernel_1px<<<(n + 255) / 256, 256, 0, b->stream>>>(b->d_accum, b->d_rgba, n, inv_n);
    else
        resolve_sum_kernel<<<((n + 1) / 2 + 255) / 256, 256, 0, b->stream>>>(b->d_accum, b->d_rgba, n, inv_n);
    SRT_CUDA(b, cudaGetLastError());
    b->launches += 1;
    b->in_flight = true;
    return SRT_OK;
}

int srt_read_rgba8(srt_backend *b, void *dst) {
    if (!b || !dst) return SRT_ERR_INVALID;
    if (!b->d_rgba) return fail(b, SRT_ERR_INVALID, "srt_read_rgba8: no image");
    SRT_CUDA(b, cudaSetDevice(b->device));
    const uint32_t *src = b->gathered_rgba ? b->d_gather_rgba : b->d_rgba;   // after srt_gather_rgba: the combined image
    SRT_CUDA(b, cudaMemcpyAsync(dst, src, (size_t)b->width * b->height * 4, cudaMemcpyDeviceToHost, b->stream));
    SRT_CUDA(b, cudaStreamSynchronize(b->stream));
    b->in_flight = false;
    return SRT_OK;
}

int srt_read_accum(srt_backend *b, float *dst) {
    if (!b || !dst) return SRT_ERR_INVALID;
    if (!b->d_accum) return fail(b, SRT_ERR_INVALID, "srt_read_accum: no image");
    SRT_CUDA(b, cudaSetDevice(b->device));
    const float4 *src = b->gathered_accum ? b->d_gather_accum : b->d_accum;
    SRT_CUDA(b, cudaMemcpyAsync(dst, src, (size_t)b->width * b->height * 16, cudaMemcpyDeviceToHost, b->stream));
    SRT_CUDA(b, cudaStreamSynchronize(b->stream));
    b->in_flight = false;
    return SRT_OK;
}

int srt_write_accum(srt_backend *b, const float *src) {
    if (!b || !src) return SRT_ERR_INVALID;
    if (!b->d_accum) return fail(b, SRT_ERR_INVALID, "srt_write_accum: no image (call srt_resize)");
    SRT_CUDA(b, cudaSetDevice(b->device));
    SRT_CUDA(b, cudaMemcpyAsync(b->d_accum, src, (size_t)b->width * b->height * 16, cudaMemcpyHostToDevice, b->stream));
    if (b->n_parts > 1u) {   // image partition: rows of other devices must stay zero (the combine step is a sum)
        const uint32_t n = b->width * b->height;
        clear_foreign_rows<<<(n + 255) / 256, 256, 0, b->stream>>>(b->d_accum, b->width, n, b->part, b->n_parts);
        SRT_CUDA(b, cudaGetLastError());
    }
    SRT_CUDA(b, cudaStreamSynchronize(b->stream));
    b->in_flight = false;
    return SRT_OK;
}

int srt_read_primary(srt_backend *b, uint32_t *prim, float *t) {
    if (!b) return SRT_ERR_INVALID;
    if (!b->tap_valid) return fail(b, SRT_ERR_INVALID, "srt_read_primary: no render with record_primary since the last reset");
    SRT_CUDA(b, cudaSetDevice(b->device));
    const size_t n = (size_t)b->width * b->height;
    if (prim) SRT_CUDA(b, cudaMemcpyAsync(prim, b->d_tap_prim, n * 4, cudaMemcpyDeviceToHost, b->stream));
    if (t) SRT_CUDA(b, cudaMemcpyAsync(t, b->d_tap_t, n * 4, cudaMemcpyDeviceToHost, b->stream));
    SRT_CUDA(b, cudaStreamSynchronize(b->stream));
    b->in_flight = false;
    return SRT_OK;
}

void *srt_accum_device_ptr(srt_backend *b) { return b ? (void *)b->d_accum : nullptr; }
void *srt_rgba_device_ptr(srt_backend *b) { return b ? (void *)b->d_rgba : nullptr; }
void *srt_cuda_stream(srt_backend *b) { return b ? (void *)b->stream : nullptr; }

// ---- multi-device: one handle per device, one host thread or process per handle ----------------------------
// The reference fans a frame out to ThreadCount workers (WorkQueueStart, base.h:174, main.cpp:851-856) that pull
// 32x32 tiles (main.cpp:9, 362-368); here the workers are GPUs and the one exchange step is an NCCL reduce on the
// handle's stream.
int srt_comm_unique_id(void *id_out) {
    if (!id_out) return fail(nullptr, SRT_ERR_INVALID, "srt_comm_unique_id: null output");
    NcclApi &N = nccl_api();
    if (!N.GetUniqueId) return fail(nullptr, SRT_ERR_COMM, "srt_comm_unique_id: " + N.error);
    static_assert(sizeof(ncclUniqueId) == SRT_COMM_ID_BYTES, "ncclUniqueId size");
    ncclUniqueId id;
    const ncclResult_t r = N.GetUniqueId(&id);
    if (r != ncclSuccess) return fail(nullptr, SRT_ERR_COMM, std::string("ncclGetUniqueId: ") + N.GetErrorString(r));
    memcpy(id_out, &id, sizeof(id));
    return SRT_OK;
}

// One host thread that drives several handles brackets each round of collective calls (and the init calls): NCCL
// then treats them as one operation instead of blocking in the first.
int srt_comm_group_begin(void) {
    NcclApi &N = nccl_api();
    if (!N.GroupStart) return fail(nullptr, SRT_ERR_COMM, "srt_comm_group_begin: " + N.error);
    const ncclResult_t r = N.GroupStart();
    return r == ncclSuccess ? SRT_OK : fail(nullptr, SRT_ERR_COMM, std::string("ncclGroupStart: ") + N.GetErrorString(r));
}
int srt_comm_group_end(void) {
    NcclApi &N = nccl_api();
    if (!N.GroupEnd) return fail(nullptr, SRT_ERR_COMM, "srt_comm_group_end: " + N.error);
    const ncclResult_t r = N.GroupEnd();
    return r == ncclSuccess ? SRT_OK : fail(nullptr, SRT_ERR_COMM, std::string("ncclGroupEnd: ") + N.GetErrorString(r));
}

int srt_comm_init_rank(srt_backend *b, int n_ranks, int rank, const void *unique_id) {
    if (!b) return SRT_ERR_INVALID;
    if (!unique_id || n_ranks < 1 || rank < 0 || rank >= n_ranks) return fail(b, SRT_ERR_INVALID, "srt_comm_init_rank: bad rank / n_ranks / id");
    NcclApi &N = nccl_api();
    if (!N.CommInitRank || !N.Reduce || !N.AllReduce || !N.CommDestroy) return fail(b, SRT_ERR_COMM, "srt_comm_init_rank: " + (N.error.empty() ? std::string("NCCL unavailable") : N.error));
    SRT_CUDA(b, cudaSetDevice(b->device));
    int rc = drain(b);
    if (rc) return rc;
    comm_release(b);
    ncclUniqueId id;
    memcpy(&id, unique_id, sizeof(id));
    ncclComm_t comm = nullptr;
    SRT_NCCL(b, N.CommInitRank(&comm, n_ranks, id, rank));
    b->comm = comm;
    b->comm_rank = rank;
    b->comm_size = n_ranks;
    return SRT_OK;
}

int srt_comm_destroy(srt_backend *b) {
    if (!b) return SRT_ERR_INVALID;
    SRT_CUDA(b, cudaSetDevice(b->device));
    int rc = drain(b);
    if (rc) return rc;
    comm_release(b);
    return SRT_OK;
}

int srt_comm_info(srt_backend *b, int *rank, int *n_ranks, int *nccl_version) {
    if (!b) return SRT_ERR_INVALID;
    if (rank) *rank = b->comm_rank;
    if (n_ranks) *n_ranks = b->comm ? b->comm_size : 0;
    if (nccl_version) { *nccl_version = 0; NcclApi &N = nccl_api(); if (N.GetVersion) N.GetVersion(nccl_version); }
    return SRT_OK;
}

int srt_set_partition(srt_backend *b, uint32_t part, uint32_t n_parts) {
    if (!b) return SRT_ERR_INVALID;
    if (n_parts == 0 || part >= n_parts || n_parts > 65535u) return fail(b, SRT_ERR_INVALID, "srt_set_partition: need part < n_parts <= 65535");
    int rc = drain(b);
    if (rc) return rc;
    const bool changed = part != b->part || n_parts != b->n_parts;
    b->part = part;
    b->n_parts = n_parts;
    // rows that now belong to another device must read as zero: the combine step is a sum
    return (changed && b->d_accum) ? srt_reset_accumulation(b) : SRT_OK;
}

// Sample sharding: every device holds a SUM over its samples of every pixel (SRT_ACCUM_SUM).
int srt_reduce_accum(srt_backend *b, int root) {
    if (!b) return SRT_ERR_INVALID;
    if (!b->comm) return fail(b, SRT_ERR_INVALID, "srt_reduce_accum: no communicator (call srt_comm_init_rank)");
    if (!b->d_accum) return fail(b, SRT_ERR_INVALID, "srt_reduce_accum: no image");
    if (root >= b->comm_size) return fail(b, SRT_ERR_INVALID, "srt_reduce_accum: bad root");
    SRT_CUDA(b, cudaSetDevice(b->device));
    NcclApi &N = nccl_api();
    const size_t n = (size_t)b->width * b->height * 4;
    if (root < 0) SRT_NCCL(b, N.AllReduce(b->d_accum, b->d_accum, n, ncclFloat32, ncclSum, (ncclComm_t)b->comm, b->stream));
    else SRT_NCCL(b, N.Reduce(b->d_accum, b->d_accum, n, ncclFloat32, ncclSum, root, (ncclComm_t)b->comm, b->stream));
    b->launches += 1;
    b->in_flight = true;
    return SRT_OK;
}

// Image partition: the rows a device does not own are zero in its buffers, so an integer (RGBA8 words) or f32
// (x + 0 = x) sum over the devices IS the gather of the row bands -- bit-identical to a single-device render.
static int gather_image(srt_backend *b, int root, bool accum, const char *who) {
    const std::string w(who);
    if (!b->comm) return fail(b, SRT_ERR_INVALID, w + ": no communicator (call srt_comm_init_rank)");
    if (!b->d_accum) return fail(b, SRT_ERR_INVALID, w + ": no image");
    if (root >= b->comm_size) return fail(b, SRT_ERR_INVALID, w + ": bad root");
    SRT_CUDA(b, cudaSetDevice(b->device));
    NcclApi &N = nccl_api();
    const size_t px = (size_t)b->width * b->height;
    const bool recv = root < 0 || root == b->comm_rank;
    void *dst = nullptr;
    if (recv) {
        if (accum) { if (!b->d_gather_accum) SRT_CUDA(b, cudaMalloc(&b->d_gather_accum, px * sizeof(float4))); dst = b->d_gather_accum; }
        else { if (!b->d_gather_rgba) SRT_CUDA(b, cudaMalloc(&b->d_gather_rgba, px * sizeof(uint32_t))); dst = b->d_gather_rgba; }
    }
    const void *src = accum ? (const void *)b->d_accum : (const void *)b->d_rgba;
    const size_t n = accum ? px * 4 : px;
    const ncclDataType_t ty = accum ? ncclFloat32 : ncclUint32;
    if (root < 0) SRT_NCCL(b, N.AllReduce(src, dst, n, ty, ncclSum, (ncclComm_t)b->comm, b->stream));
    else SRT_NCCL(b, N.Reduce(src, recv ? dst : (void *)src, n, ty, ncclSum, root, (ncclComm_t)b->comm, b->stream));
    if (recv) { if (accum) b->gathered_accum = true; else b->gathered_rgba = true; }
    b->launches += 1;
    b->in_flight = true;
    return SRT_OK;
}
int srt_gather_rgba(srt_backend *b, int root) { return b ? gather_image(b, root, false, "srt_gather_rgba") : SRT_ERR_INVALID; }
int srt_gather_accum(srt_backend *b, int root) { return b ? gather_image(b, root, true, "srt_gather_accum") : SRT_ERR_INVALID; }

int srt_last_kernel(srt_backend *b, char *name, uint32_t capacity) {
    if (!b || !name || capacity == 0) return SRT_ERR_INVALID;
    snprintf(name, capacity, "%s", b->last_kernel.c_str());
    return SRT_OK;
}

int srt_device_info(srt_backend *b, int *sm_count, int *clock_khz, uint64_t *launches) {
    if (!b) return SRT_ERR_INVALID;
    if (sm_count) *sm_count = b->sm_count;
    if (clock_khz) *clock_khz = b->clock_khz;
    if (launches) *launches = b->launches;
    return SRT_OK;
}

} // extern "C"
