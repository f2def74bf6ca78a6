// srt_pool.cuh -- CTA-level path pools of the pool kernel (sm_100a).
//
// A thread that owns its two path slots mixes fresh primary rays with secondary rays in every
// pass, and regeneration runs for a fraction of the lanes each time.  Here the path slots belong
// to the CTA, not to a thread: rings in shared memory hold the ids of the slots that need a
// NEW PRIMARY ray (ring P), of those with a SECONDARY ray ready (ring S) and -- bundle mode --
// of those that delivered the last sample of a pixel chunk and must fold it (ring F).  A warp
// repeatedly takes 64 ids of one kind, runs regenerate / scan / select / shade on them (two per
// lane, packed in the f32x2 halves) and pushes every slot back to the ring its next step needs
// (warp-aggregated, ballot + popc).  Work items are (sample chunk, pixel) pairs handed out by a
// global counter; a pixel's chunks fold in order through done[pix] (srt_backend.cu), so the
// running mean of main.cpp:484-489 keeps its order and every buffer stays bit-identical.
//
// Ring protocol (multi-producer / multi-consumer, no CTA barrier after start-up):
//   push: pos = atomicAdd(tail, n); wait until ring[pos+i] is EMPTY; write id; fence;
//         atomicAdd(avail, n)
//   take: CAS avail -= n (only if avail >= n); pos = atomicAdd(head, n); wait until
//         ring[pos+i] != EMPTY; read id; write EMPTY
// avail counts published ids, so [head, head+n) has always been reserved by producers that
// are a few instructions away from writing it: both waits are short and cannot deadlock.
// "Short" is load-bearing: an entry that stays unread for a whole lap of its ring lets the producers of two laps wait on
// it together, and both write when it is cleared.  So every lane of a warp must be past its spin before the warp moves on:
// pool_push ends in __syncwarp(), and the caller of pool_pop reconverges right after the pops (render_pool) -- without
// that, the lanes that already hold their ids may run a whole scan while the others have not even read their entries.
// The -DSRT_SPIN_LIMIT build audits all of this per entry (g_dbg_* below).
// With kPoolSlots = 64 * (warps + 2) one of the rings always holds a full batch while the
// other warps are busy.
#pragma once
#include "srt_pairs.cuh"
#include "srt_path.cuh"

namespace srt {

constexpr uint32_t kRingEmpty = 0xFFFFFFFFu;   // `no slot` in registers
// Ring entries are 16-bit slot ids (< 2,048 slots per CTA): with three rings of 2,048 entries the bundle-mode CTA
// stays under 196 KB of shared memory, the largest carve-out that leaves 60 KB (instead of 28 KB) of L1 for the
// candidate queues in local memory.
typedef uint16_t ring_t;
constexpr ring_t kRingSlotEmpty = 0xFFFFu;

struct PoolCtl {
    unsigned int head[3], tail[3], avail[3];   // ring 0 = P (needs a primary ray), ring 1 = S (secondary ready),
                                               // ring 2 = F (bundle mode: holds the last sample of a pixel chunk, must fold it)
    unsigned int inflight;                     // warps currently holding a batch
    unsigned int retired;                      // slots that ran out of work
    unsigned int spins_steady, spins_tail;     // diagnostics: acquire-loop iterations before / after work ran out
    unsigned int scanning[4];                  // warps of each SM sub-partition (warp id % 4) that are inside a per-ray slab scan
};

#ifdef SRT_SPIN_LIMIT
// Debug build (-DSRT_SPIN_LIMIT): shadow state that audits the ring protocol, indexed by (CTA, ring, physical entry).
//   g_dbg_wpos   the absolute position the entry was last written for: a pop must find its own position (site 40), a push
//                must not find its own (site 42: one position written twice)
//   g_dbg_busy   1 + lap of the consumer that is inside pool_pop on the entry: two at once (site 35) means the consumers
//                of two laps met on one entry, i.e. an entry stayed unread for a whole lap
// plus: a pop re-reads the entry after clearing it (site 41), and neither wait may last longer than kDbgWaitCycles
// (sites 36 pop, 37 push) -- the protocol's claim is that both waits are short.  Sites 16-34 are in srt_backend.cu.
constexpr unsigned int kDbgRingEntries = 2048u, kDbgCtas = 1024u;     // grids are one resident wave: <= 148 SMs x 2 CTAs here (site 15 checks)
constexpr long long kDbgWaitCycles = 400000ll;            // 0.2 ms; a warp pass is 20-50 us
__device__ unsigned int g_dbg_wpos[kDbgCtas * 3 * kDbgRingEntries];
__device__ unsigned int g_dbg_busy[kDbgCtas * 3 * kDbgRingEntries];
__device__ __forceinline__ unsigned int dbg_ring_index(uint32_t q, uint32_t entry) { return ((blockIdx.x % kDbgCtas) * 3u + q) * kDbgRingEntries + entry; }
#endif
// Inline on purpose: an out-of-line copy (one instead of seven in pool_acquire) measured 6 % slower.
__device__ __forceinline__ bool pool_try_take(PoolCtl *ctl, int q, uint32_t want_min, uint32_t &n, uint32_t &pos) {
    unsigned int old = *(volatile unsigned int *)&ctl->avail[q];
    while (old >= want_min && old > 0u) {
        const unsigned int take = old < 64u ? old : 64u;
        const unsigned int seen = atomicCAS(&ctl->avail[q], old, old - take);
        if (seen == old) {
            pos = atomicAdd(&ctl->head[q], take);
            n = take;
            atomicAdd(&ctl->inflight, 1u);
            return true;
        }
        old = seen;
    }
    return false;
}

// Lane 0 decides; returns n == 0 when every slot of the CTA has retired.  `hungry`: the caller's
// last batch held no runnable path (all its slots wait for another slot's sample chunk): it may
// take a partial batch of secondary rays, so the paths being waited for always advance.
#ifndef SRT_FOLD_MIN
#define SRT_FOLD_MIN 64
#endif
constexpr uint32_t kFoldBatchMin = SRT_FOLD_MIN;   // ring F entries that make a fold pass worth a warp's while (two chunks per lane; 64 beats 32 and 16 by 1 %)

#ifndef SRT_SCAN_TARGET
#define SRT_SCAN_TARGET 3
#endif
// Bundle-mode kernels: scanning warps an SM sub-partition wants; below it, a warp takes secondary rays first (0: off).
// Measured on the 4K / 256-spp launch: 0: 711.8 ms, 1: 705.1, 2: 693.3, 3: 684.5, 4: 688.8, 5: 687.5, always: 687.7.
constexpr uint32_t kScanTarget = SRT_SCAN_TARGET;

// An attempt = take >= min entries of ring q, one byte: 0x80 | q | code << 2 (code 0: min 1, 1: min 64, 2: min kFoldBatchMin).
// pool_acquire walks a list of attempts (one byte each, lowest byte first) with ONE inlined copy of pool_try_take: the
// nine copies of the unrolled version were 2 KB of a hot code footprint that sits right at the 32 KB of the L1.5
// instruction cache (12 % of the kernel's stall samples are instruction fetches).
__device__ __forceinline__ constexpr uint64_t pool_attempt(int q, int code) { return 0x80u | (unsigned)q | ((unsigned)code << 2); }

template <bool FOLD = false>
__device__ __forceinline__ void pool_acquire(PoolCtl *ctl, uint32_t n_slots, bool hungry, uint32_t &kind, uint32_t &n, uint32_t &pos) {
    n = 0; kind = 0; pos = 0;
    constexpr uint64_t kSteady = FOLD ? (pool_attempt(2, 2) | pool_attempt(0, 1) << 8 | pool_attempt(1, 1) << 16)   // enough chunks to fold, one per lane; then P, then S
                                      : (pool_attempt(0, 1) | pool_attempt(1, 1) << 8);
    constexpr uint64_t kDrain = FOLD ? (pool_attempt(2, 0) | pool_attempt(1, 0) << 8 | pool_attempt(0, 0) << 16)
                                     : (pool_attempt(1, 0) | pool_attempt(0, 0) << 8);
    constexpr int kSteadyBits = FOLD ? 24 : 16;
    // Only the per-ray slab scan keeps the issue port busy; refine / select / shade and the bundled primary passes are
    // latency-bound.  A sub-partition whose warps are all in those phases idles (18 % of the scheduler cycles before this),
    // so a warp that finds fewer than kScanTarget scanning warps on its sub-partition takes a batch of secondary rays
    // (= a scan) before anything else.  Bundle mode only: without it every pass scans (-3 % on the 256-sphere scene).
    uint64_t head = 0;                                    // one-shot attempts in front of the steady list
    int head_bits = 0;
    if (FOLD && kScanTarget && !hungry && *(volatile unsigned int *)&ctl->scanning[(threadIdx.x >> 5) & 3u] < kScanTarget) {
        head = pool_attempt(1, 1); head_bits = 8;
    }
    if (hungry) {                                         // the chunk being waited for may sit in ring F; else advance any secondary ray
        __nanosleep(200);
        head = FOLD ? (pool_attempt(2, 0) | pool_attempt(1, 0) << 8) : pool_attempt(1, 0);
        head_bits = FOLD ? 16 : 8;
    }
    [[maybe_unused]] uint32_t spins = 0;
    for (;;) {
        const bool quiet = *(volatile unsigned int *)&ctl->inflight == 0u;   // nobody will add to the rings: drain partial batches too
        uint64_t list = head | (kSteady << head_bits);
        if (quiet) list |= kDrain << (head_bits + kSteadyBits);
        head = 0; head_bits = 0;
#pragma unroll 1
        while (list) {
            const uint32_t a = (uint32_t)list & 0xFFu;
            list >>= 8;
            const int q = (int)(a & 3u);
            const uint32_t code = (a >> 2) & 3u;
            if (pool_try_take(ctl, q, code == 0u ? 1u : (code == 1u ? 64u : kFoldBatchMin), n, pos)) { kind = (uint32_t)q; return; }
        }
        if (quiet && *(volatile unsigned int *)&ctl->retired == n_slots) return;
        if (*(volatile unsigned int *)&ctl->retired) ctl->spins_tail += 1; else ctl->spins_steady += 1;   // racy, diagnostic only
        __nanosleep(100);
        SRT_SPIN(kSpinAcquire, spins);
    }
}

__device__ __forceinline__ uint32_t pool_pop(ring_t *ring, uint32_t cap_mask, uint32_t pos, [[maybe_unused]] uint32_t q) {
    volatile ring_t *e = ring + (pos & cap_mask);
    ring_t id;
    [[maybe_unused]] uint32_t spins = 0;
#ifdef SRT_SPIN_LIMIT
    unsigned int *busy = &g_dbg_busy[dbg_ring_index(q, pos & cap_mask)];
    const unsigned int was = atomicExch(busy, 1u + pos / (cap_mask + 1u));
    SRT_CHECKV(35, was == 0u, ((unsigned long long)was << 32) | (q << 28) | pos);
    const long long t0 = clock64();
#endif
    while ((id = *e) == kRingSlotEmpty) {
        SRT_SPIN(kSpinPop, spins);
#ifdef SRT_SPIN_LIMIT
        SRT_CHECKV(36, clock64() - t0 < kDbgWaitCycles, ((unsigned long long)q << 60) | pos);
#endif
    }
#ifdef SRT_SPIN_LIMIT
    {
        const unsigned int w = atomicAdd(&g_dbg_wpos[dbg_ring_index(q, pos & cap_mask)], 0u);
        SRT_CHECKV(40, w == pos, ((unsigned long long)q << 62) | ((unsigned long long)w << 30) | ((unsigned long long)(pos & 0x3FFFFFu) << 8) | (id & 0xFFu));
    }
#endif
    *e = kRingSlotEmpty;
#ifdef SRT_SPIN_LIMIT
    __threadfence_block();
    {
        const ring_t again = *e;
        const unsigned int w2 = atomicAdd(&g_dbg_wpos[dbg_ring_index(q, pos & cap_mask)], 0u);
        SRT_CHECKV(41, again == kRingSlotEmpty || w2 != pos, ((unsigned long long)q << 62) | ((unsigned long long)again << 40) | ((unsigned long long)id << 24) | (pos & 0xFFFFFFu));
    }
    atomicExch(busy, 0u);
#endif
    return id;
}

// Warp-aggregated push of `id` by the lanes with `doit`.
__device__ __forceinline__ void pool_push(PoolCtl *ctl, ring_t *ring, uint32_t cap_mask, int q, bool doit, uint32_t id, unsigned lane) {
    const unsigned m = __ballot_sync(0xFFFFFFFFu, doit);
    if (m == 0u) return;
    const unsigned leader = __ffs(m) - 1;
    unsigned pos = 0;
    if (lane == leader) pos = atomicAdd(&ctl->tail[q], __popc(m));
    pos = __shfl_sync(0xFFFFFFFFu, pos, leader);
    if (doit) {
        const uint32_t mine = pos + __popc(m & ((1u << lane) - 1u));
        volatile ring_t *e = ring + (mine & cap_mask);
        [[maybe_unused]] uint32_t spins = 0;
#ifdef SRT_SPIN_LIMIT
        const long long t0 = clock64();
#endif
        while (*e != kRingSlotEmpty) {
            SRT_SPIN(kSpinPush, spins);
#ifdef SRT_SPIN_LIMIT
            SRT_CHECKV(37, clock64() - t0 < kDbgWaitCycles, ((unsigned long long)q << 60) | ((unsigned long long)*e << 32) | mine);
#endif
        }
#ifdef SRT_SPIN_LIMIT
        {
            const unsigned int oldw = atomicExch(&g_dbg_wpos[dbg_ring_index((uint32_t)q, mine & cap_mask)], mine);
            SRT_CHECKV(42, oldw != mine, ((unsigned long long)q << 62) | ((unsigned long long)id << 32) | mine);
            __threadfence();
        }
#endif
        *e = (ring_t)id;
    }
    __threadfence_block();
    __syncwarp();                 // every lane has written (and is past its wait) before the ids are announced
    if (lane == leader) atomicAdd(&ctl->avail[q], __popc(m));
}

} // namespace srt
