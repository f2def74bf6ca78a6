// srt_stream.cuh -- warp-private TMA staging of sphere tables that do not fit in shared memory
// (BASELINE configs[4]: 65,536 spheres = 1 MiB per table).
//
// The pool kernel's warps are not in lock-step (each is somewhere in its own primary or
// secondary pass), so the scene is not tiled per CTA: every warp streams the table it is
// scanning through its own small ring of shared-memory stages.  One elected lane issues
// `cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes` (the 1-D TMA bulk copy)
// for the next chunk of kStreamChunk spheres; all lanes wait on the stage's mbarrier parity,
// scan the chunk with warp-uniform LDS.128, `__syncwarp()`, and the stage is refilled.  The
// tables stay L2-resident (3 x 1 MiB against 126 MB), so the copies never touch HBM after the
// first pass.
#pragma once
#include <stdint.h>
#include "srt_device.cuh"

namespace srt {

#ifndef SRT_STREAM_CHUNK
#define SRT_STREAM_CHUNK 256
#endif
#ifndef SRT_STREAM_STAGES
#define SRT_STREAM_STAGES 2
#endif
constexpr int kStreamChunk = SRT_STREAM_CHUNK;        // spheres per stage
constexpr int kStreamStageF4 = kStreamChunk * 3 / 4;  // float4 per stage: the slab table has 12 B per sphere
constexpr int kStreamStages = SRT_STREAM_STAGES;      // stages per warp

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

// Elected lane: arm the barrier with the byte count and start the bulk copy global -> shared.
__device__ __forceinline__ void stream_issue(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_addr(dst)), "l"(src), "r"(bytes), "r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
#ifdef SRT_SPIN_LIMIT
    uint32_t spins = 0;
    for (;;) {
        uint32_t ok;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(smem_addr(bar)), "r"(parity) : "memory");
        if (ok) break;
        SRT_SPIN(kSpinMbar, spins);
    }
#else
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}"
                 ::"r"(smem_addr(bar)), "r"(parity) : "memory");
#endif
}

// Per-warp stream state: the ring of stages and the parity each stage's barrier will show next.
struct WarpStream {
    float4 *stage;          // kStreamStages x kStreamStageF4 float4, this warp's
    uint64_t *bar;          // kStreamStages barriers, this warp's
    uint32_t parity;        // bit s = parity to wait for on stage s
};

} // namespace srt
