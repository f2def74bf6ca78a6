/* include/srt_backend.h -- C ABI of the B200 (sm_100a) path-tracing backend.
 *
 * This is the drop-in boundary for ONE hot path of Ne0nWinds/SIMD-Ray-Tracer: the per-tile
 * callbacks RenderTile / RenderTileScalar (main.cpp:348-640) that the platform work queue
 * runs when the application calls WorkQueueStart (main.cpp:852, base.h:174).  A maintainer
 * of the reference replaces that kick with srt_render(); WorkQueueHasCompleted /
 * WorkQueueWaitUntilCompletion / WorkQueueGetCompletionTime (base.h:175-178) map to
 * srt_poll / srt_wait.  INTEGRATION.md shows the exact binding.
 *
 * Plain C, plain pointers and sizes; no C++ or torch types.  Every function returns
 * SRT_OK (0) or a negative srt_status; srt_last_error() describes the last failure.
 * The library has NO CPU fallback: without a CUDA device srt_create fails.
 * One host thread per handle; at most one render in flight per handle.
 * Several devices: one handle per device (srt_create(.., device_id)), one host thread or
 * process each, joined by srt_comm_init_rank -- the counterpart of the ThreadCount argument of
 * WorkQueueCreate / WorkQueueStart (base.h:172-174, main.cpp:660, 851-856).
 */
#ifndef SRT_BACKEND_H
#define SRT_BACKEND_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    SRT_OK = 0,
    SRT_ERR_INVALID = -1,     /* bad argument / call order */
    SRT_ERR_CUDA = -2,        /* CUDA runtime error (see srt_last_error) */
    SRT_ERR_NO_DEVICE = -3,   /* no usable sm_100 device */
    SRT_ERR_TOO_LARGE = -4,   /* scene / image exceeds a backend limit */
    SRT_ERR_COMM = -5         /* NCCL missing or a collective failed (see srt_last_error) */
} srt_status;

/* POD mirror of `material`, main.cpp:11-16 (v3 is 16 bytes: x,y,z,_w; base.h:357-375). 48 B */
typedef struct srt_material {
    float color[4];
    float emissive[4];
    float specular;
    float index_of_refraction;
    float _pad[2];
} srt_material;

/* POD mirror of `scalar_sphere`, main.cpp:17-21. 80 B: Scene.ScalarSpheres.Data can be
 * passed as is. */
typedef struct srt_sphere {
    float position[4];
    float radius;
    float _pad[3];
    srt_material material;
} srt_sphere;

/* POD mirror of the geometric part of `camera_info`, main.cpp:270-277. */
typedef struct srt_camera {
    float camera_position[4];
    float camera_z[4];
    float camera_x[4];
    float camera_y[4];
    float film_center[4];
    float film_w;
    float film_h;
    float _pad[2];
} srt_camera;

/* Which tile callback's semantics to reproduce (render_params.EnableSIMD, base.h:157-161,
 * selected at main.cpp:852-856), and which arithmetic profile of x64_math.h. */
typedef struct srt_scene_options {
    int32_t use_sky_color;  /* scene.UseSkyColor, main.cpp:44 */
    int32_t lane_width;     /* 8 or 4: RenderTile with SIMD_WIDTH lanes (tie-break + sticky
                               inside flag of main.cpp:419-427); 1: RenderTileScalar */
    int32_t fma_dot;        /* 1: v3x8::Dot/LengthSquared use FMA (x64_math.h:409-420);
                               0: mul,mul,mul,add,add (x64_math.h:299-309) */
    int32_t _reserved;      /* 0. (non-zero selects diagnostic kernel variants: 1 = scalar-issue loop) */
} srt_scene_options;

typedef enum {
    SRT_ACCUM_MEAN = 0,  /* exact running mean of main.cpp:484-489 (single device) */
    SRT_ACCUM_SUM = 1    /* plain f32 sum of the samples' radiance; combine across devices,
                            then srt_resolve_sum() */
} srt_accum_mode;

typedef struct srt_render_args {
    uint32_t first_sample;   /* absolute index k of the first sample (PreviousRayCount, main.cpp:7) */
    uint32_t n_samples;      /* samples per pixel rendered by this call */
    uint32_t sample_stride;  /* k advances by this; 1 = contiguous; G for sample sharding over G devices */
    uint32_t accum_mode;     /* srt_accum_mode */
    uint64_t stream_base;    /* per-path RNG stream = scramble(stream_base + k*W*H + y*W + x),
                                scramble = main.cpp:667-676 on a 64-bit index, a zero shift count
                                (i % 64 == 57 / 41: the reference's expression zeroes the seed there)
                                replaced by 32 -- DESIGN.md section 3 */
    uint32_t record_primary; /* 1: record primary-ray hit/primitive/t of sample `first_sample` */
    uint32_t _reserved;
} srt_render_args;

typedef struct srt_backend srt_backend;

/* Lifetime. device_id: CUDA ordinal.  On failure *out is NULL (nothing to destroy) and
 * srt_last_error(NULL) holds the reason. */
int srt_create(srt_backend **out, int device_id);
void srt_destroy(srt_backend *b);
const char *srt_last_error(const srt_backend *b);   /* b may be NULL: last create error */

/* Scene upload (replaces ConvertScalarSpheresToSIMDSpheres' SoA, main.cpp:73-91): copies;
 * the caller keeps ownership of `spheres`. */
int srt_set_scene(srt_backend *b, const srt_sphere *spheres, uint32_t n_spheres,
                  const srt_scene_options *options);

/* (Re)allocate PreviousImage (f32x4) + CurrentImage (RGBA8) on the device and reset the
 * accumulation, like main.cpp:797-804.  width, height <= 65535 and width * height < 2^31. */
int srt_resize(srt_backend *b, uint32_t width, uint32_t height);
/* Reset only (camera moved / key R, main.cpp:791-798): zeroes the accumulator. */
int srt_reset_accumulation(srt_backend *b);
/* Optional: render into caller-owned DEVICE buffers (accum: W*H*4 f32, rgba: W*H u32),
 * e.g. to run a collective on them. NULL restores the internal buffers. */
int srt_bind_device_buffers(srt_backend *b, void *accum_dev, void *rgba_dev);

/* Kick one batch: the WorkQueueStart replacement. Asynchronous. */
int srt_render(srt_backend *b, const srt_camera *camera, const srt_render_args *args);
/* srt_render that allocates NOTHING (no cudaMalloc / cudaFree, no allocating library routine).  No counterpart in the
 * reference: added for callers that cannot allocate -- code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc.  The one piece of scratch memory the render kernels
 * need on the fly (the pool kernel's per-pixel chunk counters) lives in a file-scope device array of
 * SRT_STATIC_MAX_PIXELS x 4 bytes = 8 MiB that every GPU loading the library holds until the process exits.
 *   - images of more than SRT_STATIC_MAX_PIXELS pixels are rejected with SRT_ERR_TOO_LARGE (use srt_render);
 *   - record_primary is rejected (its buffers are allocated by srt_render);
 *   - same kernels and bit-identical buffers as srt_render; launches that srt_render runs in bundle mode (staging
 *     rows it allocates) run the plain pool kernel here, so large scenes at >= 8 samples per call are slower;
 *   - the array is shared by every handle and call on a device and keeps the last call's contents: the part a
 *     launch reads is reset on the handle's stream first, and calls that OVERLAP on the GPU (several handles of
 *     one device rendering at once) race -- serialise them.
 * srt_set_scene / srt_resize (which do allocate) must have been called before, as for srt_render. */
#define SRT_STATIC_MAX_PIXELS (2u << 20)
int srt_render_static(srt_backend *b, const srt_camera *camera, const srt_render_args *args);
/* 1 = finished, 0 = still running (WorkQueueHasCompleted). */
int srt_poll(srt_backend *b);
/* Block until finished (WorkQueueWaitUntilCompletion); out_ms = device time of the last
 * batch (CUDA events), out_rays = path segments cast by it, counted like main.cpp:390. */
int srt_wait(srt_backend *b, double *out_ms, uint64_t *out_rays);

/* Diagnostic: soundness audit of the ray-pair kernel's conservative sphere filters
 * (csrc/srt_pairs.cuh: level 1 slab test, level 2 fused 7-FMA test).  Traces the same paths as srt_render(first_sample, n_samples) would,
 * with the reference's per-lane arithmetic, and for EVERY (path segment, sphere) pair evaluates
 * both the reference's hit predicate (main.cpp:409 / :557) and the filter predicate.  Blocking.
 * out_counts = { tests, exact hits, candidates (pass both levels), violations, level-1 survivors,
 *                primary-ray tests of bundled pixels, level-0 (pixel bundle, csrc/srt_bundle.cuh) candidates,
 *                level-0 violations };
 * the filters are sound iff both violation counts (an exact hit a level would have dropped) are 0.
 * Touches no image buffer. */
int srt_filter_audit(srt_backend *b, const srt_camera *camera, const srt_render_args *args, uint64_t out_counts[8]);

/* accum holds a SUM over `total_samples`: mean = sum * (1/N), tonemap + pack into rgba
 * (main.cpp:312-346). Asynchronous on the backend stream; ordered after prior renders. */
int srt_resolve_sum(srt_backend *b, uint32_t total_samples);

/* Read-backs into HOST buffers (row 0 = image bottom, as in the reference). */
int srt_read_rgba8(srt_backend *b, void *dst /* W*H*4 bytes */);
int srt_read_accum(srt_backend *b, float *dst /* W*H*4 floats */);
int srt_read_primary(srt_backend *b, uint32_t *prim /* 0xFFFFFFFF = miss */, float *t /* 1e30 = miss */);
/* Under an image partition (srt_set_partition, n_parts > 1) the pixels of the other parts read back with all bits set
 * in both buffers (prim 0xFFFFFFFF, t = the NaN 0xFFFFFFFF); a handle that owns no row at all still answers. */
/* Upload a HOST accumulator (W*H*4 floats, as srt_read_accum returned it) into PreviousImage:
 * resume an interrupted accumulation -- the next srt_render continues with
 * first_sample = the number of samples the buffer holds (checkpoint/resume; the reference keeps
 * PreviousImage only in its arena, main.cpp:801-804). */
int srt_write_accum(srt_backend *b, const float *src /* W*H*4 floats */);   /* a partitioned handle keeps its own rows only */

/* ---- several devices (one handle each) -----------------------------------------------------
 * The reference fans a frame out to ThreadCount worker threads that pull 32x32 tiles
 * (WorkQueueStart base.h:174, main.cpp:851-856; tiles main.cpp:9, 362-368, 824-827).  Here the
 * workers are GPUs.  Two partitions of a frame are supported:
 *   image partition  srt_set_partition(part, n_parts): the handle renders every n_parts-th row of
 *                    8x4-pixel blocks (rows part, part + n_parts, ...) with ALL samples and the exact
 *                    running mean; rows of other devices stay zero, so srt_gather_rgba /
 *                    srt_gather_accum (an NCCL sum) return an image that is BIT-IDENTICAL to a
 *                    single-device render;
 *   sample partition srt_render(first_sample = rank, sample_stride = n_ranks, SRT_ACCUM_SUM) on every
 *                    device, srt_reduce_accum(root), srt_resolve_sum(total) on the root (sum-then-divide:
 *                    within the stated tolerance of the running mean, not bit-identical).
 * NCCL is loaded at run time (the libnccl.so.2 already in the process, else the loader's); the
 * collectives run on the handle's stream, ordered after its renders.  Every rank must make the same
 * sequence of srt_comm_init_rank / srt_reduce_accum / srt_gather_* calls. */
#define SRT_COMM_ID_BYTES 128
int srt_comm_unique_id(void *id_out /* SRT_COMM_ID_BYTES; made by one rank, given to all */);
int srt_comm_init_rank(srt_backend *b, int n_ranks, int rank, const void *unique_id);   /* collective */
int srt_comm_destroy(srt_backend *b);
/* A host thread that drives SEVERAL handles brackets the init calls, and each round of srt_reduce_accum /
 * srt_gather_* calls over its handles, with these (ncclGroupStart / ncclGroupEnd). */
int srt_comm_group_begin(void);
int srt_comm_group_end(void);
int srt_comm_info(srt_backend *b, int *rank, int *n_ranks /* 0: no communicator */, int *nccl_version);
/* Image partition of this handle; resets the accumulation when it changes. Default (0, 1): whole image. */
int srt_set_partition(srt_backend *b, uint32_t part, uint32_t n_parts);
/* In-place f32 sum of the accumulators onto `root` (root < 0: onto every rank). */
int srt_reduce_accum(srt_backend *b, int root);
/* Combine the row bands of an image partition on `root` (root < 0: on every rank).  Afterwards
 * srt_read_rgba8 / srt_read_accum on a receiving rank return the combined image, until its next
 * render or reset. */
int srt_gather_rgba(srt_backend *b, int root);
int srt_gather_accum(srt_backend *b, int root);

/* Diagnostic: run the device functions the kernels are built from on caller-provided records (host
 * buffers), n records, one per thread.  Record layouts (in -> out):
 *   SRT_KAT_PCG            u64 state                      -> 16 x u32 draws        (base.h:954-963)
 *   SRT_KAT_RANDOM_FLOAT   {u64 state, f32 InverseMaxInt, f32 Min} -> 16 x f32     (base.h:983-989)
 *   SRT_KAT_STREAM_SEED    u64 index                      -> u64                    (main.cpp:667-676)
 *   SRT_KAT_NORMALIZE      3 x f32                        -> 3 x f32                (x64_math.h:234-245)
 *   SRT_KAT_REFLECTANCE    {f32 cos, f32 eta}             -> f32                    (main.cpp:292-300)
 *   SRT_KAT_LINEAR_TO_SRGB f32                            -> f32                    (main.cpp:312-322)
 *   SRT_KAT_PACK           4 x f32 {r, g, b, -} linear    -> u32 RGBA8              (main.cpp:324-346) */
typedef enum {
    SRT_KAT_PCG = 0, SRT_KAT_RANDOM_FLOAT, SRT_KAT_STREAM_SEED, SRT_KAT_NORMALIZE, SRT_KAT_REFLECTANCE,
    SRT_KAT_LINEAR_TO_SRGB, SRT_KAT_PACK, SRT_KAT_COUNT
} srt_kat_op;
int srt_device_kat(srt_backend *b, uint32_t op, const void *in, uint32_t n, void *out);

/* Device pointers / stream of the current buffers (for collectives and graphs). */
void *srt_accum_device_ptr(srt_backend *b);
void *srt_rgba_device_ptr(srt_backend *b);
void *srt_cuda_stream(srt_backend *b);

/* Introspection for benchmarks: the render kernel the last srt_render / srt_render_static launched, spelled as ncu and
 * cuobjdump print it (e.g. "render_pool<8, 1, 896, 1, 0, 1>"); "" before the first render. */
int srt_last_kernel(srt_backend *b, char *name, uint32_t capacity);
/* Introspection for benchmarks: SM count, clock (kHz), kernel launches issued so far. */
int srt_device_info(srt_backend *b, int *sm_count, int *clock_khz, uint64_t *launches);

#ifdef __cplusplus
}
#endif
#endif /* SRT_BACKEND_H */
