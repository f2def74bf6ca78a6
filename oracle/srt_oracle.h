/* oracle/srt_oracle.h -- CPU restatement of the reference's path-tracing hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this.  The product
 * (simd-ray-tracer_b200/, include/) never links, loads or calls it.
 *
 * Parity status: PINNED.  The reference ships no tests or golden vectors (SURVEY.md
 * section 4), so this restatement is pinned against the reference's own code compiled
 * here (oracle/_ref, see oracle/build_ref.py): bit-exact accumulators, RGBA8 images,
 * primary-ray primitive ids / t, scene tables, camera and RNG streams
 * (tests/test_oracle_vs_ref.py), plus committed goldens under tests/golden/.
 *
 * All citations are file:line in /root/reference.
 */
#ifndef SRT_ORACLE_H
#define SRT_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* POD mirrors of main.cpp:11-21 (v3 is 16 bytes: x,y,z,_w -- base.h:357-375). */
typedef struct { float color[4]; float emissive[4]; float specular; float ior; float _pad[2]; } orc_material; /* 48 B */
typedef struct { float position[4]; float radius; float _pad[3]; orc_material material; } orc_sphere;      /* 80 B */
/* camera_info main.cpp:270-277 */
typedef struct { float position[4], z[4], x[4], y[4], film_center[4]; float film_w, film_h; float _pad[2]; } orc_camera;

/* A7  base.h:954-963, 983-989 */
uint32_t orc_pcg(uint64_t *state);
float orc_random_float(uint64_t *state, float mn, float mx);
/* A10 main.cpp:667-676 on a 64-bit index; a zero shift count (i % 64 == 57 / 41) replaced by 32 */
uint64_t orc_stream_seed(uint64_t index);
/* A2  x64_math.h:234-245 (IEEE; NormalizeFast == Normalize per wasm_math.h:168-170) */
void orc_normalize(const float in[3], float out[3]);
/* A6  main.cpp:292-300 */
float orc_reflectance(float cos_theta, float eta);
/* A9  main.cpp:312-346 */
float orc_linear_to_srgb(float l);
uint32_t orc_pack(const float rgba[4]);

/* Built-in scenes, main.cpp:93-268.  header = LookAt.xyz, DefaultDistanceFromLookAt,
 * DefaultXAngle, DefaultYHeight.  Returns sphere count (5 / 256 / 482) or <0. */
int orc_builtin_scene(int scene, orc_sphere *out, uint32_t capacity, int *use_sky, float header[6]);
/* A12 main.cpp:776-780, 811-822 */
void orc_camera_from_orbit(const float look_at[3], float distance, float x_angle, float y_height,
                           uint32_t width, uint32_t height, orc_camera *out);

/* A0..A9: renders samples [first_sample, first_sample+n_samples) of every pixel with
 * the per-(sample,pixel) stream contract:
 *     state = orc_stream_seed(stream_base + sample*W*H + y*W + x)
 * lane_width 4 or 8 = RenderTile SIMD-lane semantics (main.cpp:348-495),
 * lane_width 1      = RenderTileScalar semantics   (main.cpp:497-640);
 * fma_dot != 0      = profile F dot/len^2 (x64_math.h:409-420) in the intersection loop.
 * accum (W*H*4 f32, running mean, in/out) must hold the mean of samples
 * [0, first_sample) on entry (ignored when first_sample == 0).  rgba, tap_prim, tap_t
 * may be NULL; taps record the primary ray of sample `tap_sample`
 * (prim 0xFFFFFFFF and t 1e30 on miss).  Row 0 is the image bottom. */
int orc_render(const orc_sphere *spheres, uint32_t n_spheres, int use_sky, int lane_width, int fma_dot,
               const orc_camera *camera, uint32_t width, uint32_t height,
               uint32_t first_sample, uint32_t n_samples, uint64_t stream_base,
               float *accum, uint32_t *rgba, uint32_t *tap_prim, float *tap_t, uint32_t tap_sample,
               uint64_t *rays, int threads);

/* Multi-GPU restatement (no counterpart in the reference): sharded f32 sum and
 * its resolve. */
int orc_render_sum(const orc_sphere *spheres, uint32_t n_spheres, int use_sky, int lane_width, int fma_dot,
                   const orc_camera *camera, uint32_t width, uint32_t height,
                   uint32_t first_sample, uint32_t n_samples, uint32_t sample_stride, uint64_t stream_base,
                   float *accum, uint64_t *rays, int threads);
void orc_resolve_sum(float *accum, uint32_t *rgba, uint32_t n_pixels, uint32_t total_samples);

#ifdef __cplusplus
}
#endif
#endif
