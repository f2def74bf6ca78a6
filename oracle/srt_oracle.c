/* oracle/srt_oracle.c -- CPU restatement of the reference's path-tracing hot path.
 *
 * TEST INFRASTRUCTURE ONLY (see srt_oracle.h).  Plain C, scalar, one function per
 * reference function, each citing the file:line it follows.  Build with
 *     gcc -O2 -ffp-contract=off -mfma   (contraction OFF: every rounding is explicit;
 *                                        fmaf() appears only where profile F has one)
 * Parity status: pinned against oracle/_ref (the reference's own code compiled here).
 */
#include "srt_oracle.h"
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#define ORC_EPS 1e-4f     /* F32Epsilon base.h:889 */
#define ORC_MAX 1e30f     /* F32Max     base.h:891 */
#define ORC_PI 3.14159265358979323846f /* PI32 base.h:892 */
#define ORC_WORLD_SCALE 0.0625f        /* main.cpp:56 */

typedef struct { float x, y, z; } v3;

/* ---- v3 operators: x64_math.h:265-280 are 4-lane SSE add/sub/mul/div, i.e. one IEEE
 *      op per component; scalars broadcast through v3(f32) (x64_math.h:110-113). ------ */
static inline v3 V3(float x, float y, float z) { v3 r = { x, y, z }; return r; }
static inline v3 v3_add(v3 a, v3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 v3_sub(v3 a, v3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 v3_mul(v3 a, v3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline v3 v3_scale(float s, v3 a) { return V3(s * a.x, s * a.y, s * a.z); }
static inline v3 v3_neg(v3 a) { return V3(-a.x, -a.y, -a.z); }                 /* x64_math.h:219-222 */
static inline float v3_dot(v3 a, v3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; } /* x64_math.h:224-227 */
static inline v3 v3_cross(v3 a, v3 b) {                                         /* x64_math.h:258-264 */
    return V3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
/* x64_math.h:234-245: divide by sqrt(len^2); all-zero when len^2 <= 1e-4 */
static inline v3 v3_normalize(v3 a) {
    float lsq = v3_dot(a, a);
    if (!(lsq > ORC_EPS)) return V3(0.0f, 0.0f, 0.0f);
    float len = sqrtf(lsq);
    return V3(a.x / len, a.y / len, a.z / len);
}
void orc_normalize(const float in[3], float out[3]) {
    v3 r = v3_normalize(V3(in[0], in[1], in[2]));
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
static inline float min_ss(float a, float b) { return a < b ? a : b; }          /* _mm_min_ss x64_math.h:79-82 */
static inline float saturate(float v) {                                         /* x64_math.h:97-101 */
    if (v < 0.0f) return 0.0f;
    if (v > 1.0f) return 1.0f;
    return v;
}

/* ---- A7: PRNG base.h:954-963, 983-989; RotateRight32 x64_math.h:168-170 ------------ */
uint32_t orc_pcg(uint64_t *state) {
    uint64_t old = *state;
    *state = old * 6364136223846793005ULL + 1442695040888963407ULL;
    uint32_t v = (uint32_t)(old >> 32) ^ (uint32_t)old;
    uint32_t rot = (uint32_t)(old >> 59);
    return (v >> rot) | (v << ((32u - rot) & 31u));
}
float orc_random_float(uint64_t *state, float mn, float mx) {
    uint32_t n = orc_pcg(state);
    float inv = (float)((double)(mx - mn) / (double)4294967295u);   /* f32 sub, f64 div, round to f32 */
    float r = (float)n * inv;
    return r + mn;
}
/* ---- A10: main.cpp:667-676, index widened to 64 bits; a shift count of 0 (i % 64 == 57 / == 41, where the
 * reference's `s ^= s >> 0` zeroes the seed and every such index would share one stream) is replaced by 32:
 * the stream contract of DESIGN.md section 3.  Equal to the reference for every other index. ------------- */
uint64_t orc_stream_seed(uint64_t i) {
    uint64_t s = 0x420247153476526ULL * i;
    s += 0x8442885C91A5C8DULL;
    unsigned k1 = (unsigned)((7 + i) % 64), k2 = (unsigned)((0x29 ^ i) % 64);
    s ^= s >> (k1 ? k1 : 32);
    s ^= s << 23;
    s ^= s >> (k2 ? k2 : 32);
    s = (s * 0x11C19226CEB4769AULL) + 0x1105404122082911ULL;
    s ^= s << 19;
    s ^= s >> 13;
    return s;
}

/* ---- A6: main.cpp:292-300 ----------------------------------------------------------- */
float orc_reflectance(float c, float eta) {
    float r0 = (1.0f - eta) / (1.0f + eta);
    r0 = r0 * r0;
    float r1 = 1.0f - c;
    r1 = r1 * r1 * r1 * r1 * r1;
    return r0 + (1.0f - r0) * r1;
}

/* ---- A9: main.cpp:312-346 ----------------------------------------------------------- */
float orc_linear_to_srgb(float l) {
    l = saturate(l);
    if (l < 0.0031308f) return l * 12.92f;
    return sqrtf(l);                       /* the "bad but fast" branch, main.cpp:322-325 */
}
static inline uint32_t to_u8(float v) {    /* u8 r = Saturate(v) * 255.0f  main.cpp:341 */
    float s = saturate(v) * 255.0f;
    if (!(s == s)) return 0;               /* cvttss2si(NaN) = 0x80000000 -> low byte 0 */
    return (uint32_t)(int32_t)s & 0xFFu;
}
uint32_t orc_pack(const float rgba[4]) {
    uint32_t r = to_u8(orc_linear_to_srgb(rgba[0]));
    uint32_t g = to_u8(orc_linear_to_srgb(rgba[1]));
    uint32_t b = to_u8(orc_linear_to_srgb(rgba[2]));
    return r | (g << 8) | (b << 16) | (255u << 24);
}

/* ---- scenes: main.cpp:57-71, 93-268 -------------------------------------------------- */
static void make_sphere(orc_sphere *s, v3 p, float radius, v3 color, float spec, float ior, v3 emissive, int scale) {
    memset(s, 0, sizeof(*s));
    if (scale) {
        s->position[0] = p.x * ORC_WORLD_SCALE; s->position[1] = p.y * ORC_WORLD_SCALE;
        s->position[2] = p.z * ORC_WORLD_SCALE; s->radius = radius * ORC_WORLD_SCALE;
    } else {
        s->position[0] = p.x; s->position[1] = p.y; s->position[2] = p.z; s->radius = radius;
    }
    s->material.color[0] = color.x; s->material.color[1] = color.y; s->material.color[2] = color.z;
    s->material.emissive[0] = emissive.x; s->material.emissive[1] = emissive.y; s->material.emissive[2] = emissive.z;
    s->material.specular = spec;
    s->material.ior = ior;
}
static v3 rand_v3_seq(uint64_t *st, float mn, float mx) {   /* P8: first draw -> x */
    float x = orc_random_float(st, mn, mx);
    float y = orc_random_float(st, mn, mx);
    float z = orc_random_float(st, mn, mx);
    return V3(x, y, z);
}

static int scene_rgb(orc_sphere *o, float header[6]) {       /* main.cpp:171-191 */
    v3 zero = V3(0, 0, 0), red = V3(1.0f, 0.0f, 0.0f);
    make_sphere(o + 0, V3(0.0f, -256 - 2.0f, -15.0f), 256.0f, V3(0.2f, 0.2f, 0.2f), 0.0f, 0.0f, zero, 1);
    make_sphere(o + 1, V3(0.0f, 0, -10.0f), 2.0f, V3(1.0f, 1.0f, 1.0f), 0.0f, 1.5f, zero, 1);
    make_sphere(o + 2, V3(-4.0f, 1.0f, -15.0f), 1.5f, red, 0.0f, 0, V3(8.0f, 0.0f, 0.0f), 1);
    make_sphere(o + 3, V3(0.0f, 1.0f, -15.0f), 1.5f, red, 0.0f, 0, V3(0.0f, 8.0f, 0.0f), 1);
    make_sphere(o + 4, V3(4.0f, 1.0f, -15.0f), 1.5f, red, 0.0f, 0, V3(0.0f, 0.0f, 8.0f), 1);
    header[0] = o[1].position[0]; header[1] = o[1].position[1]; header[2] = o[1].position[2];
    header[3] = 16.0f * ORC_WORLD_SCALE;
    header[4] = (float)((double)ORC_PI / 3.0);
    header[5] = 4.0f * ORC_WORLD_SCALE;
    return 5;
}

static int scene_randomized(orc_sphere *o, float header[6]) { /* main.cpp:96-167 */
    uint64_t st = 0x29D7A0A514F22432ULL;
    orc_material mats[28];
    memset(mats, 0, sizeof(mats));
    for (int i = 0; i < 28; ++i) {
        v3 color;
        color.x = orc_random_float(&st, 0.15f, 1.0f);
        color.y = orc_random_float(&st, 0.1f, 0.75f);
        color.z = orc_random_float(&st, 0.15f, 1.0f);
        v3 emissive = V3(0, 0, 0);
        float spec = 0.0f;
        if (orc_random_float(&st, 0.0f, 1.0f) < 0.125f) {
            emissive = v3_scale(orc_random_float(&st, 2.0f, 5.0f), color);
        } else {
            float r = orc_random_float(&st, 0.0f, 1.0f);
            if (r < 0.65f) spec = 1.0f;
        }
        mats[i].color[0] = color.x; mats[i].color[1] = color.y; mats[i].color[2] = color.z;
        mats[i].emissive[0] = emissive.x; mats[i].emissive[1] = emissive.y; mats[i].emissive[2] = emissive.z;
        mats[i].ior = 0.0f; mats[i].specular = spec;
    }
#define MAT_ARGS(m) V3((m).color[0], (m).color[1], (m).color[2]), (m).specular, (m).ior, V3((m).emissive[0], (m).emissive[1], (m).emissive[2])
    float radius0 = orc_random_float(&st, 2.0f, 8.0f);
    make_sphere(o + 0, V3(1.0f, 0.0f, 0.0f), radius0, MAT_ARGS(mats[0]), 0);
    make_sphere(o + 1, V3(8.0f, -1.0f, 8.0f), radius0, MAT_ARGS(mats[0]), 0);
    make_sphere(o + 2, V3(-20.0f, -4.0f, -20.0f), radius0, MAT_ARGS(mats[0]), 0);
    for (int i = 3; i < 256; ++i) {
        v3 vec;
        vec.x = orc_random_float(&st, -1.0f, 1.0f);
        vec.y = orc_random_float(&st, -1.0f, 1.0f);
        vec.z = orc_random_float(&st, -1.0f, 1.0f);
        v3 n = v3_normalize(vec);
        float r_prev = o[i - 3].radius;
        v3 p_prev = V3(o[i - 3].position[0], o[i - 3].position[1], o[i - 3].position[2]);
        float radius = orc_random_float(&st, 1.0f, 4.0f);
        float dist = orc_random_float(&st, 1.0f, 8.0f) + radius + r_prev;
        v3 pos = v3_add(p_prev, v3_scale(dist, n));   /* NormalizedVector * (..): commutative per component */
        make_sphere(o + i, pos, radius, MAT_ARGS(mats[i % 28]), 0);
    }
#undef MAT_ARGS
    for (int i = 0; i < 256; ++i) {
        o[i].radius *= ORC_WORLD_SCALE;
        o[i].position[0] *= ORC_WORLD_SCALE; o[i].position[1] *= ORC_WORLD_SCALE; o[i].position[2] *= ORC_WORLD_SCALE;
    }
    header[0] = 2.0f * ORC_WORLD_SCALE; header[1] = 0.0f * ORC_WORLD_SCALE; header[2] = 2.0f * ORC_WORLD_SCALE;
    header[3] = 48.0f * ORC_WORLD_SCALE;
    header[4] = (float)((double)(ORC_PI * 2.65f) / 2.0);
    header[5] = 0.0f;                      /* never assigned: static zero (main.cpp:54) */
    return 256;
}

static int scene_rtweekend(orc_sphere *o488, float header[6]) { /* main.cpp:196-268 */
    v3 zero = V3(0, 0, 0);
    int idx = 0;
    make_sphere(o488 + idx++, V3(0, -1000, 0), 1000, V3(0.5f, 0.5f, 0.5f), 0.0f, 0.0f, zero, 1);
    make_sphere(o488 + idx++, V3(0, 1, 0), 1, V3(1.0f, 1.0f, 1.0f), 0.0f, 1.5f, zero, 1);
    make_sphere(o488 + idx++, V3(-4, 1, 0), 1, V3(0.4f, 0.2f, 0.1f), 0.0f, 0.0f, zero, 1);
    make_sphere(o488 + idx++, V3(4, 1, 0), 1, V3(0.7f, 0.6f, 0.5f), 1.0f, 0.0f, zero, 1);
    uint64_t st = 0xCD46749A57ACB371ULL;
    for (int i = -11; i < 11; ++i) {
        for (int j = -11; j < 11; ++j) {
            float m = orc_random_float(&st, 0.0f, 1.0f);
            v3 c;
            int t1, t2, t3;
            do {
                c.x = (float)i + orc_random_float(&st, -1.0f, 1.0f);
                c.y = 0.2f;
                c.z = (float)j + orc_random_float(&st, -1.0f, 1.0f);
                v3 d1 = v3_sub(c, V3(4, 0.2f, 0)), d2 = v3_sub(c, V3(0, 0.2f, 0)), d3 = v3_sub(c, V3(-4, 0.2f, 0));
                t1 = (double)sqrtf(v3_dot(d1, d1)) > 0.9;   /* f32 length compared with a double literal */
                t2 = (double)sqrtf(v3_dot(d2, d2)) > 0.9;
                t3 = (double)sqrtf(v3_dot(d3, d3)) > 0.9;
            } while (!t1 || !t2 || !t3);
            v3 color = zero;
            float spec = 0.0f, ior = 0.0f;
            if ((double)m < 0.8) {
                color = rand_v3_seq(&st, 0.0f, 1.0f);
            } else if ((double)m < 0.95) {
                color = rand_v3_seq(&st, 0.0f, 1.0f);
                spec = orc_random_float(&st, 0.5f, 1.0f);
            } else {
                color = V3(1.0f, 1.0f, 1.0f);
                ior = 1.5f;
            }
            make_sphere(o488 + idx++, c, 0.2f, color, spec, ior, zero, 1);
        }
    }
    header[0] = o488[1].position[0]; header[1] = o488[1].position[1]; header[2] = o488[1].position[2];
    header[3] = 12.0f * ORC_WORLD_SCALE;
    header[4] = ORC_PI / 8;
    header[5] = 2.0f * ORC_WORLD_SCALE;
    return 482;                            /* Count stays 482 of the 488 generated */
}

int orc_builtin_scene(int scene, orc_sphere *out, uint32_t capacity, int *use_sky, float header[6]) {
    static orc_sphere tmp[488];
    float h[6];
    int n;
    if (scene == 0) n = scene_rgb(tmp, h);
    else if (scene == 1) n = scene_randomized(tmp, h);
    else if (scene == 2) n = scene_rtweekend(tmp, h);
    else return -1;
    if (use_sky) *use_sky = (scene == 2);   /* main.cpp:203 */
    if (header) memcpy(header, h, sizeof(h));
    if (out) memcpy(out, tmp, sizeof(orc_sphere) * ((uint32_t)n < capacity ? (uint32_t)n : capacity));
    return n;
}

/* ---- A12: camera main.cpp:776-780, 811-822 ------------------------------------------- */
void orc_camera_from_orbit(const float look_at[3], float distance, float x_angle, float y_height,
                           uint32_t width, uint32_t height, orc_camera *out) {
    memset(out, 0, sizeof(*out));
    v3 la = V3(look_at[0], look_at[1], look_at[2]);
    float cx = cosf(x_angle) * distance, cz = sinf(x_angle) * distance;
    v3 pos = v3_add(V3(cx, y_height, cz), la);
    v3 z = v3_normalize(v3_sub(pos, la));
    v3 x = v3_normalize(v3_cross(V3(0.0f, 1.0f, 0.0f), z));
    v3 y = v3_normalize(v3_cross(z, x));
    v3 fc = v3_sub(pos, z);
    float fw = 1.0f, fh = 1.0f;
    if (width > height) fh = (float)height / (float)width;
    else fw = (float)width / (float)height;
    out->position[0] = pos.x; out->position[1] = pos.y; out->position[2] = pos.z;
    out->z[0] = z.x; out->z[1] = z.y; out->z[2] = z.z;
    out->x[0] = x.x; out->x[1] = x.y; out->x[2] = x.z;
    out->y[0] = y.x; out->y[1] = y.y; out->y[2] = y.z;
    out->film_center[0] = fc.x; out->film_center[1] = fc.y; out->film_center[2] = fc.z;
    out->film_w = fw; out->film_h = fh;
}

/* =====================================================================================
 * The tile pipeline A0-A9: RenderTile main.cpp:348-495 / RenderTileScalar main.cpp:497-640
 * ===================================================================================== */
typedef struct {
    const orc_sphere *spheres;
    uint32_t n_spheres;
    int use_sky, lane_width, fma_dot;
    const orc_camera *cam;
    uint32_t width, height, first_sample, n_samples, tap_sample;
    uint32_t sample_stride;     /* 0/1: contiguous running mean; >1 with sum_mode: sharded sum */
    int sum_mode;               /* 1: accum += radiance (multi-GPU shard), no mean, no pack */
    uint64_t stream_base;
    float *accum;
    uint32_t *rgba, *tap_prim;
    float *tap_t;
    volatile uint32_t next_row;
    uint64_t rays_total;
    pthread_mutex_t lock;
} orc_job;

typedef struct { float t; uint32_t prim; int inside; v3 normal, next_origin; int hit; } orc_hit;

/* v3x::Dot / LengthSquared in the intersection loop: profile I x64_math.h:299-309
 * (mul,mul,mul,add,add) or profile F x64_math.h:409-420 (mul,fma,fma). */
static inline float lane_dot(v3 a, v3 b, int fma_dot) {
    if (fma_dot) return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x));
    return (a.x * b.x + a.y * b.y) + a.z * b.z;
}

/* A3 + A4, SIMD-lane semantics: main.cpp:392-455.  Sphere s lives in group s / W,
 * lane s % W.  Each lane keeps its own MinT (strict <, so the earliest group wins a tie
 * inside a lane), group index and a STICKY inside flag (main.cpp:425 only ever sets it).
 * The group-level early-outs (main.cpp:411, 421) cannot change any lane's result because
 * every conditional move is masked by MoveMask.  Padding lanes hold zero spheres
 * (r = 0, so d2 < 0 never holds) and are skipped. */
static orc_hit intersect_lanes(const orc_job *J, v3 o, v3 d) {
    const int W = J->lane_width;
    float min_t[8];
    uint32_t group[8];
    int inside[8];
    v3 normal[8], next_o[8];
    for (int l = 0; l < W; ++l) {
        min_t[l] = ORC_MAX; group[l] = 0; inside[l] = 0;
        normal[l] = V3(0, 0, 0); next_o[l] = V3(0, 0, 0);
    }
    for (uint32_t s = 0; s < J->n_spheres; ++s) {
        const orc_sphere *sp = &J->spheres[s];
        const int l = (int)(s % (uint32_t)W);
        v3 sc = v3_sub(V3(sp->position[0], sp->position[1], sp->position[2]), o);   /* :401 */
        float t = lane_dot(sc, d, J->fma_dot);                                       /* :402 */
        v3 pp = v3_scale(t, d);                                                      /* :403 */
        float r2 = sp->radius * sp->radius;                                          /* :406 */
        v3 diff = v3_sub(sc, pp);
        float d2 = lane_dot(diff, diff, J->fma_dot);                                 /* :407 */
        if (!(d2 < r2)) continue;                                                    /* :409 HitMask */
        float x = sqrtf(r2 - d2);                                                    /* :413 */
        float it = t - x;                                                            /* :415 */
        int test = it < ORC_EPS;                                                     /* :416 */
        if (test) it = t + x;                                                        /* :417 */
        if (!((it < min_t[l]) && (it > ORC_EPS))) continue;                          /* :419-421 */
        v3 ip = v3_scale(it, d);                                                     /* :423 */
        if (test) inside[l] = 1;                                                     /* :425 sticky */
        group[l] = s / (uint32_t)W;                                                  /* :426 */
        min_t[l] = it;                                                               /* :427 */
        normal[l] = v3_sub(ip, sc);                                                  /* :424,428 */
        next_o[l] = v3_add(o, ip);                                                   /* :429 */
    }
    /* HorizontalMin x64_math.h:446-452 / 579-585: a min tree; with no NaNs (MinT is only
     * ever written with a value that compared < the old one) it is the plain minimum. */
    float mv = min_t[0];
    for (int l = 1; l < W; ++l) if (min_t[l] < mv) mv = min_t[l];
    orc_hit h;
    memset(&h, 0, sizeof(h));
    h.t = mv;
    if (mv == ORC_MAX) { h.hit = 0; h.prim = 0xFFFFFFFFu; return h; }               /* :433 */
    int idx = 0;
    while (!(min_t[idx] == mv)) ++idx;                                              /* FindFirstIndex :441 */
    h.hit = 1;
    h.prim = group[idx] * (uint32_t)W + (uint32_t)idx;                              /* :443 */
    h.inside = inside[idx];                                                         /* :455 */
    h.normal = normal[idx];
    h.next_origin = next_o[idx];
    return h;
}

/* A3 + A4, scalar semantics: main.cpp:541-588.  Differences from the lane version are
 * deliberate and the reference's own: d2 == r2 counts as a hit (:557), a later sphere
 * wins a tie (:568), t == 1e-4 is accepted (:569), inside is overwritten (:574). */
static orc_hit intersect_scalar(const orc_job *J, v3 o, v3 d) {
    orc_hit h;
    memset(&h, 0, sizeof(h));
    float min_t = ORC_MAX;
    uint32_t prim = 0;
    int inside = 0;
    v3 normal = V3(0, 0, 0), next_o = V3(0, 0, 0);
    for (uint32_t s = 0; s < J->n_spheres; ++s) {
        const orc_sphere *sp = &J->spheres[s];
        v3 sc = v3_sub(V3(sp->position[0], sp->position[1], sp->position[2]), o);
        float t = v3_dot(sc, d);
        v3 pp = v3_scale(t, d);
        float r2 = sp->radius * sp->radius;
        v3 diff = v3_sub(sc, pp);
        float d2 = v3_dot(diff, diff);
        if (d2 > r2) continue;
        float x = sqrtf(r2 - d2);
        float it = t - x;
        int test = it < ORC_EPS;
        if (test) it = t + x;
        if (it > min_t) continue;
        if (it < ORC_EPS) continue;
        v3 ip = v3_scale(it, d);
        inside = test;
        prim = s;
        min_t = it;
        normal = v3_sub(ip, sc);
        next_o = v3_add(o, ip);
    }
    h.t = min_t;
    if (min_t == ORC_MAX) { h.hit = 0; h.prim = 0xFFFFFFFFu; return h; }
    h.hit = 1; h.prim = prim; h.inside = inside; h.normal = normal; h.next_origin = next_o;
    return h;
}

/* One path = one sample of one pixel: main.cpp:375-482 (527-627). Returns OutputColor. */
static v3 trace_path(const orc_job *J, uint32_t px, uint32_t py, uint64_t *rng, uint64_t *rays,
                     uint32_t *tap_prim, float *tap_t) {
    const orc_camera *C = J->cam;
    v3 atten = V3(1.0f, 1.0f, 1.0f), out = V3(0.0f, 0.0f, 0.0f);
    /* A1 main.cpp:378-385 */
    float jx = orc_random_float(rng, -0.5f, 0.5f);
    float jy = orc_random_float(rng, -0.5f, 0.5f);
    float film_x = -1.0f + (((float)px + jx) * 2.0f) / (float)J->width;
    float film_y = -1.0f + (((float)py + jy) * 2.0f) / (float)J->height;
    float ax = (film_x * C->film_w) * 0.5f, ay = (film_y * C->film_h) * 0.5f;
    v3 cam_x = V3(C->x[0], C->x[1], C->x[2]), cam_y = V3(C->y[0], C->y[1], C->y[2]);
    v3 fc = V3(C->film_center[0], C->film_center[1], C->film_center[2]);
    v3 film_p = v3_add(v3_add(fc, v3_scale(ax, cam_x)), v3_scale(ay, cam_y));
    v3 o = V3(C->position[0], C->position[1], C->position[2]);
    v3 d = v3_normalize(v3_sub(film_p, o));

    for (uint32_t i = 0; i < 5; ++i) {                                   /* MaxRayBounce :387 */
        *rays += 1;                                                      /* :390 */
        orc_hit h = (J->lane_width == 1) ? intersect_scalar(J, o, d) : intersect_lanes(J, o, d);
        if (i == 0) { if (tap_t) *tap_t = h.t; if (tap_prim) *tap_prim = h.prim; }
        if (!h.hit) {                                                    /* A5 :433-440 */
            if (J->use_sky) {
                float a = (d.y + 1.0f) * 0.5f;
                float ia = 1.0f - a;
                v3 sky = v3_add(v3_scale(ia, V3(1.0f, 1.0f, 1.0f)), v3_scale(a, V3(0.5f, 0.7f, 1.0f)));
                out = v3_add(out, v3_mul(sky, atten));
            }
            break;
        }
        const orc_material *M = &J->spheres[h.prim].material;            /* A6 :443-481 */
        v3 emissive = V3(M->emissive[0], M->emissive[1], M->emissive[2]);
        v3 color = V3(M->color[0], M->color[1], M->color[2]);
        out = v3_add(out, v3_mul(emissive, atten));
        atten = v3_mul(atten, color);
        o = h.next_origin;
        float spec = M->specular;
        v3 n = v3_normalize(h.normal);
        v3 pure = v3_sub(d, v3_scale(2.0f * v3_dot(d, n), n));           /* :453 */
        if (h.inside) n = v3_neg(n);
        if (M->ior == 0.0f) {
            v3 r = v3_normalize(rand_v3_seq(rng, -1.0f, 1.0f));          /* :461 (P8 order, P9 IEEE) */
            v3 rb = v3_add(n, r);
            float oms = (float)(1.0 - (double)spec);                     /* (1.0 - Specular) is f64 :463 */
            d = v3_normalize(v3_add(v3_scale(oms, rb), v3_scale(spec, pure)));
        } else {
            float eta = h.inside ? M->ior : 1.0f / M->ior;               /* :466 */
            float c = min_ss(v3_dot(v3_neg(d), n), 1.0f);                /* :468 */
            float cc = c * c;
            float s = sqrtf((float)(1.0 - (double)cc));                  /* f64 subtract :469 */
            int cant = eta * s > 1.0f;                                   /* :471 */
            v3 perp = v3_scale(eta, v3_add(d, v3_scale(c, n)));          /* :472 */
            float k = -sqrtf(fabsf(1.0f - v3_dot(perp, perp)));          /* :473 */
            v3 par = v3_scale(k, n);
            v3 refr = v3_normalize(v3_add(perp, par));                   /* :474 */
            /* :476 -- the draw is skipped on total internal reflection (|| short-circuit),
             * and taken even for inside rays, which then always refract */
            int reflect = cant || (orc_reflectance(c, eta) > orc_random_float(rng, 0.0f, 1.0f));
            if (reflect && !h.inside) d = pure; else d = refr;
        }
    }
    return out;
}

static void *orc_worker(void *arg) {
    orc_job *J = (orc_job *)arg;
    uint64_t rays = 0;
    const uint64_t W = J->width, H = J->height;
    for (;;) {
        uint32_t y = __atomic_fetch_add(&J->next_row, 1, __ATOMIC_RELAXED);
        if (y >= J->height) break;
        for (uint32_t x = 0; x < J->width; ++x) {
            const uint64_t pix = (uint64_t)y * W + x;
            float *acc = J->accum + pix * 4;
            float prev[4] = { acc[0], acc[1], acc[2], acc[3] };
            if (J->sum_mode) {
                /* sample-sharded partial sum: plain f32 adds in sample order */
                const uint32_t stride = J->sample_stride ? J->sample_stride : 1;
                for (uint32_t i = 0; i < J->n_samples; ++i) {
                    const uint32_t k = J->first_sample + i * stride;
                    uint64_t rng = orc_stream_seed(J->stream_base + (uint64_t)k * W * H + pix);
                    v3 c = trace_path(J, x, y, &rng, &rays, 0, 0);
                    prev[0] = prev[0] + c.x; prev[1] = prev[1] + c.y; prev[2] = prev[2] + c.z;
                }
                acc[0] = prev[0]; acc[1] = prev[1]; acc[2] = prev[2]; acc[3] = 1.0f;
                continue;
            }
            if (J->first_sample == 0) { prev[0] = prev[1] = prev[2] = prev[3] = 0.0f; }
            for (uint32_t k = J->first_sample; k < J->first_sample + J->n_samples; ++k) {
                uint64_t rng = orc_stream_seed(J->stream_base + (uint64_t)k * W * H + pix);   /* P10 */
                int tap = (k == J->tap_sample);
                v3 c = trace_path(J, x, y, &rng, &rays,
                                  (tap && J->tap_prim) ? &J->tap_prim[pix] : 0,
                                  (tap && J->tap_t) ? &J->tap_t[pix] : 0);
                /* A8 main.cpp:484-489 */
                uint32_t total = k + 1;
                float wa = 1.0f / (float)total;
                float wb = (float)k / (float)total;
                prev[0] = c.x * wa + prev[0] * wb;
                prev[1] = c.y * wa + prev[1] * wb;
                prev[2] = c.z * wa + prev[2] * wb;
                prev[3] = 1.0f;
            }
            acc[0] = prev[0]; acc[1] = prev[1]; acc[2] = prev[2]; acc[3] = prev[3];
            if (J->rgba) J->rgba[pix] = orc_pack(prev);                  /* A9 main.cpp:491-492 */
        }
    }
    pthread_mutex_lock(&J->lock);
    J->rays_total += rays;
    pthread_mutex_unlock(&J->lock);
    return 0;
}

int orc_render(const orc_sphere *spheres, uint32_t n_spheres, int use_sky, int lane_width, int fma_dot,
               const orc_camera *camera, uint32_t width, uint32_t height,
               uint32_t first_sample, uint32_t n_samples, uint64_t stream_base,
               float *accum, uint32_t *rgba, uint32_t *tap_prim, float *tap_t, uint32_t tap_sample,
               uint64_t *rays, int threads) {
    if (!spheres || !camera || !accum || width == 0 || height == 0) return -1;
    if (lane_width != 1 && lane_width != 4 && lane_width != 8) return -2;
    orc_job J;
    memset(&J, 0, sizeof(J));
    J.spheres = spheres; J.n_spheres = n_spheres; J.use_sky = use_sky;
    J.lane_width = lane_width; J.fma_dot = fma_dot; J.cam = camera;
    J.width = width; J.height = height; J.first_sample = first_sample; J.n_samples = n_samples;
    J.tap_sample = tap_sample; J.stream_base = stream_base;
    J.accum = accum; J.rgba = rgba; J.tap_prim = tap_prim; J.tap_t = tap_t;
    pthread_mutex_init(&J.lock, 0);
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    pthread_t tid[256];
    for (int i = 1; i < threads; ++i) pthread_create(&tid[i], 0, orc_worker, &J);
    orc_worker(&J);
    for (int i = 1; i < threads; ++i) pthread_join(tid[i], 0);
    pthread_mutex_destroy(&J.lock);
    if (rays) *rays = J.rays_total;
    return 0;
}

/* Sample-sharded partial SUM (the multi-GPU path): accum += radiance of samples
 * first, first+stride, ... (n of them); then orc_resolve_sum() after the cross-rank sum. */
int orc_render_sum(const orc_sphere *spheres, uint32_t n_spheres, int use_sky, int lane_width, int fma_dot,
                   const orc_camera *camera, uint32_t width, uint32_t height,
                   uint32_t first_sample, uint32_t n_samples, uint32_t sample_stride, uint64_t stream_base,
                   float *accum, uint64_t *rays, int threads) {
    if (!spheres || !camera || !accum || width == 0 || height == 0 || sample_stride == 0) return -1;
    if (lane_width != 1 && lane_width != 4 && lane_width != 8) return -2;
    orc_job J;
    memset(&J, 0, sizeof(J));
    J.spheres = spheres; J.n_spheres = n_spheres; J.use_sky = use_sky;
    J.lane_width = lane_width; J.fma_dot = fma_dot; J.cam = camera;
    J.width = width; J.height = height; J.first_sample = first_sample; J.n_samples = n_samples;
    J.sample_stride = sample_stride; J.sum_mode = 1; J.stream_base = stream_base; J.accum = accum;
    pthread_mutex_init(&J.lock, 0);
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    pthread_t tid[256];
    for (int i = 1; i < threads; ++i) pthread_create(&tid[i], 0, orc_worker, &J);
    orc_worker(&J);
    for (int i = 1; i < threads; ++i) pthread_join(tid[i], 0);
    pthread_mutex_destroy(&J.lock);
    if (rays) *rays = J.rays_total;
    return 0;
}

/* mean = sum * (1/N) (one f32 reciprocal, one multiply per channel), then A9. */
void orc_resolve_sum(float *accum, uint32_t *rgba, uint32_t n_pixels, uint32_t total_samples) {
    const float inv = 1.0f / (float)total_samples;
    for (uint32_t i = 0; i < n_pixels; ++i) {
        float *a = accum + (size_t)i * 4;
        a[0] = a[0] * inv; a[1] = a[1] * inv; a[2] = a[2] * inv; a[3] = 1.0f;
        if (rgba) rgba[i] = orc_pack(a);
    }
}
