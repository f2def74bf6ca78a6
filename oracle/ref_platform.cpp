// oracle/ref_platform.cpp -- headless Linux platform layer for the UNMODIFIED-LOGIC
// reference (patched at build time by oracle/build_ref.py; see its docstring).
//
// TEST INFRASTRUCTURE ONLY: this file plus the reference's own base.h / x64_math.h /
// main.cpp form oracle/_ref/libsrt_ref_w{4,8}.so, the parity anchor for oracle/srt_oracle.c
// and the timed "reference" CPU baseline.  The product never links or loads it.
//
// It plays the role wasm/wasm.cpp plays in the reference: it implements the platform
// half of the base.h contract (base.h:105-115 arena, base.h:172-178 work queue,
// base.h:318 IsDown, base.h:1000 timer) and drives OnInit / OnRender (base.h:163-164).
#include <immintrin.h>
#include <x86intrin.h>
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

// P2: g++ rejects vector_size on struct types; the aligned attribute survives.
#define __vector_size__(x)
// P4: MSVC/clang-cl rotate intrinsics -> GNU spellings.
#define _rotr64 __rorq
#define _rotl64 __rolq

// Run-time switches consulted by the patched sources (defaults = as shipped).
static bool SrtRefIEEENormalizeFast = false;   // P9
static bool SrtRefParityRng = false;           // P10
static uint64_t SrtRefStreamBase = 0;          // P10
static uint32_t *SrtRefTapPrim = 0;            // debug tap: primary-ray primitive id
static float *SrtRefTapT = 0;                  // debug tap: primary-ray MinT
static uint32_t SrtRefTapSample = 0;           // which sample index the taps record

// Per-stream seed: main.cpp:667-676 evaluated on a 64-bit index.  A shift count of zero
// (i % 64 == 57 / == 41: `S ^= S >> 0` zeroes the seed, so all those indices would share ONE stream) is
// replaced by 32; every other index gets the reference's value (checked against srt_ref_thread_seed).
static inline uint64_t SrtRefStreamSeed(uint64_t i) {
    uint64_t S = 0x420247153476526ULL * i;
    S += 0x8442885C91A5C8DULL;
    const unsigned K1 = (unsigned)((7 + i) % 64), K2 = (unsigned)((0x29 ^ i) % 64);
    S ^= S >> (K1 ? K1 : 32);
    S ^= S << 23;
    S ^= S >> (K2 ? K2 : 32);
    S = (S * 0x11C19226CEB4769AULL) + 0x1105404122082911ULL;
    S ^= S << 19;
    S ^= S >> 13;
    return S;
}

#define SRT_REF_PIXEL_PROLOGUE(x, y)                                                          \
    if (SrtRefParityRng) {                                                                    \
        RandomState.Seed = SrtRefStreamSeed(SrtRefStreamBase +                                \
            (u64)PreviousRayCount * (u64)CurrentImage.Width * (u64)CurrentImage.Height +      \
            (u64)(y) * (u64)CurrentImage.Width + (u64)(x));                                   \
    }
#define SRT_REF_TAP_T(i, x, y, v)                                                             \
    if ((i) == 0 && SrtRefTapT && PreviousRayCount == SrtRefTapSample)                        \
        SrtRefTapT[(y) * CurrentImage.Width + (x)] = (v);
#define SRT_REF_TAP_PRIM(i, x, y, v)                                                          \
    if ((i) == 0 && SrtRefTapPrim && PreviousRayCount == SrtRefTapSample)                     \
        SrtRefTapPrim[(y) * CurrentImage.Width + (x)] = (v);

#include "base.h"   // patched copy (temp dir on the include path)

// P8 helper: three draws, strictly sequenced, first draw -> x (clang's order).
static inline v3 SrtRefRandV3(u32_random_state &S, f32 Min, f32 Max) {
    f32 X = S.RandomFloat(Min, Max);
    f32 Y = S.RandomFloat(Min, Max);
    f32 Z = S.RandomFloat(Min, Max);
    return v3(X, Y, Z);
}

// P6: members declared (base.h:503-504,546) but defined only in wasm_math.h:286-323.
// Semantics: first lane equal to Value; lane extract.
inline u32 f32x4::FindFirstIndex(const f32x4 &A, f32 Value) {
    xmm Comparison = A == f32x4(Value);
    u32 MoveMask = _mm_movemask_ps(Comparison);
    return _tzcnt_u32(MoveMask);
}
inline f32 f32x4::Extract(u32 Index) { return Value[Index]; }
inline u32 u32x4::Extract(u32 Index) { return Value[Index]; }
#if SIMD_WIDTH >= 8
inline u32 f32x8::FindFirstIndex(const f32x8 &A, f32 Value) {
    ymm Comparison = A == f32x8(Value);
    u32 MoveMask = _mm256_movemask_ps(Comparison);
    return _tzcnt_u32(MoveMask);
}
inline f32 f32x8::Extract(u32 Index) { return Value[Index]; }
inline u32 u32x8::Extract(u32 Index) { return Value[Index]; }
#endif

#include "main.cpp" // patched copy; unity build gives access to its file statics

// ===================================================================================
// Platform half of the base.h contract
// ===================================================================================

// ---- arena (base.h:105-115; zero-filling Push like wasm/wasm.cpp:52) ----------------
memory_arena AllocateArenaFromOS(u32 Size, u64 /*StartingAddress*/) {
    memory_arena Arena = {};
    void *Mem = 0;
    if (posix_memalign(&Mem, 4096, (size_t)Size + 4096) != 0) Mem = 0;
    Arena.Start = Mem;
    Arena.Offset = Mem;
    Arena.Size = Size;
    return Arena;
}
void *memory_arena::Push(u64 PushSize, u32 Alignment) {
    uintptr_t P = (uintptr_t)Offset;
    P = (P + (Alignment - 1)) & ~(uintptr_t)(Alignment - 1);
    if (P + PushSize > (uintptr_t)Start + Size) return 0;
    Offset = (void *)(P + PushSize);
    memset((void *)P, 0, PushSize);
    return (void *)P;
}
void memory_arena::Pop(void *Ptr) { Offset = Ptr; }
void memory_arena::Reset() { Offset = Start; }
memory_arena memory_arena::CreateScratch() {
    memory_arena Result = {};
    Result.Start = Offset;
    Result.Offset = Offset;
    Result.Size = Size - (u32)((uintptr_t)Offset - (uintptr_t)Start);
    return Result;
}

// ---- timer (base.h:1000) -------------------------------------------------------------
f64 QueryTimestampInMilliseconds() {
    timespec Ts;
    clock_gettime(CLOCK_MONOTONIC, &Ts);
    return (f64)Ts.tv_sec * 1000.0 + (f64)Ts.tv_nsec * 1e-6;
}

// ---- input (base.h:318): headless; R can be "pressed" once to force a reset ----------
static bool SrtRefPressR = false;
bool IsDown(key Key) { return Key == key::R ? SrtRefPressR : false; }

// ---- work queue (base.h:172-178; semantics of wasm/wasm.cpp:604-694) ------------------
static u32 SrtRefMaxThreads = 1;
static pthread_t *SrtRefThreads = 0;
static pthread_mutex_t SrtRefLock = PTHREAD_MUTEX_INITIALIZER;
static pthread_cond_t SrtRefKick = PTHREAD_COND_INITIALIZER;
static pthread_cond_t SrtRefDone = PTHREAD_COND_INITIALIZER;
static thread_callback SrtRefCallback = 0;
static u32 SrtRefItemCount = 0;
static u32 SrtRefActiveThreads = 0;    // threads taking part in the current batch
static u64 SrtRefGeneration = 0;       // bumped per WorkQueueStart
static volatile u32 SrtRefWorkIndex = 0;
static volatile u32 SrtRefWorkCompleted = 0;
static u32 SrtRefThreadsRunning = 0;
static f64 SrtRefCompletionTime = 0.0;
static bool SrtRefQueueCreated = false;

static void *SrtRefWorker(void *Arg) {
    u32 ThreadIndex = (u32)(uintptr_t)Arg;
    u64 SeenGeneration = 0;
    for (;;) {
        pthread_mutex_lock(&SrtRefLock);
        while (SrtRefGeneration == SeenGeneration || ThreadIndex >= SrtRefActiveThreads) {
            if (SrtRefGeneration != SeenGeneration) SeenGeneration = SrtRefGeneration; // sit this one out
            pthread_cond_wait(&SrtRefKick, &SrtRefLock);
        }
        SeenGeneration = SrtRefGeneration;
        thread_callback Callback = SrtRefCallback;
        u32 ItemCount = SrtRefItemCount;
        pthread_mutex_unlock(&SrtRefLock);

        for (;;) {
            u32 WorkEntry = __atomic_fetch_add(&SrtRefWorkIndex, 1, __ATOMIC_RELAXED);
            if (WorkEntry >= ItemCount) break;
            work_queue_context Context = { WorkEntry, ThreadIndex };
            Callback(&Context);
            __atomic_fetch_add(&SrtRefWorkCompleted, 1, __ATOMIC_RELEASE);
        }

        pthread_mutex_lock(&SrtRefLock);
        SrtRefThreadsRunning -= 1;
        if (SrtRefThreadsRunning == 0) {
            SrtRefCompletionTime = QueryTimestampInMilliseconds(); // last finisher stamps
            pthread_cond_broadcast(&SrtRefDone);
        }
        pthread_mutex_unlock(&SrtRefLock);
    }
    return 0;
}

u32 GetProcessorThreadCount() { return SrtRefMaxThreads; }

void WorkQueueCreate() {
    if (SrtRefQueueCreated) return;
    SrtRefThreads = (pthread_t *)calloc(SrtRefMaxThreads, sizeof(pthread_t));
    for (u32 i = 0; i < SrtRefMaxThreads; ++i) {
        pthread_create(&SrtRefThreads[i], 0, SrtRefWorker, (void *)(uintptr_t)i);
    }
    SrtRefQueueCreated = true;
}
void WorkQueueStart(thread_callback ThreadCallback, u32 WorkItemCount, u32 ThreadCount) {
    if (ThreadCount < 1) ThreadCount = 1;
    if (ThreadCount > SrtRefMaxThreads) ThreadCount = SrtRefMaxThreads;
    pthread_mutex_lock(&SrtRefLock);
    SrtRefCallback = ThreadCallback;
    SrtRefItemCount = WorkItemCount;
    SrtRefWorkIndex = 0;
    SrtRefWorkCompleted = 0;
    SrtRefActiveThreads = ThreadCount;
    SrtRefThreadsRunning = ThreadCount;
    SrtRefGeneration += 1;
    pthread_cond_broadcast(&SrtRefKick);
    pthread_mutex_unlock(&SrtRefLock);
}
void WorkQueueWaitUntilCompletion() {
    pthread_mutex_lock(&SrtRefLock);
    while (SrtRefThreadsRunning != 0) pthread_cond_wait(&SrtRefDone, &SrtRefLock);
    pthread_mutex_unlock(&SrtRefLock);
}
bool WorkQueueIsReady() { return SrtRefQueueCreated; }
bool WorkQueueHasCompleted() {
    pthread_mutex_lock(&SrtRefLock);
    bool Result = SrtRefThreadsRunning == 0;
    pthread_mutex_unlock(&SrtRefLock);
    return Result;
}
f64 WorkQueueGetCompletionTime() { return SrtRefCompletionTime; }

// ===================================================================================
// C entry points (ctypes-friendly).  All of them are test/baseline plumbing.
// ===================================================================================
extern "C" {
uint64_t srt_ref_thread_seed(uint32_t i);

static bool SrtRefInitDone = false;
static void *SrtRefPresentBuffer = 0;   // the platform-owned `image` OnRender copies into
static u64 SrtRefPresentCapacity = 0;

int srt_ref_simd_width(void) { return SIMD_WIDTH; }

// OnInit (main.cpp:645-679) once per process.  max_threads <= 0 -> online CPUs.
int srt_ref_init(int max_threads) {
    if (SrtRefInitDone) return (int)SrtRefMaxThreads;
    long Online = sysconf(_SC_NPROCESSORS_ONLN);
    if (max_threads <= 0) max_threads = Online > 0 ? (int)Online : 1;
    SrtRefMaxThreads = (u32)max_threads;
    init_params Params = {};
    OnInit(&Params);
    SrtRefInitDone = true;
    return (int)SrtRefMaxThreads;
}

// Renders `spp` frames (1 sample per pixel each, main.cpp:484) of built-in scene
// `scene` at w x h through OnRender (main.cpp:705-859), starting from a reset.
//   flags bit0: P10 per-(sample,pixel) RNG streams   bit1: P9 IEEE NormalizeFast
// Outputs (any may be NULL): rgba = CurrentImage (w*h u32, row 0 = image bottom),
// accum = PreviousImage (w*h*4 f32), tap_prim/tap_t = primary-ray primitive id
// (0xFFFFFFFF on miss) and MinT for sample `tap_sample`; rays = sum of
// RaysCastInThread over all frames; ms = sum over frames of (last-worker completion
// stamp - kick time), the reference's own frame-time definition (main.cpp:848-849).
int srt_ref_render(int scene, uint32_t w, uint32_t h, uint32_t spp, int enable_simd,
                   uint32_t threads, int flags, uint64_t stream_base,
                   uint32_t *rgba, float *accum, uint32_t *tap_prim, float *tap_t,
                   uint32_t tap_sample, uint64_t *rays, double *ms) {
    if (!SrtRefInitDone) srt_ref_init(0);
    if (scene < 0 || scene > 2 || w == 0 || h == 0 || spp == 0) return -1;
    u64 Pixels = (u64)w * h;
    if (Pixels * 20 > MB(512) - 4096) return -2;   // RenderData arena (main.cpp:658)
    if (SrtRefPresentCapacity < Pixels * 4) {
        free(SrtRefPresentBuffer);
        SrtRefPresentBuffer = malloc(Pixels * 4);
        SrtRefPresentCapacity = Pixels * 4;
    }
    SrtRefParityRng = (flags & 1) != 0;
    SrtRefIEEENormalizeFast = (flags & 2) != 0;
    SrtRefStreamBase = stream_base;
    SrtRefTapPrim = tap_prim;
    SrtRefTapT = tap_t;
    SrtRefTapSample = tap_sample;
    if (tap_prim) memset(tap_prim, 0xFF, Pixels * 4);
    if (tap_t) for (u64 i = 0; i < Pixels; ++i) tap_t[i] = F32Max;

    image Image = {};
    Image.Data = SrtRefPresentBuffer;
    Image.Width = w;
    Image.Height = h;
    Image.Format = format::R8G8B8A8_U32;
    render_params Params = {};
    Params.ThreadCount = threads ? threads : SrtRefMaxThreads;
    Params.EnableSIMD = enable_simd != 0;
    Params.SceneIndex = (u32)scene;

    // The as-shipped RNG is per worker thread and sequential; re-seed exactly as OnInit
    // does (main.cpp:667-678) so that every call is reproducible at 1 thread.
    for (u32 i = 0; i < SrtRefMaxThreads; ++i) {
        ThreadContexts[i].RandomState.Seed = srt_ref_thread_seed(i);   // the reference's own expression, zero shifts included
        ThreadContexts[i].RaysCastInThread = 0;
    }

    u64 TotalRays = 0;
    f64 TotalMs = 0.0;
    WorkQueueWaitUntilCompletion();
    for (u32 Frame = 0; Frame < spp; ++Frame) {
        u64 RaysOut = 0;
        f64 TimeOut = 0.0;
        SrtRefPressR = (Frame == 0);   // force the reset branch (main.cpp:791-804)
        f64 Kick = QueryTimestampInMilliseconds();
        OnRender(Image, Params, &RaysOut, &TimeOut);
        SrtRefPressR = false;
        WorkQueueWaitUntilCompletion();
        TotalMs += WorkQueueGetCompletionTime() - Kick;
        TotalRays += GetTotalRayCastCount();   // main.cpp:696-703 (zeroed by the next OnRender)
    }
    SrtRefTapPrim = 0;
    SrtRefTapT = 0;
    if (rgba) memcpy(rgba, CurrentImage.Data, Pixels * 4);
    if (accum) memcpy(accum, PreviousImage.Data, Pixels * 16);
    if (rays) *rays = TotalRays;
    if (ms) *ms = TotalMs;
    return 0;
}

// camera_info of the most recent frame (main.cpp:270-277):
// out[0..19] = CameraPosition, CameraZ, CameraX, CameraY, FilmCenter (4 floats each,
// lane 3 is the v3 padding), out[20] = FilmW, out[21] = FilmH.
void srt_ref_camera(float *out) {
    memcpy(out + 0, &CameraInfo.CameraPosition, 16);
    memcpy(out + 4, &CameraInfo.CameraZ, 16);
    memcpy(out + 8, &CameraInfo.CameraX, 16);
    memcpy(out + 12, &CameraInfo.CameraY, 16);
    memcpy(out + 16, &CameraInfo.FilmCenter, 16);
    out[20] = CameraInfo.FilmW;
    out[21] = CameraInfo.FilmH;
}

// Scene dump: scalar_sphere array (80 B each, main.cpp:17-21) + scene header fields.
// header = LookAt.xyz, DefaultDistanceFromLookAt, DefaultXAngle, DefaultYHeight.
int srt_ref_scene(int scene, void *spheres, uint32_t capacity, uint32_t *count,
                  int *use_sky, float *header) {
    if (!SrtRefInitDone) srt_ref_init(0);
    if (scene < 0 || scene > 2) return -1;
    const struct scene &S = Scenes[scene];
    if (count) *count = S.ScalarSpheres.Count;
    if (use_sky) *use_sky = S.UseSkyColor ? 1 : 0;
    if (header) {
        header[0] = S.LookAt.x; header[1] = S.LookAt.y; header[2] = S.LookAt.z;
        header[3] = S.DefaultDistanceFromLookAt;
        header[4] = S.DefaultXAngle;
        header[5] = S.DefaultYHeight;
    }
    if (spheres) {
        u32 N = S.ScalarSpheres.Count < capacity ? S.ScalarSpheres.Count : capacity;
        memcpy(spheres, S.ScalarSpheres.Data, (size_t)N * sizeof(scalar_sphere));
    }
    return (int)sizeof(scalar_sphere);
}

// Known-answer taps on the reference's own functions.
void srt_ref_rng(uint64_t seed, uint32_t n, float mn, float mx, int as_float, void *out) {
    u32_random_state S = { seed };
    for (u32 i = 0; i < n; ++i) {
        if (as_float) ((float *)out)[i] = S.RandomFloat(mn, mx);   // base.h:983-989
        else ((uint32_t *)out)[i] = S.PCG();                       // base.h:954-963
    }
}
uint64_t srt_ref_stream_seed(uint64_t i) { return SrtRefStreamSeed(i); }   // the P10 stream contract
uint64_t srt_ref_thread_seed(uint32_t i) {   // main.cpp:667-676 with the reference's u32 index
    u64 InitialSeed = 0x420247153476526ULL * (u64)i;
    InitialSeed += 0x8442885C91A5C8DULL;
    InitialSeed ^= InitialSeed >> ((7 + i) % 64);
    InitialSeed ^= InitialSeed << 23;
    InitialSeed ^= InitialSeed >> ((0x29 ^ i) % 64);
    InitialSeed = (InitialSeed * 0x11C19226CEB4769AULL) + 0x1105404122082911ULL;
    InitialSeed ^= InitialSeed << 19;
    InitialSeed ^= InitialSeed >> 13;
    return InitialSeed;
}
float srt_ref_linear_to_srgb(float l) { return LinearToSRGB(l); }            // main.cpp:312-329
uint32_t srt_ref_pack(const float *rgba) {                                   // main.cpp:331-346
    v4 V = v4(rgba[0], rgba[1], rgba[2], rgba[3]);
    return ColorFromV4(LinearToSRGB(V));
}
float srt_ref_reflectance(float c, float eta) { return Reflectance(c, eta); } // main.cpp:292-300
void srt_ref_normalize(const float *in, int fast, float *out) {              // x64_math.h:234-257
    v3 V = v3(in[0], in[1], in[2]);
    v3 R = fast ? v3::NormalizeFast(V) : v3::Normalize(V);
    out[0] = R.x; out[1] = R.y; out[2] = R.z;
}
void srt_ref_set_ieee_normalize_fast(int on) { SrtRefIEEENormalizeFast = on != 0; }

// Replace built-in scene `slot` (0..2) by a caller-supplied sphere table so that the
// reference's own RenderTile / RenderTileScalar can be timed and checked on the synthetic
// stress scenes (BASELINE.json configs 3 and 5).  Builds the SoA exactly like the reference
// does (ConvertScalarSpheresToSIMDSpheres, main.cpp:73-91); padding lanes are zero spheres.
// header = LookAt.xyz, DefaultDistanceFromLookAt, DefaultXAngle, DefaultYHeight.
int srt_ref_set_scene(int slot, const void *spheres, uint32_t n, int use_sky, const float *header) {
    if (!SrtRefInitDone) srt_ref_init(0);
    if (slot < 0 || slot > 2 || !spheres || n == 0 || !header) return -1;
    WorkQueueWaitUntilCompletion();
    scene &S = Scenes[slot];
    u32 Groups = (n + (SIMD_WIDTH - 1)) / SIMD_WIDTH;
    void *A = 0, *B = 0, *C = 0;
    if (posix_memalign(&A, 64, (size_t)n * sizeof(scalar_sphere))) return -2;
    if (posix_memalign(&B, 64, (size_t)Groups * sizeof(sphere_group))) return -2;
    if (posix_memalign(&C, 64, (size_t)(n + 1) * sizeof(material))) return -2;
    memcpy(A, spheres, (size_t)n * sizeof(scalar_sphere));
    memset(B, 0, (size_t)Groups * sizeof(sphere_group));
    memset(C, 0, (size_t)(n + 1) * sizeof(material));
    S.ScalarSpheres.Data = (scalar_sphere *)A; S.ScalarSpheres.Count = n;
    S.SIMDSpheres.Data = (sphere_group *)B; S.SIMDSpheres.Count = Groups;
    S.Materials.Data = (material *)C; S.Materials.Count = n + 1;
    S.UseSkyColor = use_sky != 0;
    S.LookAt = v3(header[0], header[1], header[2]);
    S.DefaultDistanceFromLookAt = header[3];
    S.DefaultXAngle = header[4];
    S.DefaultYHeight = header[5];
    ConvertScalarSpheresToSIMDSpheres(&S);
    SceneIndex = (u32)-1;      // force the scene-switch branch (main.cpp:718-727) on the next frame
    return 0;
}

} // extern "C"
