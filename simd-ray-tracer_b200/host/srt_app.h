// srt_app.h -- host-side mirror of the reference's application interface for the hot path.
//
// Keeps the names and meaning of the reference's types and entry points so a user of
// SIMD-Ray-Tracer finds the same surface (file:line in the reference):
//   material / scalar_sphere / scene / camera_info     main.cpp:11-51, 270-282
//   image / format / init_params / render_params        base.h:117-161
//   OnInit / OnRender                                    base.h:163-164, main.cpp:645-859
// What changed: the work-queue kick (main.cpp:852) launches the CUDA backend through the
// C ABI in include/srt_backend.h; there is no window, no GL and no keyboard -- camera keys
// are injected with SetKeyDown() so scripted orbits exercise the same code path.
#pragma once
#include <stdint.h>
#include "srt_backend.h"

typedef uint8_t u8;
typedef uint32_t u32;
typedef int32_t s32;
typedef uint64_t u64;
typedef float f32;
typedef double f64;

struct v3 { f32 x, y, z, _w; };                 // 16 bytes, base.h:357-375

// Binary-identical to srt_material / srt_sphere (and to the reference's structs).
struct material { v3 Color; v3 Emissive; f32 Specular; f32 IndexOfRefraction; f32 _pad[2]; };
struct scalar_sphere { v3 Position; f32 Radius; f32 _pad[3]; material Material; };
static_assert(sizeof(material) == sizeof(srt_material), "material layout");
static_assert(sizeof(scalar_sphere) == sizeof(srt_sphere), "scalar_sphere layout");

struct scene {
    v3 LookAt;
    bool UseSkyColor;
    f32 DefaultDistanceFromLookAt;
    f32 DefaultXAngle;
    f32 DefaultYHeight;
    scalar_sphere *ScalarSpheres;
    u32 SphereCount;
    const char *Name;
};

enum class format : u32 { R32B32G32A32_F32 = 0x1, R8G8B8A8_U32 };
struct image { void *Data; u32 Width, Height; format Format; };

struct init_params { u32 WindowWidth, WindowHeight; const char *WindowTitle; };

// render_params base.h:157-161.  ThreadCount = workers used for the frame (WorkQueueStart(.., ThreadCount),
// main.cpp:851-856): here the workers are the GPUs OnInit created; 0 or too many = all of them.
// EnableSIMD picks RenderTile (lanes) vs RenderTileScalar semantics as in main.cpp:852-856.
struct render_params { u32 ThreadCount; bool EnableSIMD; u32 SceneIndex; };

// Backend knobs that have no counterpart in the reference (set before OnInit).
constexpr u32 MaxDeviceCount = 16;
struct backend_params {
    int DeviceId = 0;         // first CUDA ordinal
    u32 DeviceCount = 1;      // GPUs DeviceId .. DeviceId + DeviceCount - 1 become the frame's workers (WorkQueueCreate, main.cpp:660)
    u32 LaneWidth = 8;        // SIMD_WIDTH the lanes emulate (base.h:53-64): 8 (AVX2) or 4 (SSE/wasm)
    bool FmaDot = true;       // x64_math.h:409-420 (AVX2 profile) vs 299-309
    u32 SamplesPerFrame = 1;  // the reference renders exactly 1 spp per frame (main.cpp:484)
    u64 StreamBase = 0;
    bool Blocking = false;    // true: OnRender waits for its own frame (deterministic drivers)
};

enum class key : u32 { None = 0, W, S, A, D, Space, C, LeftControl, R, ArrowUp, ArrowDown, ArrowLeft, ArrowRight, Count };
void SetKeyDown(key Key, bool Down);
bool IsDown(key Key);

constexpr u32 BuiltinSceneCount = 3;
constexpr u32 SyntheticScene1K = 3;    // BASELINE.json config 3: 1,024 random spheres + ground
constexpr u32 SyntheticScene64K = 4;   // BASELINE.json config 5: 65,536 spheres + ground
constexpr u32 SceneCount = 5;

void SetBackendParams(const backend_params &Params);
bool OnInit(init_params *Params);     // false: backend could not be created (see LastError)
bool OnRender(const image &Image, render_params RenderParams, u64 *OutTotalRaysCast, f64 *OutTimeElapsed);
void OnShutdown();
const char *LastError();

// Introspection used by the headless driver and the tests.
const scene *GetScene(u32 SceneIndex);
u32 GetAccumulatedSampleCount();                       // PreviousRayCount + completed frame
void WaitForFrame();                                   // WorkQueueWaitUntilCompletion
// Publish the frame in flight without kicking another one (end of a batch render): waits, gathers the workers'
// rows, copies the RGBA8 image to Image.Data (if any) and returns the frame's statistics.
bool PublishFrame(const image &Image, u64 *OutTotalRaysCast, f64 *OutTimeElapsed);
u32 GetDeviceCount();
bool ReadAccumulator(f32 *Dst);                        // PreviousImage, W*H*4 floats
srt_backend *GetBackend();
void ComputeCamera(const scene &Scene, f32 Distance, f32 XAngle, f32 YHeight, u32 Width, u32 Height, srt_camera *Out);

// Files either side of the path (srt_files.cpp): accumulator checkpoints and scene descriptions.
#include <vector>
struct checkpoint_header { u32 Width, Height, Samples, SceneIndex; u64 StreamBase; u32 LaneWidth, FmaDot; u64 SceneHash; };   // SceneHash 0: unknown
u64 HashScene(const scene &Scene);                      // FNV-1a of the sphere table + sky flag
u32 GetSceneGeneration(u32 SceneIndex);                 // bumped by SetScene / ResetScene
bool SaveCheckpoint(const char *Path, const checkpoint_header &Header, const f32 *Accum);
bool LoadCheckpoint(const char *Path, checkpoint_header *Header, std::vector<f32> *Accum);
bool SaveSceneFile(const char *Path, const scene &Scene);
// Presented image (RGBA8, row 0 = image bottom) and linear accumulator to files; rows are flipped to top-down on the way out.
bool WritePpm(const char *Path, const u32 *Pixels, u32 W, u32 H);
bool WritePng(const char *Path, const u32 *Pixels, u32 W, u32 H);
bool WriteImage(const char *Path, const u32 *Pixels, u32 W, u32 H);   // by extension: .png -> PNG, anything else PPM
bool WritePfm(const char *Path, const f32 *Accum, u32 W, u32 H);
bool LoadSceneFile(const char *Path, scene *Out);       // allocates Out->ScalarSpheres
// Replace scene slot `SceneIndex` (0..SceneCount-1) by a loaded scene; the next OnRender with
// that index uploads it.  Resume: continue an accumulation from a checkpoint (the next
// OnRender must use the checkpoint's size and scene).
bool SetScene(u32 SceneIndex, const scene &Scene);
void ResetScene(u32 SceneIndex);                        // back to the built-in scene of that slot
bool ResumeFromCheckpoint(const checkpoint_header &Header, const f32 *Accum);

// Scene construction (srt_scenes.cpp).
void BuildScene(u32 SceneIndex, scene *Out);           // allocates Out->ScalarSpheres
f64 QueryTimestampInMilliseconds();
