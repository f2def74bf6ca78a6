// srt_app.cpp -- OnInit / OnRender on top of the CUDA backend.
//
// Same frame protocol as the reference's OnRender (main.cpp:705-859):
//   * non-blocking: a call publishes the previous frame (if finished) and kicks the next;
//     it returns false while the previous frame is still rendering;
//   * scene switch, resize, camera movement or key R reset the accumulation
//     (main.cpp:718-727, 791-804); otherwise PreviousRayCount advances by one frame;
//   * OutTotalRaysCast / OutTimeElapsed describe the frame that was just published.
// The kick itself -- WorkQueueStart(RenderTile | RenderTileScalar, ...) at main.cpp:852 --
// is srt_render(); completion polling maps to srt_poll / srt_wait.
#include "srt_app.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <time.h>
#include <vector>

namespace {

constexpr f32 kWorldScale = 1.0f / 16.0f;
constexpr f32 kPi = 3.14159265358979323846f;

backend_params g_Backend;
srt_backend *g_Handle = nullptr;
scene g_Scenes[SceneCount];
bool g_SceneBuilt[SceneCount];
bool g_Keys[(u32)key::Count];
std::string g_Error;

u32 g_SceneIndex = 0xFFFFFFFFu;      // static u32 SceneIndex = -1  (main.cpp:53)
bool g_SceneUploadedSimd = false;
u32 g_PreviousRayCount = 0;          // samples already in the accumulator (main.cpp:7)
u32 g_Width = 0, g_Height = 0;       // PreviousImage / CurrentImage size
bool g_FrameInFlight = false;
u32 g_InFlightSamples = 0;
bool g_ResumePending = false;        // ResumeFromCheckpoint armed: the next OnRender uploads g_ResumeAccum
checkpoint_header g_ResumeHeader;
std::vector<f32> g_ResumeAccum;
f32 g_Distance, g_XAngle, g_YHeight; // orbit state (function statics at main.cpp:712-714)

struct vec { f32 x, y, z; };
vec Sub(vec a, vec b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
vec Cross(vec a, vec b) { return { a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x }; }
vec Normalize(vec a) {                                                   // x64_math.h:234-245
    const f32 L2 = (a.x * a.x + a.y * a.y) + a.z * a.z;
    if (!(L2 > 1e-4f)) return { 0, 0, 0 };
    const f32 L = sqrtf(L2);
    return { a.x / L, a.y / L, a.z / L };
}

bool Fail(const char *What) {
    g_Error = What;
    if (g_Handle) { g_Error += ": "; g_Error += srt_last_error(g_Handle); }
    return false;
}

} // namespace

f64 QueryTimestampInMilliseconds() {
    timespec Ts;
    clock_gettime(CLOCK_MONOTONIC, &Ts);
    return (f64)Ts.tv_sec * 1e3 + (f64)Ts.tv_nsec * 1e-6;
}

void SetKeyDown(key Key, bool Down) { g_Keys[(u32)Key] = Down; }
bool IsDown(key Key) { return g_Keys[(u32)Key]; }
void SetBackendParams(const backend_params &Params) { g_Backend = Params; }
const char *LastError() { return g_Error.c_str(); }
srt_backend *GetBackend() { return g_Handle; }
u32 GetAccumulatedSampleCount() { return g_PreviousRayCount + g_InFlightSamples; }   // valid after WaitForFrame()

const scene *GetScene(u32 SceneIndex) {
    if (SceneIndex >= SceneCount) return nullptr;
    if (!g_SceneBuilt[SceneIndex]) {      // the two synthetic scenes are built on first use
        BuildScene(SceneIndex, &g_Scenes[SceneIndex]);
        g_SceneBuilt[SceneIndex] = true;
    }
    return &g_Scenes[SceneIndex];
}

// main.cpp:776-780 (orbit position) and 811-822 (basis + film).
void ComputeCamera(const scene &Scene, f32 Distance, f32 XAngle, f32 YHeight, u32 Width, u32 Height, srt_camera *Out) {
    memset(Out, 0, sizeof(*Out));
    const vec LookAt = { Scene.LookAt.x, Scene.LookAt.y, Scene.LookAt.z };
    const f32 Px = cosf(XAngle) * Distance, Pz = sinf(XAngle) * Distance;
    const vec Position = { Px + LookAt.x, YHeight + LookAt.y, Pz + LookAt.z };
    const vec Z = Normalize(Sub(Position, LookAt));
    const vec X = Normalize(Cross({ 0.0f, 1.0f, 0.0f }, Z));
    const vec Y = Normalize(Cross(Z, X));
    const vec FilmCenter = Sub(Position, Z);
    f32 FilmW = 1.0f, FilmH = 1.0f;
    if (Width > Height) FilmH = (f32)Height / (f32)Width;
    else FilmW = (f32)Width / (f32)Height;
    const vec *Src[5] = { &Position, &Z, &X, &Y, &FilmCenter };
    f32 *Dst[5] = { Out->camera_position, Out->camera_z, Out->camera_x, Out->camera_y, Out->film_center };
    for (int i = 0; i < 5; ++i) { Dst[i][0] = Src[i]->x; Dst[i][1] = Src[i]->y; Dst[i][2] = Src[i]->z; }
    Out->film_w = FilmW;
    Out->film_h = FilmH;
}

bool OnInit(init_params *Params) {
    if (Params) {                         // main.cpp:646-650
        Params->WindowTitle = "SIMD Ray Tracer";
        Params->WindowWidth = 1280;
        Params->WindowHeight = 720;
    }
    for (u32 i = 0; i < BuiltinSceneCount; ++i) GetScene(i);    // main.cpp:652-654
    if (!g_Handle) {
        if (srt_create(&g_Handle, g_Backend.DeviceId) != SRT_OK) {
            g_Error = std::string("srt_create: ") + srt_last_error(g_Handle);
            if (g_Handle) { srt_destroy(g_Handle); g_Handle = nullptr; }
            return false;
        }
    }
    g_SceneIndex = 0xFFFFFFFFu;
    g_PreviousRayCount = 0;
    g_FrameInFlight = false;
    g_InFlightSamples = 0;
    g_Width = g_Height = 0;
    return true;
}

void OnShutdown() {
    if (g_Handle) { srt_destroy(g_Handle); g_Handle = nullptr; }
}

void WaitForFrame() {
    if (g_Handle && g_FrameInFlight) srt_wait(g_Handle, nullptr, nullptr);
}

// Replace a scene slot by a caller-provided scene (copied).  A slot that is currently
// uploaded is re-uploaded by the next OnRender (which resets the accumulation, as a scene
// switch does at main.cpp:718-727).
bool SetScene(u32 SceneIndex, const scene &Scene) {
    if (SceneIndex >= SceneCount || !Scene.ScalarSpheres || Scene.SphereCount == 0) return false;
    scalar_sphere *Copy = (scalar_sphere *)malloc(sizeof(scalar_sphere) * Scene.SphereCount);
    if (!Copy) return false;
    memcpy(Copy, Scene.ScalarSpheres, sizeof(scalar_sphere) * Scene.SphereCount);
    if (g_SceneBuilt[SceneIndex]) free(g_Scenes[SceneIndex].ScalarSpheres);
    g_Scenes[SceneIndex] = Scene;
    g_Scenes[SceneIndex].ScalarSpheres = Copy;
    g_SceneBuilt[SceneIndex] = true;
    if (g_SceneIndex == SceneIndex) g_SceneIndex = 0xFFFFFFFFu;
    return true;
}

// Drop a replaced slot: the next GetScene(SceneIndex) rebuilds the built-in / synthetic scene.
void ResetScene(u32 SceneIndex) {
    if (SceneIndex >= SceneCount || !g_SceneBuilt[SceneIndex]) return;
    free(g_Scenes[SceneIndex].ScalarSpheres);
    g_Scenes[SceneIndex].ScalarSpheres = nullptr;
    g_SceneBuilt[SceneIndex] = false;
    if (g_SceneIndex == SceneIndex) g_SceneIndex = 0xFFFFFFFFu;
}

// Arm a resume: the next OnRender that would start a fresh accumulation (first call, scene
// switch or resize) with the checkpoint's size and scene uploads the accumulator instead and
// continues at sample index Header.Samples.  The sample streams are pure functions of
// (stream_base, sample, pixel), so the continued image is bit-identical to an
// uninterrupted one.
bool ResumeFromCheckpoint(const checkpoint_header &Header, const f32 *Accum) {
    if (!Accum || Header.Width == 0 || Header.Height == 0 || Header.SceneIndex >= SceneCount) return false;
    g_ResumeHeader = Header;
    g_ResumeAccum.assign(Accum, Accum + (size_t)Header.Width * Header.Height * 4);
    g_ResumePending = true;
    return true;
}

bool ReadAccumulator(f32 *Dst) { return g_Handle && srt_read_accum(g_Handle, Dst) == SRT_OK; }

bool OnRender(const image &Image, render_params RenderParams, u64 *OutTotalRaysCast, f64 *OutTimeElapsed) {
    if (!g_Handle) return Fail("OnRender: OnInit has not succeeded");          // WorkQueueIsReady, main.cpp:707
    if (RenderParams.SceneIndex >= SceneCount) return Fail("OnRender: bad SceneIndex");

    const bool Complete = !g_FrameInFlight || srt_poll(g_Handle) == 1;        // WorkQueueHasCompleted
    bool Moved = false;
    const bool SceneSwitch = g_SceneIndex != RenderParams.SceneIndex;
    const bool ModeSwitch = !SceneSwitch && g_SceneUploadedSimd != RenderParams.EnableSIMD;
    if (SceneSwitch || ModeSwitch) {                                           // main.cpp:718-727
        if (!Complete) return false;
        const scene *S = GetScene(RenderParams.SceneIndex);
        srt_scene_options Opt = {};
        Opt.use_sky_color = S->UseSkyColor ? 1 : 0;
        Opt.lane_width = RenderParams.EnableSIMD ? (int32_t)g_Backend.LaneWidth : 1;   // main.cpp:852-856
        Opt.fma_dot = (RenderParams.EnableSIMD && g_Backend.FmaDot) ? 1 : 0;
        if (srt_set_scene(g_Handle, (const srt_sphere *)S->ScalarSpheres, S->SphereCount, &Opt) != SRT_OK)
            return Fail("srt_set_scene");
        g_SceneUploadedSimd = RenderParams.EnableSIMD;
        if (SceneSwitch) {
            g_SceneIndex = RenderParams.SceneIndex;
            g_Distance = S->DefaultDistanceFromLookAt;
            g_XAngle = S->DefaultXAngle;
            g_YHeight = S->DefaultYHeight;
        }
        Moved = true;
    }
    const scene &Scene = *GetScene(g_SceneIndex);

    {   // orbit controls, main.cpp:730-774
        const f32 Speed = 1.0f * kWorldScale;
        if (IsDown(key::W) || IsDown(key::ArrowUp)) { g_Distance -= Speed; Moved = true; }
        if (IsDown(key::S) || IsDown(key::ArrowDown)) { g_Distance += Speed; Moved = true; }
        if (IsDown(key::D) || IsDown(key::ArrowRight)) { g_XAngle -= 1.0f / 16.0f; Moved = true; }
        if (IsDown(key::A) || IsDown(key::ArrowLeft)) { g_XAngle += 1.0f / 16.0f; Moved = true; }
        if (IsDown(key::Space)) { g_YHeight += Speed; Moved = true; }
        if (IsDown(key::C) || IsDown(key::LeftControl)) { g_YHeight -= Speed; Moved = true; }
        if (g_XAngle > kPi * 2.0f) g_XAngle -= kPi * 4.0f;
        if (g_XAngle < -(kPi * 2.0f)) g_XAngle += kPi * 4.0f;
        if (g_Distance < 0.5f * kWorldScale) g_Distance = 0.5f * kWorldScale;
        if (g_YHeight > g_Distance) g_YHeight = g_Distance;
    }

    const bool Resize = Image.Width != g_Width || Image.Height != g_Height;
    bool CopyToOutput = Complete && !Resize && g_Width != 0;

    // publish the finished frame (CopyImage, main.cpp:787-789) and its statistics (840-849)
    auto Publish = [&]() -> bool {
        f64 Ms = 0.0;
        u64 Rays = 0;
        if (srt_wait(g_Handle, &Ms, &Rays) != SRT_OK) return Fail("srt_wait");
        if (Image.Data && srt_read_rgba8(g_Handle, Image.Data) != SRT_OK) return Fail("srt_read_rgba8");
        if (OutTotalRaysCast) *OutTotalRaysCast = Rays;
        if (OutTimeElapsed) *OutTimeElapsed = Ms;
        g_FrameInFlight = false;
        return true;
    };

    if (Resize || Moved || IsDown(key::R)) {                                   // main.cpp:791-804
        if (g_FrameInFlight || (!Resize && g_Width != 0)) {
            if (!Resize && g_Width != 0) { if (!Publish()) return false; CopyToOutput = true; }
            else { srt_wait(g_Handle, nullptr, nullptr); g_FrameInFlight = false; }
        }
        g_PreviousRayCount = 0;
        g_InFlightSamples = 0;
        if (Resize) {
            if (srt_resize(g_Handle, Image.Width, Image.Height) != SRT_OK) return Fail("srt_resize");
            g_Width = Image.Width;
            g_Height = Image.Height;
        } else if (srt_reset_accumulation(g_Handle) != SRT_OK) return Fail("srt_reset_accumulation");
        if (g_ResumePending) {
            g_ResumePending = false;
            if (g_ResumeHeader.Width != g_Width || g_ResumeHeader.Height != g_Height || g_ResumeHeader.SceneIndex != g_SceneIndex ||
                g_ResumeHeader.StreamBase != g_Backend.StreamBase)
                return Fail("OnRender: checkpoint does not match the image size, scene or stream base");
            if (srt_write_accum(g_Handle, g_ResumeAccum.data()) != SRT_OK) return Fail("srt_write_accum");
            g_PreviousRayCount = g_ResumeHeader.Samples;
            g_ResumeAccum.clear();
            g_ResumeAccum.shrink_to_fit();
        }
    } else if (Complete) {
        if (CopyToOutput && !Publish()) return false;
        g_PreviousRayCount += g_InFlightSamples;                                // PreviousRayCount += 1, main.cpp:806
        g_InFlightSamples = 0;
    } else {
        return false;                                                          // still rendering, main.cpp:807-809
    }

    srt_camera Camera;
    ComputeCamera(Scene, g_Distance, g_XAngle, g_YHeight, Image.Width, Image.Height, &Camera);

    srt_render_args Args = {};
    Args.first_sample = g_PreviousRayCount;
    Args.n_samples = g_Backend.SamplesPerFrame ? g_Backend.SamplesPerFrame : 1;
    Args.sample_stride = 1;
    Args.accum_mode = SRT_ACCUM_MEAN;
    Args.stream_base = g_Backend.StreamBase;
    if (srt_render(g_Handle, &Camera, &Args) != SRT_OK) return Fail("srt_render");   // the kick, main.cpp:852
    g_FrameInFlight = true;
    g_InFlightSamples = Args.n_samples;
    if (g_Backend.Blocking) srt_wait(g_Handle, nullptr, nullptr);
    return CopyToOutput;
}
