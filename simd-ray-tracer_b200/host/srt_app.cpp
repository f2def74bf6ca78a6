// srt_app.cpp -- OnInit / OnRender on top of the CUDA backend.
//
// Same frame protocol as the reference's OnRender (main.cpp:705-859):
//   * non-blocking: a call publishes the previous frame (if finished) and kicks the next;
//     it returns false while the previous frame is still rendering;
//   * scene switch, resize, camera movement or key R reset the accumulation
//     (main.cpp:718-727, 791-804); otherwise PreviousRayCount advances by one frame;
//   * OutTotalRaysCast / OutTimeElapsed describe the frame that was just published.
// The kick itself -- WorkQueueStart(RenderTile | RenderTileScalar, TileCount, ThreadCount) at
// main.cpp:851-856 -- is one srt_render() per worker; the workers are GPUs
// (backend_params.DeviceCount handles, created at OnInit like WorkQueueCreate at main.cpp:660).
// Each worker owns every DeviceCount-th row of 8x4-pixel blocks of the frame (the reference's
// workers pull 32x32 tiles, main.cpp:9, 362-368) and accumulates it with the exact running mean;
// publishing a frame gathers the rows on worker 0 with one NCCL reduce issued by the library
// (srt_gather_rgba).  The published image is bit-identical for every worker count.
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
srt_backend *g_Handles[MaxDeviceCount];
u32 g_DeviceCount = 0;               // handles created by OnInit
u32 g_Workers = 0;                   // handles the current accumulation is partitioned over
bool g_Keys[(u32)key::Count];
std::string g_Error;

u32 g_SceneIndex = 0xFFFFFFFFu;      // static u32 SceneIndex = -1  (main.cpp:53)
u32 g_SceneGeneration = 0;           // GetSceneGeneration() of the uploaded slot: SetScene / ResetScene force a re-upload
bool g_SceneUploadedSimd = false;
u32 g_PreviousRayCount = 0;          // samples already in the accumulator (main.cpp:7)
u32 g_Width = 0, g_Height = 0;       // PreviousImage / CurrentImage size
bool g_FrameInFlight = false;
u32 g_InFlightSamples = 0;
bool g_ResumePending = false;        // ResumeFromCheckpoint armed: the next OnRender uploads g_ResumeAccum
checkpoint_header g_ResumeHeader;
std::vector<f32> g_ResumeAccum;
f32 g_Distance, g_XAngle, g_YHeight; // orbit state (function statics at main.cpp:712-714)

bool Fail(const char *What, srt_backend *Handle = nullptr) {
    g_Error = What;
    if (!Handle && g_DeviceCount) Handle = g_Handles[0];
    if (Handle) { g_Error += ": "; g_Error += srt_last_error(Handle); }
    return false;
}

void DestroyHandles() {
    for (u32 i = 0; i < g_DeviceCount; ++i) if (g_Handles[i]) { srt_destroy(g_Handles[i]); g_Handles[i] = nullptr; }
    g_DeviceCount = 0;
    g_Workers = 0;
}

// Wait for every worker; device time of the frame = the slowest worker, rays = the sum (main.cpp:840-849).
bool WaitAll(f64 *OutMs, u64 *OutRays) {
    f64 MaxMs = 0.0;
    u64 Rays = 0;
    for (u32 i = 0; i < g_Workers; ++i) {
        f64 Ms = 0.0;
        u64 R = 0;
        if (srt_wait(g_Handles[i], &Ms, &R) != SRT_OK) return Fail("srt_wait", g_Handles[i]);
        if (Ms > MaxMs) MaxMs = Ms;
        Rays += R;
    }
    if (OutMs) *OutMs = MaxMs;
    if (OutRays) *OutRays = Rays;
    return true;
}

// One host thread drives every worker, so the collective calls of a step are issued as one NCCL group.
bool GatherOnWorker0(bool Accum) {
    if (g_Workers <= 1) return true;
    if (srt_comm_group_begin() != SRT_OK) return Fail("srt_comm_group_begin");
    bool Ok = true;
    for (u32 i = 0; i < g_Workers; ++i) {
        const int Rc = Accum ? srt_gather_accum(g_Handles[i], 0) : srt_gather_rgba(g_Handles[i], 0);
        if (Rc != SRT_OK) { Ok = false; Fail(Accum ? "srt_gather_accum" : "srt_gather_rgba", g_Handles[i]); break; }
    }
    if (srt_comm_group_end() != SRT_OK && Ok) return Fail("srt_comm_group_end");
    return Ok;
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
srt_backend *GetBackend() { return g_DeviceCount ? g_Handles[0] : nullptr; }
u32 GetDeviceCount() { return g_DeviceCount; }
u32 GetAccumulatedSampleCount() { return g_PreviousRayCount + g_InFlightSamples; }   // valid after WaitForFrame()

bool OnInit(init_params *Params) {
    if (Params) {                         // main.cpp:646-650
        Params->WindowTitle = "SIMD Ray Tracer";
        Params->WindowWidth = 1280;
        Params->WindowHeight = 720;
    }
    g_Error.clear();
    for (u32 i = 0; i < BuiltinSceneCount; ++i) GetScene(i);    // main.cpp:652-654
    if (g_DeviceCount == 0) {
        // the workers: WorkQueueCreate(ThreadCount) at main.cpp:660 -> one backend handle per GPU
        const u32 Count = g_Backend.DeviceCount ? g_Backend.DeviceCount : 1u;
        if (Count > MaxDeviceCount) { g_Error = "OnInit: DeviceCount exceeds MaxDeviceCount"; return false; }
        for (u32 i = 0; i < Count; ++i) {
            if (srt_create(&g_Handles[i], g_Backend.DeviceId + (int)i) != SRT_OK) {
                g_Error = std::string("srt_create: ") + srt_last_error(nullptr);
                DestroyHandles();
                return false;
            }
            g_DeviceCount = i + 1;
        }
        if (Count > 1) {
            unsigned char Id[SRT_COMM_ID_BYTES];
            bool Ok = srt_comm_unique_id(Id) == SRT_OK && srt_comm_group_begin() == SRT_OK;
            if (!Ok) g_Error = std::string("OnInit: NCCL: ") + srt_last_error(nullptr);
            for (u32 i = 0; Ok && i < Count; ++i)
                if (srt_comm_init_rank(g_Handles[i], (int)Count, (int)i, Id) != SRT_OK) { Ok = false; Fail("srt_comm_init_rank", g_Handles[i]); }
            if (srt_comm_group_end() != SRT_OK && Ok) { Ok = false; g_Error = std::string("OnInit: NCCL: ") + srt_last_error(nullptr); }
            if (!Ok) { DestroyHandles(); return false; }
        }
    }
    g_SceneIndex = 0xFFFFFFFFu;
    g_PreviousRayCount = 0;
    g_FrameInFlight = false;
    g_InFlightSamples = 0;
    g_Width = g_Height = 0;
    g_Workers = 0;
    return true;
}

void OnShutdown() { DestroyHandles(); }

void WaitForFrame() {
    if (g_FrameInFlight) for (u32 i = 0; i < g_Workers; ++i) srt_wait(g_Handles[i], nullptr, nullptr);
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

bool PublishFrame(const image &Image, u64 *OutTotalRaysCast, f64 *OutTimeElapsed) {
    if (!g_DeviceCount || !g_Workers) return Fail("PublishFrame: nothing rendered");
    if (!WaitAll(OutTimeElapsed, OutTotalRaysCast)) return false;
    g_FrameInFlight = false;
    if (Image.Data) {
        if (Image.Width != g_Width || Image.Height != g_Height) return Fail("PublishFrame: image size differs from the accumulation");
        if (!GatherOnWorker0(false)) return false;
        if (srt_read_rgba8(g_Handles[0], Image.Data) != SRT_OK) return Fail("srt_read_rgba8");
    }
    return true;
}

bool ReadAccumulator(f32 *Dst) {
    if (!g_DeviceCount || !g_Workers) return false;
    WaitForFrame();
    if (!GatherOnWorker0(true)) return false;
    return srt_read_accum(g_Handles[0], Dst) == SRT_OK;
}

bool OnRender(const image &Image, render_params RenderParams, u64 *OutTotalRaysCast, f64 *OutTimeElapsed) {
    if (!g_DeviceCount) return Fail("OnRender: OnInit has not succeeded");     // WorkQueueIsReady, main.cpp:707
    if (RenderParams.SceneIndex >= SceneCount) return Fail("OnRender: bad SceneIndex");

    bool Complete = true;                                                       // WorkQueueHasCompleted
    if (g_FrameInFlight) for (u32 i = 0; i < g_Workers; ++i) Complete = Complete && srt_poll(g_Handles[i]) == 1;
    bool Moved = false;
    // render_params.ThreadCount (base.h:157-161) = workers used for the frame, as in WorkQueueStart(.., ThreadCount)
    // at main.cpp:851-856; 0 or more than OnInit created: all of them.  Changing it re-partitions the image.
    const u32 Workers = (RenderParams.ThreadCount == 0 || RenderParams.ThreadCount > g_DeviceCount) ? g_DeviceCount : RenderParams.ThreadCount;
    const bool Repartition = Workers != g_Workers;
    const bool SceneSwitch = g_SceneIndex != RenderParams.SceneIndex || g_SceneGeneration != GetSceneGeneration(RenderParams.SceneIndex);
    // EnableSIMD toggled: the reference only swaps the tile callback (main.cpp:852-856) and keeps accumulating; here the
    // two callbacks are different arithmetic profiles of the scene upload, and one accumulation never mixes profiles
    // (that is what keeps a checkpoint resumable bit for bit), so the toggle restarts the accumulation like a scene switch.
    const bool ModeSwitch = !SceneSwitch && g_SceneUploadedSimd != RenderParams.EnableSIMD;
    if (SceneSwitch || ModeSwitch) {                                           // main.cpp:718-727
        if (!Complete) return false;
        const scene *S = GetScene(RenderParams.SceneIndex);
        srt_scene_options Opt = {};
        Opt.use_sky_color = S->UseSkyColor ? 1 : 0;
        Opt.lane_width = RenderParams.EnableSIMD ? (int32_t)g_Backend.LaneWidth : 1;   // main.cpp:852-856
        Opt.fma_dot = (RenderParams.EnableSIMD && g_Backend.FmaDot) ? 1 : 0;
        for (u32 i = 0; i < g_DeviceCount; ++i)
            if (srt_set_scene(g_Handles[i], (const srt_sphere *)S->ScalarSpheres, S->SphereCount, &Opt) != SRT_OK)
                return Fail("srt_set_scene", g_Handles[i]);
        g_SceneUploadedSimd = RenderParams.EnableSIMD;
        if (SceneSwitch) {
            g_SceneIndex = RenderParams.SceneIndex;
            g_SceneGeneration = GetSceneGeneration(g_SceneIndex);
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
        if (!WaitAll(&Ms, &Rays)) return false;
        if (Image.Data) {
            if (!GatherOnWorker0(false)) return false;
            if (srt_read_rgba8(g_Handles[0], Image.Data) != SRT_OK) return Fail("srt_read_rgba8");
        }
        if (OutTotalRaysCast) *OutTotalRaysCast = Rays;
        if (OutTimeElapsed) *OutTimeElapsed = Ms;
        g_FrameInFlight = false;
        return true;
    };

    if (Resize || Moved || Repartition || IsDown(key::R)) {                    // main.cpp:791-804
        if (g_FrameInFlight || (!Resize && g_Width != 0)) {
            if (!Resize && g_Width != 0 && g_Workers != 0) { if (!Publish()) return false; CopyToOutput = true; }
            else { WaitForFrame(); g_FrameInFlight = false; }
        }
        g_PreviousRayCount = 0;
        g_InFlightSamples = 0;
        g_Workers = Workers;
        for (u32 i = 0; i < g_DeviceCount; ++i) {
            srt_backend *Handle = g_Handles[i];
            // worker i renders block rows i, i + Workers, ...; a handle that sits this accumulation out keeps the default
            if (srt_set_partition(Handle, i < Workers ? i : 0u, i < Workers ? Workers : 1u) != SRT_OK) return Fail("srt_set_partition", Handle);
            if (Resize) { if (srt_resize(Handle, Image.Width, Image.Height) != SRT_OK) return Fail("srt_resize", Handle); }
            else if (srt_reset_accumulation(Handle) != SRT_OK) return Fail("srt_reset_accumulation", Handle);
        }
        if (Resize) { g_Width = Image.Width; g_Height = Image.Height; }
        if (g_ResumePending) {
            g_ResumePending = false;
            const u32 Lanes = RenderParams.EnableSIMD ? g_Backend.LaneWidth : 1u, Fma = (RenderParams.EnableSIMD && g_Backend.FmaDot) ? 1u : 0u;
            if (g_ResumeHeader.Width != g_Width || g_ResumeHeader.Height != g_Height || g_ResumeHeader.SceneIndex != g_SceneIndex ||
                g_ResumeHeader.StreamBase != g_Backend.StreamBase)
                return Fail("OnRender: checkpoint does not match the image size, scene or stream base");
            if (g_ResumeHeader.LaneWidth != Lanes || g_ResumeHeader.FmaDot != Fma)
                return Fail("OnRender: checkpoint was written with another arithmetic profile (lane width / FMA)");
            if (g_ResumeHeader.SceneHash != 0 && g_ResumeHeader.SceneHash != HashScene(Scene))
                return Fail("OnRender: checkpoint was written for another sphere table");
            for (u32 i = 0; i < g_Workers; ++i)     // every worker keeps its own rows of the uploaded accumulator (srt_write_accum)
                if (srt_write_accum(g_Handles[i], g_ResumeAccum.data()) != SRT_OK) return Fail("srt_write_accum", g_Handles[i]);
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
    for (u32 i = 0; i < g_Workers; ++i)                                         // the kick, main.cpp:851-856
        if (srt_render(g_Handles[i], &Camera, &Args) != SRT_OK) return Fail("srt_render", g_Handles[i]);
    g_FrameInFlight = true;
    g_InFlightSamples = Args.n_samples;
    if (g_Backend.Blocking) WaitForFrame();
    return CopyToOutput;
}
