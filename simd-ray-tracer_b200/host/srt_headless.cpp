// srt_headless.cpp -- headless Linux driver: the platform layer of the host mirror.
//
// Plays the role of wasm/wasm.cpp / win32/win32.cpp in the reference (main(), the
// platform-owned `image`, the per-frame OnRender loop and the statistics read-out of
// wasm/wasm.cpp:176-219) without a window: the presented image goes to a PPM file
// (rows flipped: the reference's row 0 is the image bottom, GL convention).
#include "srt_app.h"

#include <algorithm>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>

static void Usage() {
    fprintf(stderr,
            "usage: srt_headless [--scene 0..4] [--width W] [--height H] [--spp N] [--spf K]\n"
            "                    [--simd 0|1] [--lanes 4|8] [--fma 0|1] [--device D] [--gpus N] [--stream-base B]\n"
            "                    [--orbit FRAMES] [--out image.ppm|image.png] [--accum accum.f32]\n"
            "                    [--scene-file in.scene] [--dump-scene out.scene]\n"
            "                    [--checkpoint out.ckpt] [--resume in.ckpt] [--pfm accum.pfm] [--threads N]\n"
            "  scenes: 0 RGB Glass, 1 Floating Spheres, 2 One Weekend, 3 synthetic 1,024, 4 synthetic 65,536\n"
            "  --scene-file replaces slot --scene by the file's scene; --dump-scene writes slot --scene as text.\n"
            "  --resume continues a checkpoint (same size, scene, stream base): --spp more samples are added.\n"
            "  --pfm writes the linear accumulator as a PFM image.\n"
            "  --gpus N: devices D .. D+N-1 render interleaved rows of 8x4 blocks of every frame and worker 0 gathers them (NCCL);\n"
            "  the image is bit-identical for every N.  --threads is render_params.ThreadCount: workers used per frame (0 = all).\n"
            "  --spf: samples per OnRender frame (reference: 1).  --orbit: hold key A for FRAMES frames first.\n"
            "  --latency FRAMES: after the render, time FRAMES bare kicks (srt_render + srt_wait, --spf samples each, no read-back):\n"
            "  device and wall p50 / p99 per frame (BASELINE configs[3] measured from C++, without the Python binding).\n");
}

int main(int argc, char **argv) {
    u32 Scene = 0, W = 1280, H = 720, Spp = 1, Spf = 1, Orbit = 0;
    bool Simd = true;
    backend_params BP;
    const char *Out = nullptr, *AccumOut = nullptr, *SceneFile = nullptr, *DumpScene = nullptr, *CkptOut = nullptr, *CkptIn = nullptr, *PfmOut = nullptr;
    u32 ThreadCount = 0, LatencyFrames = 0;
    for (int i = 1; i < argc; ++i) {
        std::string A = argv[i];
        auto Next = [&]() -> const char * { if (i + 1 >= argc) { Usage(); exit(2); } return argv[++i]; };
        if (A == "--scene") Scene = atoi(Next());
        else if (A == "--width") W = atoi(Next());
        else if (A == "--height") H = atoi(Next());
        else if (A == "--spp") Spp = atoi(Next());
        else if (A == "--spf") Spf = atoi(Next());
        else if (A == "--simd") Simd = atoi(Next()) != 0;
        else if (A == "--lanes") BP.LaneWidth = atoi(Next());
        else if (A == "--fma") BP.FmaDot = atoi(Next()) != 0;
        else if (A == "--device") BP.DeviceId = atoi(Next());
        else if (A == "--gpus") BP.DeviceCount = atoi(Next());
        else if (A == "--stream-base") BP.StreamBase = strtoull(Next(), nullptr, 0);
        else if (A == "--orbit") Orbit = atoi(Next());
        else if (A == "--out") Out = Next();
        else if (A == "--accum") AccumOut = Next();
        else if (A == "--scene-file") SceneFile = Next();
        else if (A == "--dump-scene") DumpScene = Next();
        else if (A == "--checkpoint") CkptOut = Next();
        else if (A == "--resume") CkptIn = Next();
        else if (A == "--pfm") PfmOut = Next();
        else if (A == "--threads") ThreadCount = atoi(Next());
        else if (A == "--latency") LatencyFrames = atoi(Next());
        else { Usage(); return 2; }
    }
    if (Spf == 0) Spf = 1;
    if (Spp == 0) Spp = 1;
    BP.SamplesPerFrame = Spf;
    SetBackendParams(BP);

    if (Scene >= SceneCount) { Usage(); return 2; }
    if (SceneFile) {
        scene Loaded;
        if (!LoadSceneFile(SceneFile, &Loaded) || !SetScene(Scene, Loaded)) { fprintf(stderr, "cannot load scene file %s\n", SceneFile); return 1; }
        free(Loaded.ScalarSpheres);
    }
    if (DumpScene) {
        if (!SaveSceneFile(DumpScene, *GetScene(Scene))) { fprintf(stderr, "cannot write %s\n", DumpScene); return 1; }
        if (!Out && !AccumOut && !CkptOut) return 0;       // dump only: no GPU needed
    }
    u32 ResumedSamples = 0;
    if (CkptIn) {
        checkpoint_header Hd;
        std::vector<f32> Acc;
        if (!LoadCheckpoint(CkptIn, &Hd, &Acc)) { fprintf(stderr, "cannot read checkpoint %s\n", CkptIn); return 1; }
        if (Hd.SceneIndex != Scene || Hd.LaneWidth != (Simd ? BP.LaneWidth : 1u) || Hd.FmaDot != (u32)(Simd && BP.FmaDot)) {
            fprintf(stderr, "checkpoint %s was written for scene %u, lanes %u, fma %u\n", CkptIn, Hd.SceneIndex, Hd.LaneWidth, Hd.FmaDot);
            return 1;
        }
        W = Hd.Width; H = Hd.Height; BP.StreamBase = Hd.StreamBase;
        SetBackendParams(BP);
        if (!ResumeFromCheckpoint(Hd, Acc.data())) { fprintf(stderr, "bad checkpoint %s\n", CkptIn); return 1; }
        ResumedSamples = Hd.Samples;
    }

    init_params Init = {};
    if (!OnInit(&Init)) { fprintf(stderr, "OnInit failed: %s\n", LastError()); return 1; }

    std::vector<u32> Present((size_t)W * H);
    image Image = { Present.data(), W, H, format::R8G8B8A8_U32 };
    render_params RP = { ThreadCount, Simd, Scene };

    // Optional scripted orbit: hold A so every frame resets, like a user dragging the camera.
    u64 Rays = 0, TotalRays = 0;
    f64 Ms = 0.0, TotalMs = 0.0;
    for (u32 f = 0; f < Orbit; ++f) {
        SetKeyDown(key::A, true);
        while (!OnRender(Image, RP, &Rays, &Ms) && f != 0) {}
    }
    SetKeyDown(key::A, false);

    // Frame loop: (frames + 1) calls publish `frames` frames (SURVEY appendix A protocol).
    const u32 Frames = (Spp + Spf - 1) / Spf;
    std::vector<f64> FrameMs;
    SetKeyDown(key::R, true);                          // start from a clean accumulation
    const f64 Wall0 = QueryTimestampInMilliseconds();
    u32 Kicked = 0;
    bool First = true;
    while (Kicked < Frames) {
        const bool Published = OnRender(Image, RP, &Rays, &Ms);
        if (LastError()[0]) { fprintf(stderr, "OnRender failed: %s\n", LastError()); return 1; }
        SetKeyDown(key::R, false);
        if (Published && !First) { FrameMs.push_back(Ms); TotalRays += Rays; TotalMs += Ms; }
        if (Published || First) { Kicked += 1; First = false; }
        else WaitForFrame();                           // not ready: block instead of spinning
    }
    f64 LastMs = 0.0;
    u64 LastRays = 0;
    if (!PublishFrame(Image, &LastRays, &LastMs)) { fprintf(stderr, "PublishFrame failed: %s\n", LastError()); return 1; }
    FrameMs.push_back(LastMs); TotalRays += LastRays; TotalMs += LastMs;
    const f64 Wall = QueryTimestampInMilliseconds() - Wall0;

    if (Out && !WriteImage(Out, Present.data(), W, H)) { fprintf(stderr, "cannot write %s\n", Out); return 1; }
    if (CkptOut) {
        std::vector<f32> Acc((size_t)W * H * 4);
        ReadAccumulator(Acc.data());
        checkpoint_header Hd = { W, H, GetAccumulatedSampleCount(), Scene, BP.StreamBase, Simd ? BP.LaneWidth : 1u, (u32)(Simd && BP.FmaDot), HashScene(*GetScene(Scene)) };
        if (!SaveCheckpoint(CkptOut, Hd, Acc.data())) { fprintf(stderr, "cannot write %s\n", CkptOut); return 1; }
    }
    if (PfmOut) {
        std::vector<f32> Acc((size_t)W * H * 4);
        ReadAccumulator(Acc.data());
        if (!WritePfm(PfmOut, Acc.data(), W, H)) { fprintf(stderr, "cannot write %s\n", PfmOut); return 1; }
    }
    if (AccumOut) {
        std::vector<f32> Acc((size_t)W * H * 4);
        ReadAccumulator(Acc.data());
        FILE *F = fopen(AccumOut, "wb");
        if (!F) { fprintf(stderr, "cannot write %s\n", AccumOut); return 1; }
        fwrite(Acc.data(), sizeof(f32), Acc.size(), F);
        fclose(F);
    }
    // The bare kick of BASELINE configs[3]: what replaces WorkQueueStart + WorkQueueWaitUntilCompletion (main.cpp:852, base.h:175),
    // timed from C++: wall = srt_render + srt_wait, device = CUDA events around the kernel.
    f64 LatDevP50 = 0, LatDevP99 = 0, LatWallP50 = 0, LatWallP99 = 0, LatOverP50 = 0;
    if (LatencyFrames && GetDeviceCount() == 1) {
        const scene *Sc = GetScene(Scene);
        srt_camera Cam;
        ComputeCamera(*Sc, Sc->DefaultDistanceFromLookAt, Sc->DefaultXAngle, Sc->DefaultYHeight, W, H, &Cam);
        srt_render_args Args = {};
        Args.n_samples = Spf; Args.sample_stride = 1; Args.accum_mode = SRT_ACCUM_MEAN; Args.stream_base = BP.StreamBase;
        std::vector<f64> Dev, Wal, Over;
        u32 Sample = GetAccumulatedSampleCount();
        for (u32 f = 0; f < LatencyFrames + 50; ++f, Sample += Spf) {
            Args.first_sample = Sample;
            const f64 T0 = QueryTimestampInMilliseconds();
            f64 FrameDev = 0;
            if (srt_render(GetBackend(), &Cam, &Args) != SRT_OK || srt_wait(GetBackend(), &FrameDev, nullptr) != SRT_OK) {
                fprintf(stderr, "latency loop failed: %s\n", srt_last_error(GetBackend()));
                return 1;
            }
            const f64 T1 = QueryTimestampInMilliseconds();
            if (f >= 50) { Dev.push_back(FrameDev); Wal.push_back(T1 - T0); Over.push_back(T1 - T0 - FrameDev); }
        }
        auto Pct = [](std::vector<f64> &V, f64 P) { std::sort(V.begin(), V.end()); return V[std::min(V.size() - 1, (size_t)(V.size() * P))]; };
        LatDevP50 = Pct(Dev, 0.5); LatDevP99 = Pct(Dev, 0.99); LatWallP50 = Pct(Wal, 0.5); LatWallP99 = Pct(Wal, 0.99); LatOverP50 = Pct(Over, 0.5);
    }
    std::sort(FrameMs.begin(), FrameMs.end());
    const f64 P50 = FrameMs[FrameMs.size() / 2], P99 = FrameMs[std::min(FrameMs.size() - 1, (size_t)(FrameMs.size() * 0.99))];
    const scene *S = GetScene(Scene);
    printf("{\"scene\": \"%s\", \"spheres\": %u, \"width\": %u, \"height\": %u, \"spp\": %u, \"frames\": %zu, "
           "\"rays\": %llu, \"device_ms\": %.4f, \"wall_ms\": %.3f, \"mrays_per_s\": %.2f, "
           "\"frame_ms_p50\": %.4f, \"frame_ms_p99\": %.4f, \"samples_accumulated\": %u, \"resumed_from\": %u, \"gpus\": %u, "
           "\"latency_frames\": %u, \"latency_device_ms_p50\": %.5f, \"latency_device_ms_p99\": %.5f, \"latency_wall_ms_p50\": %.5f, "
           "\"latency_wall_ms_p99\": %.5f, \"latency_host_overhead_ms_p50\": %.5f}\n",
           S->Name, S->SphereCount, W, H, Frames * Spf, FrameMs.size(), (unsigned long long)TotalRays, TotalMs, Wall,
           TotalMs > 0 ? TotalRays / TotalMs / 1e3 : 0.0, P50, P99, GetAccumulatedSampleCount(), ResumedSamples, GetDeviceCount(),
           LatencyFrames, LatDevP50, LatDevP99, LatWallP50, LatWallP99, LatOverP50);
    OnShutdown();
    return 0;
}
