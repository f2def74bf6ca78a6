#!/usr/bin/env python3
"""bench.py -- headline benchmark of the B200 path-tracing backend.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload NAME]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

Metric (BASELINE.json): Mrays/s -- path segments per second, a "ray" counted exactly like the
reference's RaysCastInThread (main.cpp:390).  Default workload = BASELINE.json configs[2], the
configuration the multi-GPU / %-of-FP32-peak metric is quoted on: synthetic 1,024-sphere
scene, 3840x2160, 256 spp, samples sharded over the N GPUs (strong scaling: the job is
fixed, rank g renders samples k = g mod N into an f32 SUM buffer, one NCCL reduce to rank 0,
which resolves mean + tonemap).  At N = 1 the exact running mean of main.cpp:484-489 is used.
One "step" = one complete render of the workload.  Prints ONE JSON line on rank 0.

Other workloads (--workload): demo1080 (configs[1]), stress64k (configs[4]), and the
progressive-latency protocol of configs[3] is always measured as `frame_latency_1080p`.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (scene index in the host mirror, width, height, spp, description)
    "synth1k_4k_256spp": (3, 3840, 2160, 256, "BASELINE configs[2]: synthetic 1,024-sphere scene, 3840x2160, 256 spp, sample-sharded"),
    "demo1080_64spp": (0, 1920, 1080, 64, "BASELINE configs[1]: built-in demo scene 0 (RGB glass, 5 spheres), 1920x1080, 64 spp"),
    "stress64k_1080_16spp": (4, 1920, 1080, 16, "BASELINE configs[4]: 65,536-sphere brute-force scene, 1920x1080, 16 spp"),
    "smoke": (3, 480, 270, 8, "tiny smoke workload (not a benchmark)"),
}
FLOP_PER_TEST = 20          # always-executed FP32 flops of one ray-sphere test
LANE_WIDTH, FMA_DOT = 8, True   # the x64 AVX2 profile (SIMD_WIDTH 8, FMA dot)


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}


def load_measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except OSError:
        return {}


# ---------------------------------------------------------------------------------------------
# reference arm / CPU baseline: the reference's OWN compiled CPU code (oracle/_ref), all host
# threads, on a bounded sample of the same workload.  (bench.py may execute oracle/ only here.)
# ---------------------------------------------------------------------------------------------
def cpu_reference_run(srt, scene_idx, width, height, budget_s, repeats=1):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import orclib
    if not orclib.have_ref():
        raise RuntimeError("oracle/_ref is missing (build it where /root/reference exists: python oracle/build_ref.py)")
    sp, sky, hdr = srt.host_scene(scene_idx)
    cores = os.cpu_count() or 1
    best = None
    # The reference ships two lane widths and two tile callbacks; time them all on a small
    # probe and keep the fastest as "the reference CPU path" (BASELINE.md section 4).
    probe_w, probe_h = 160, 90
    for tag in ("w8", "w4"):
        R = orclib.Ref(tag, cores)
        R.set_scene(1, sp, sky, hdr)
        for simd in (1, 0):
            r = R.render(1, probe_w, probe_h, 1, simd=simd, threads=cores, flags=0)
            rate = r["rays"] / max(r["ms"], 1e-6)      # rays per ms
            if best is None or rate > best[0]:
                best = (rate, tag, simd, R)
    rate, tag, simd, R = best
    # Size the sample: same scene and camera, the workload's aspect ratio at reduced
    # resolution / spp so that one run costs about budget_s of wall time.
    # Enough 32x32 tiles (main.cpp:9) to keep every worker thread busy.
    div = 4
    while div > 1 and ((width // div + 31) // 32) * ((height // div + 31) // 32) < 4 * cores:
        div //= 2
    w, h = max(32, width // div), max(32, height // div)
    rays_per_spp = 2.2 * w * h
    spp = int(max(1, min(64, budget_s * 1000.0 * rate / rays_per_spp)))
    results = []
    for _ in range(repeats):
        r = R.render(1, w, h, spp, simd=simd, threads=cores, flags=0)
        results.append((r["rays"], r["ms"]))
    rays = sum(x[0] for x in results)
    ms = sum(x[1] for x in results)
    kind = f"reference ({'RenderTile SIMD_WIDTH=' + tag[1:] if simd else 'RenderTileScalar'}, g++ -O3 {'-mavx2 -mfma' if tag == 'w8' else '-msse4.2'}, as-shipped RNG)"
    return dict(mrays=rays / ms / 1e3, rays=rays, ms=ms, cores=cores, sample=f"{w}x{h}, {spp} spp, same scene/camera, {len(results)} run(s)", variant=kind,
                per_run=[dict(rays=a, ms=b) for a, b in results])


def run_reference_arm(args):
    rank, _, world = dist_env()
    if rank != 0:
        return 0
    srt = importlib.import_module("simd-ray-tracer_b200")
    scene_idx, w, h, spp, desc = WORKLOADS[args.workload]
    for _ in range(args.warmup):
        cpu_reference_run(srt, scene_idx, w, h, budget_s=1.0)
    t0 = time.perf_counter()
    r = cpu_reference_run(srt, scene_idx, w, h, budget_s=max(2.0, 40.0 / max(1, args.steps)), repeats=args.steps)
    wall = time.perf_counter() - t0
    sp, _, _ = srt.host_scene(scene_idx)
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": r["mrays"], "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms"] / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "description": desc, "spheres": int(len(sp)), "width": w, "height": h, "spp": spp,
                   "step": "bounded sample of the workload on the host CPU: " + r["sample"]},
        "cpu_baseline": {"value": r["mrays"], "unit": "Mrays/s", "cores": r["cores"], "kind": "reference", "sample": r["sample"], "variant": r["variant"]},
        "e2e": {"value": r["mrays"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_s": wall,
    }
    print(json.dumps(line), file=_STDOUT, flush=True)
    return 0


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
def run_gpu_arm(args):
    import torch
    rank, local_rank, world = dist_env()
    if world != args.gpus and world != 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    n_gpus = world
    srt = importlib.import_module("simd-ray-tracer_b200")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the backend has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if n_gpus > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    scene_idx, W, H, SPP, desc = WORKLOADS[args.workload]
    shard = srt.sharding.plan(rank, n_gpus, SPP)
    spheres, sky, hdr = srt.host_scene(scene_idx)
    camera = srt.host_camera(scene_idx, W, H)
    n_pad = (len(spheres) + 7) // 8 * 8

    b = srt.Backend(local_rank)
    info0 = b.device_info()
    b.set_scene(spheres, sky, LANE_WIDTH, FMA_DOT)
    b.resize(W, H)
    # torch owns the framebuffers so that NCCL can reduce them in place
    accum = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    rgba = torch.zeros((H, W), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    b.bind_device_buffers(accum.data_ptr(), rgba.data_ptr())
    ext_stream = torch.cuda.ExternalStream(b.stream_ptr)      # NCCL ops ordered on the backend's stream

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    kernel_ms, kernel_rays = [], []

    def step():
        """One complete render of the workload, inputs resident in HBM."""
        with torch.cuda.stream(ext_stream):
            accum.zero_()
        if n_gpus == 1:
            b.render(camera, 0, SPP, stride=1, mode=srt.ACCUM_MEAN)
            ms, rays = b.wait()
        else:
            b.render(camera, shard["first"], shard["count"], stride=shard["stride"], mode=srt.ACCUM_SUM)
            with torch.cuda.stream(ext_stream):
                srt.sharding.reduce_sum(dist, accum, dst=0)
            if rank == 0:
                b.resolve_sum(SPP)
            ms, rays = b.wait()
        kernel_ms.append(ms)
        kernel_rays.append(rays)
        return rays

    for _ in range(args.warmup):
        step()
    barrier()
    launches0 = b.device_info()["launches"]
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    kernel_ms.clear(); kernel_rays.clear()
    # Device-side clock: CUDA events on the stream every launch of a step goes to (the backend's
    # own stream, which also carries the memset and the NCCL reduce); wall clock kept beside it.
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    ev0.record(ext_stream)
    rays_local = 0
    for _ in range(args.steps):
        rays_local += step()
    ev1.record(ext_stream)
    barrier()
    wall_elapsed = time.perf_counter() - t0
    elapsed = ev0.elapsed_time(ev1) * 1e-3
    clocks = sampler.stop() if rank == 0 else None
    launches = b.device_info()["launches"] - launches0

    t = torch.tensor([elapsed, float(rays_local), float(launches), float(np.mean(kernel_ms)), wall_elapsed], dtype=torch.float64, device="cuda")
    if dist is not None:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        elapsed, rays_total, launches_total, kms, wall_elapsed = tmax[0].item(), tsum[1].item(), tsum[2].item(), tmax[3].item(), tmax[4].item()
    else:
        elapsed, rays_total, launches_total, kms = t[0].item(), t[1].item(), t[2].item(), t[3].item()
    value = rays_total / elapsed / 1e6
    ms_per_step = elapsed * 1e3 / args.steps

    # ---- end to end through the C ABI with HOST buffers (rank-local pinned memory) ------------
    pin_scene = torch.empty(spheres.nbytes, dtype=torch.uint8).pin_memory()
    pin_scene.numpy()[:] = spheres.view(np.uint8).reshape(-1)
    pin_cam = torch.empty(camera.nbytes, dtype=torch.uint8).pin_memory()
    pin_cam.numpy()[:] = camera.view(np.uint8).reshape(-1)
    pin_img = torch.empty(H * W * 4, dtype=torch.uint8).pin_memory()
    scene_host = np.frombuffer(pin_scene.numpy(), dtype=srt.sphere_dtype)
    cam_host = np.frombuffer(pin_cam.numpy(), dtype=srt.camera_dtype)
    img_host = pin_img.numpy().view(np.uint32).reshape(H, W)

    def e2e_step():
        b.set_scene(scene_host, sky, LANE_WIDTH, FMA_DOT)         # H2D: scene table
        with torch.cuda.stream(ext_stream):
            accum.zero_()
        if n_gpus == 1:
            b.render(cam_host, 0, SPP, stride=1, mode=srt.ACCUM_MEAN)   # H2D: camera (kernel params)
        else:
            b.render(cam_host, shard["first"], shard["count"], stride=shard["stride"], mode=srt.ACCUM_SUM)
            with torch.cuda.stream(ext_stream):
                srt.sharding.reduce_sum(dist, accum, dst=0)
            if rank == 0:
                b.resolve_sum(SPP)
        _, rays = b.wait()
        if rank == 0:
            b.read_rgba8(img_host)                                 # D2H: the presented RGBA8 image
        return rays

    e2e_steps = max(1, min(args.steps, 2))
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    e2e_rays = 0
    for _ in range(e2e_steps):
        e2e_rays += e2e_step()
    barrier()
    e2e_elapsed = time.perf_counter() - t0
    t = torch.tensor([e2e_elapsed, float(e2e_rays)], dtype=torch.float64, device="cuda")
    if dist is not None:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        e2e_elapsed, e2e_rays = tmax[0].item(), tsum[1].item()
    e2e_value = e2e_rays / e2e_elapsed / 1e6
    checksum = int(img_host.view(np.uint8)[::997].astype(np.uint64).sum()) if rank == 0 else 0

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return 0

    # ---- roofline of the dominant kernel (render_paths): FP32 CUDA-core pipe ------------------
    peaks = load_measured_peaks()
    sm_max_mhz = float(peaks.get("sm_max_mhz", 1965.0))
    peak_tflops = info0["sm_count"] * 128 * 2 * sm_max_mhz * 1e6 / 1e12
    rays_per_launch = float(np.mean(kernel_rays))
    achieved = FLOP_PER_TEST * rays_per_launch * n_pad / (kms * 1e-3) / 1e12
    pair_kernel = n_pad >= 64
    primary_per_launch = float(W * H * SPP) / max(1, n_gpus)   # one render launch per step per rank
    issue_cycles_per_sphere_pass = (info0["sm_count"] * 4 * sm_max_mhz * 1e6 * 64) / (rays_per_launch * n_pad / (kms * 1e-3))
    # bundle mode (csrc/srt_bundle.cuh; auto from 384 spheres and 8 samples per launch): the primary passes run the
    # level-1 test once per (pixel, sphere) for all samples of the pixel, 1/32 of the spheres per lane
    bundle = pair_kernel and n_pad >= 384 and SPP // max(1, n_gpus) >= 8 and os.environ.get("SRT_BUNDLE", "1") != "0"
    f_prim = min(1.0, primary_per_launch / max(1.0, rays_per_launch))
    floor_pass = (8.06 * (1.0 - f_prim) + 0.35 * f_prim) if bundle else (8.06 if pair_kernel else 30.0)
    exec_flop = (4.0 * (1.0 - f_prim) + 6.0 / 64.0 * f_prim) if bundle else (4.0 if pair_kernel else 20.0)
    roofline = {
        "bound": "fp32",
        "kernel": ("render_pool<8,true> (CTA-owned path slots, warp passes of 64 rays, two rays per thread: level-1 slab test of every "
                   "sphere (2 FFMA2 + 2 FSETP per sphere per ray pair"
                   + ("; bundle mode: a primary pass holds the samples of one pixel and runs the test once per (pixel, sphere), "
                      "1/32 of the spheres per lane" if bundle else "")
                   + "), level-2 fused 7-FMA filter on the survivors, the reference's arithmetic on what is left"
                   + ("; tables streamed through per-warp TMA stages)" if n_pad > 8192 else "; tables resident in shared memory)")
                   if pair_kernel else "render_paths<8,true,true> (packed FFMA2 intersection loop)"),
        "achieved": achieved, "peak": peak_tflops, "unit": "TFLOP/s", "frac": achieved / peak_tflops,
        "peak_source": f"{info0['sm_count']} SMs x 128 FP32 lanes x 2 x {sm_max_mhz:.0f} MHz (sm_max_mhz of MEASURED_PEAKS.json); "
                       "no tensor-core or HBM roof applies: the loop has no dense contraction and its scene is shared-memory resident",
        "algorithmic_flop_per_test": FLOP_PER_TEST, "tests_per_launch": rays_per_launch * n_pad,
        "executed_flop_per_test": exec_flop,
        "frac_executed": exec_flop * rays_per_launch * n_pad / (kms * 1e-3) / 1e12 / peak_tflops,
        "bundle_mode": bool(bundle), "primary_ray_fraction": f_prim,
        "issue_cycles_per_sphere_per_warp_pass": issue_cycles_per_sphere_pass,
        "issue_floor_cycles_per_sphere_per_warp_pass": floor_pass,
        "issue_slot_frac": floor_pass / issue_cycles_per_sphere_pass,
        "issue_bound_note": ("`achieved`/`frac` count the reference's 20 flop per ray-sphere test (SURVEY 8d) and exceed the FP32 peak "
                             "because the kernel does not execute them: every sphere is tested against every ray with a conservative "
                             "2-FMA slab test (4 executed flop), ~2 % go on to a 7-FMA filter, ~0.2 % to the reference's 20-flop "
                             "sequence (bit-exact results; soundness proof DESIGN.md 5.3, on-device audit srt_filter_audit).  The "
                             "binding resource is the SM sub-partition's issue port: a packed FFMA2 holds it for 2 cycles and LDS / "
                             "FSETP / LOP3 are not issued in its shadow (scripts/microbench/issue_mix.cu), so the level-1 floor is "
                             "2x2 + 2 + 1 + 0.75 + loop = 8.06 issue cycles per sphere per 64 rays; `issue_slot_frac` = that floor / "
                             "the measured whole-kernel cycles per sphere per 64 rays (regenerate, level 2, select, shade included)"
                             + ("; in bundle mode the primary passes' share of the floor is 0.35 cycles per sphere (one 6-flop test per "
                                "(pixel, sphere) for the 64 samples of the pass, DESIGN.md 5.2d), weighted by primary_ray_fraction" if bundle else "")
                             if pair_kernel else
                             "15 of the 20 flops' instructions are mandatory separate roundings (parity), so 15 FP32 issue cycles/test "
                             "is the hard floor: frac <= 20/30 = 0.667; see DESIGN.md"),
        "kernel_ms": kms,
        "traffic": None, "hbm_gbs_measured_peak": peaks.get("hbm_gbs"),
    }
    prof = os.path.join(ROOT, "profiles", "r1_roofline_traffic.json")
    if os.path.exists(prof):
        try:
            with open(prof) as f:
                roofline.update(json.load(f))
        except (OSError, ValueError):
            pass

    # ---- HBM roofline of accumulate/tonemap standalone (A8-A9): resolve_sum_kernel -------------
    # 36 algorithmic bytes per pixel (16 R + 16 W accumulator, 4 W RGBA8); at N = 1 the
    # path kernel does this work itself, so the kernel is timed here on the step's own buffers.
    hbm = None
    if not args.no_resolve_bench:
        reps = 20
        for _ in range(3):
            b.resolve_sum(SPP)
        r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        r0.record(ext_stream)
        for _ in range(reps):
            b.resolve_sum(SPP)
        r1.record(ext_stream)
        b.wait()
        torch.cuda.synchronize()
        r_ms = r0.elapsed_time(r1) / reps
        gbs = 36.0 * W * H / (r_ms * 1e-3) / 1e9
        peak_gbs = float(peaks.get("hbm_gbs", 6451.8))
        hbm = {"bound": "hbm", "kernel": "resolve_sum_kernel (mean + tonemap + pack after the cross-GPU sum)", "achieved": gbs, "peak": peak_gbs,
               "unit": "GB/s", "frac": gbs / peak_gbs, "bytes_per_pixel": 36, "pixels": W * H, "launch_us": r_ms * 1e3, "launches": reps,
               "l2": f"working set {20 * W * H / 1e6:.0f} MB " + ("> 126 MB L2" if 20 * W * H > 126e6 else "fits the 126 MB L2: not an HBM number"),
               "peak_source": "MEASURED_PEAKS.json hbm_gbs (copy bandwidth, burst)", "traffic": None}

    # ---- secondary metric: ms/frame @1080p, progressive 1 spp/frame (BASELINE configs[3]) --------
    lat = None
    if n_gpus == 1 and not args.no_latency:
        b.bind_device_buffers(None, None)
        s0, sky0, _ = srt.host_scene(0)
        cam0 = srt.host_camera(0, 1920, 1080)
        b.set_scene(s0, sky0, LANE_WIDTH, FMA_DOT)
        b.resize(1920, 1080)
        dev, wall = [], []
        for k in range(args.latency_frames + 50):
            tw = time.perf_counter()
            b.render(cam0, k, 1)
            ms, _ = b.wait()
            if k >= 50:
                dev.append(ms); wall.append((time.perf_counter() - tw) * 1e3)
        lat = {"frames": len(dev), "device_ms_p50": float(np.percentile(dev, 50)), "device_ms_p99": float(np.percentile(dev, 99)),
               "wall_ms_p50": float(np.percentile(wall, 50)), "wall_ms_p99": float(np.percentile(wall, 99)),
               "protocol": "BASELINE configs[3]: scene 0, 1920x1080, 1 spp/frame, 50 warm-up frames, exact running mean on device, read-back excluded"}

    # ---- CPU baseline: the reference's own code on this box's host cores, bounded sample ---------
    cpu = None
    if n_gpus == 1 and not args.no_cpu:
        try:
            r = cpu_reference_run(srt, scene_idx, W, H, budget_s=args.cpu_seconds)
            cpu = {"value": r["mrays"], "unit": "Mrays/s", "cores": r["cores"], "kind": "reference", "sample": r["sample"], "variant": r["variant"]}
        except Exception as e:   # noqa: BLE001 - the baseline is reported, never required
            cpu = {"value": None, "unit": "Mrays/s", "cores": os.cpu_count(), "kind": "reference", "sample": f"unavailable: {e}"}

    line = {
        "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": n_gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "timing": "CUDA events on the launching stream around the K steps, max over ranks; wall clock of the same region in wall_ms_per_step",
        "wall_ms_per_step": wall_elapsed * 1e3 / args.steps,
        "config": {"workload": args.workload, "description": desc, "spheres": int(len(spheres)), "width": W, "height": H, "spp": SPP,
                   "max_bounces": 5, "lane_width": LANE_WIDTH, "fma_dot": FMA_DOT,
                   "parallelism": f"sample-sharded x{n_gpus}" + (" + NCCL reduce(sum) to rank 0" if n_gpus > 1 else " (exact running mean)"),
                   "l2": (f"accumulator + RGBA8 working set {H * W * 20 / 1e6:.0f} MB" + (" > 126 MB L2" if H * W * 20 > 126e6 else " (fits L2)")
                          + "; the step is FP32-issue-bound with the sphere tables in shared memory, HBM traffic is < 0.1 % of the "
                            "step (ncu: dram 7 MB per 42 ms launch), so L2 residency cannot change the number")},
        "rays_per_step": rays_total / args.steps, "gpu_launches": int(launches_total),
        "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": int(spheres.nbytes + camera.nbytes),
                "d2h_bytes_per_step": int(H * W * 4), "steps": e2e_steps, "ms_per_step": e2e_elapsed * 1e3 / e2e_steps,
                "image_checksum": checksum},
        "clocks": clocks, "roofline": roofline,
    }
    if cpu is not None:
        line["cpu_baseline"] = cpu
    if hbm is not None:
        line["roofline_hbm"] = hbm
    if lat is not None:
        line["frame_latency_1080p"] = lat
    print(json.dumps(line), file=_STDOUT, flush=True)
    if dist is not None:
        dist.destroy_process_group()
    return 0


_STDOUT = sys.stdout


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="synth1k_4k_256spp", choices=sorted(WORKLOADS))
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="wall-time budget of the CPU baseline sample")
    ap.add_argument("--latency-frames", type=int, default=1000)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-latency", action="store_true")
    ap.add_argument("--no-resolve-bench", action="store_true")
    args = ap.parse_args()
    # stdout carries exactly ONE JSON line: libraries that write to fd 1 (NCCL prints its version
    # banner there) are sent to stderr while the benchmark runs.
    global _STDOUT
    sys.stdout.flush()
    _STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
