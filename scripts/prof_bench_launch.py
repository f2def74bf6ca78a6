"""The launch bench.py times, alone, for ncu:  python scripts/prof_bench_launch.py [workload] [launches]
Renders the workload `launches` times (default 2: profile the second with `ncu -k regex:render_ -s 1 -c 1`)."""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

srt = importlib.import_module("simd-ray-tracer_b200")
name = sys.argv[1] if len(sys.argv) > 1 else "synth1k_4k_256spp"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2
scene, w, h, spp, desc = bench.WORKLOADS[name]
sp, sky, hdr = srt.host_scene(scene)
cam = srt.host_camera(scene, w, h)
b = srt.Backend(0)
b.set_scene(sp, sky, bench.LANE_WIDTH, bench.FMA_DOT)
b.resize(w, h)
for i in range(n):
    b.render(cam, 0, spp)
    ms, rays = b.wait()
    print(f"{name}: launch {i}: {ms:.3f} ms, {rays} rays, {rays / ms / 1e3:.1f} Mrays/s", flush=True)
b.close()
