"""GPU: the -DSRT_SPIN_LIMIT debug build of the library (libsrt_cuda_dbg.so: every wait loop of the pool kernel traps
with a site code after N iterations instead of hanging the GPU) has the same ABI and renders the same bits."""
import hashlib
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import hashlib, importlib, sys
sys.path.insert(0, %r)
srt = importlib.import_module("simd-ray-tracer_b200")
out = []
for scene, w, h, spp in ((3, 320, 180, 20), (4, 96, 54, 8), (0, 200, 100, 3)):
    sp, sky, hdr = srt.host_scene(scene)
    b = srt.Backend(0)
    b.set_scene(sp, sky, 8, True)
    b.resize(w, h)
    b.render(srt.host_camera(scene, w, h, hdr), 0, spp)
    ms, rays = b.wait()
    out.append((rays, hashlib.sha256(b.read_accum().tobytes()).hexdigest(), hashlib.sha256(b.read_rgba8().tobytes()).hexdigest()))
    b.close()
print(out)
"""


def test_debug_library_renders_the_same_bits(srt):
    dbg = os.path.join(srt.PKG_DIR, "libsrt_cuda_dbg.so")
    assert os.path.exists(dbg), "run __graft_entry__.build()"
    res = []
    for lib in (None, dbg):
        env = dict(os.environ)
        env.pop("SRT_CUDA_LIB", None)
        if lib:
            env["SRT_CUDA_LIB"] = lib
        r = subprocess.run([sys.executable, "-c", SCRIPT % ROOT], capture_output=True, text=True, env=env, timeout=600)
        assert r.returncode == 0, r.stderr
        res.append(r.stdout.strip().splitlines()[-1])
    assert res[0] == res[1]


def test_audited_ring_protocol_under_load(srt):
    """scripts/soak.py on the debug library: jobs that fill the machine (1,024 and 482 spheres, bundle mode, several frames),
    every ring push / pop audited (slot ownership, one consumer per entry, the position an entry was written for, no wait
    longer than 0.2 ms), every job twice, then the hashes against the release library's.  Before the __syncwarp() after the
    ring pops this failed within the first 20 jobs."""
    dbg = os.path.join(srt.PKG_DIR, "libsrt_cuda_dbg.so")
    out = {}
    for tag, lib in (("release", None), ("debug", dbg)):
        env = dict(os.environ)
        env.pop("SRT_CUDA_LIB", None)
        if lib:
            env["SRT_CUDA_LIB"] = lib
        hashes = []
        for n, (fixed, cases) in enumerate((("1024,943,409,64,3", 12), ("482,687,446,33,2", 30), ("", 60))):
            path = os.path.join(ROOT, "gpurun_out", f"soak_test_{tag}_{n}.json")
            cmd = [sys.executable, os.path.join(ROOT, "scripts", "soak.py"), "--cases", str(cases), "--modes", "8:1", "--out", path]
            if fixed:
                cmd += ["--fixed", fixed]
            r = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=900)
            assert r.returncode == 0, (tag, fixed, r.stdout[-600:], r.stderr[-600:])
            with open(path) as f:
                hashes.append(__import__("json").load(f)["cases"])
        out[tag] = hashes
    assert out["release"] == out["debug"]
