"""GPU: CUDA path vs the committed goldens produced by the reference's own compiled code
(no oracle in the loop)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
MODES = {"w8f": (8, 1), "w4i": (4, 0), "scalar": (1, 0)}


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(G, "ref_goldens.npz"))


@pytest.mark.parametrize("scene", [0, 1, 2])
@pytest.mark.parametrize("mode", list(MODES))
def test_cuda_matches_reference_goldens(backend, srt, gold, scene, mode):
    W, fma = MODES[mode]
    spheres = np.frombuffer(gold[f"scene{scene}_spheres"].tobytes(), srt.sphere_dtype)
    sky = bool(gold[f"scene{scene}_sky"][0])
    h, w = gold[f"img_s{scene}_{mode}_rgba"].shape
    rc = gold[f"img_s{scene}_{mode}_camera"]
    cam = np.zeros(1, srt.camera_dtype)
    cam["position"][0] = rc[0:4]; cam["z"][0] = rc[4:8]; cam["x"][0] = rc[8:12]; cam["y"][0] = rc[12:16]
    cam["film_center"][0] = rc[16:20]; cam["film_w"][0] = rc[20]; cam["film_h"][0] = rc[21]
    backend.set_scene(spheres, sky, W, bool(fma))
    backend.resize(w, h)
    backend.render(cam, 0, 2, record_primary=True)
    ms, rays = backend.wait()
    prim, t = backend.read_primary()
    assert np.array_equal(prim, gold[f"img_s{scene}_{mode}_prim"])                       # bit-exact primary ids
    assert np.array_equal(t.view(np.uint32), gold[f"img_s{scene}_{mode}_t"].view(np.uint32))
    assert np.array_equal(backend.read_accum().view(np.uint32), gold[f"img_s{scene}_{mode}_accum"].view(np.uint32))
    assert np.array_equal(backend.read_rgba8(), gold[f"img_s{scene}_{mode}_rgba"])
    assert rays == int(gold[f"img_s{scene}_{mode}_rays"][0])
