"""CPU: the C++ host mirror (host/srt_scenes.cpp, srt_app.cpp) against goldens dumped from
the reference's own compiled code, and the properties of the synthetic stress scenes."""
import os

import numpy as np
import pytest

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(G, "ref_goldens.npz"))


@pytest.mark.parametrize("scene", [0, 1, 2])
def test_builtin_scene_tables_match_reference(srt, gold, scene):
    sp, sky, hdr = srt.host_scene(scene)
    assert sp.view(np.uint8).reshape(len(sp), 80).tobytes() == gold[f"scene{scene}_spheres"].tobytes()
    assert int(sky) == int(gold[f"scene{scene}_sky"][0])
    assert hdr.tobytes() == gold[f"scene{scene}_header"].tobytes()


@pytest.mark.parametrize("scene", [0, 1, 2])
def test_camera_matches_reference(srt, gold, scene):
    h, w = gold[f"img_s{scene}_w8f_rgba"].shape
    cam = srt.host_camera(scene, w, h).view(np.float32).reshape(-1)
    ref = gold[f"img_s{scene}_w8f_camera"]
    for off in (0, 4, 8, 12, 16):
        assert cam[off:off + 3].tobytes() == ref[off:off + 3].tobytes()
    assert cam[20:22].tobytes() == ref[20:22].tobytes()


@pytest.mark.parametrize("scene,count", [(3, 1024), (4, 65536)])
def test_synthetic_scenes(srt, scene, count):
    sp, sky, hdr = srt.host_scene(scene)
    assert len(sp) == count and sky
    sp2, _, _ = srt.host_scene(scene)
    assert sp.tobytes() == sp2.tobytes()                       # deterministic
    p = sp["position"][1:, :3].astype(np.float64)
    r = sp["radius"][1:].astype(np.float64)
    assert (p[:, 1] - r >= -1e-6).all()                        # resting on / above the ground plane
    # non-overlap (checked on a sample; the generator enforces it for all pairs)
    idx = np.random.default_rng(0).choice(len(p), 400, replace=False)
    d = np.linalg.norm(p[idx, None, :] - p[None, :, :], axis=2)
    rr = r[idx, None] + r[None, :]
    d[np.arange(len(idx)), idx] = np.inf
    assert (d > rr).all()
    assert (sp["ior"] == 0).all()
    frac_em = (sp["emissive"][1:, :3].sum(1) > 0).mean()
    assert 0.08 < frac_em < 0.17


# ---- files either side of the path (host/srt_files.cpp): scene text format, checkpoints -----
@pytest.mark.parametrize("scene", [0, 1, 2, 3])
def test_scene_file_round_trip_is_bit_exact(srt, tmp_path, scene):
    sp, sky, hdr = srt.host_scene(scene)
    path = tmp_path / "s.scene"
    srt.save_scene_file(scene, path)
    srt.load_scene_file(4, path)                               # into the 64K slot (replaced)
    sp2, sky2, hdr2 = srt.host_scene(4)
    srt.reset_scene(4)
    assert len(sp2) == len(sp) and sky2 == sky and hdr2.tobytes() == hdr.tobytes()
    for f in ("position", "radius", "color", "emissive", "specular", "ior"):
        a, b = np.ascontiguousarray(sp[f]), np.ascontiguousarray(sp2[f])
        if a.ndim == 2:
            a, b = np.ascontiguousarray(a[:, :3]), np.ascontiguousarray(b[:, :3])
        assert a.tobytes() == b.tobytes(), f


def test_scene_file_hand_written_and_malformed(srt, tmp_path):
    good = tmp_path / "g.scene"
    good.write_text("# two spheres\nsky 1\nlookat 0 0.5 0\ncamera 3 0.25 1\n"
                    "sphere 0 -100 0 100  0.5 0.5 0.5  0 0 0  0 0\n"
                    "  sphere 0 0.5 0 0.5  1 1 1  0 0 0  1 1.5\n\n")
    srt.load_scene_file(3, good)
    sp, sky, hdr = srt.host_scene(3)
    assert len(sp) == 2 and sky and hdr.tolist() == [0.0, 0.5, 0.0, 3.0, 0.25, 1.0]
    assert sp["radius"].tolist() == [100.0, 0.5] and sp["ior"].tolist() == [0.0, 1.5] and sp["specular"].tolist() == [0.0, 1.0]
    for text in ("sphere 1 2 3\n", "sky 1\n", "bogus 1\nsphere 0 0 0 1 1 1 1 0 0 0 0 0\n"):
        bad = tmp_path / "b.scene"
        bad.write_text(text)
        with pytest.raises(srt.SrtError):
            srt.load_scene_file(3, bad)
    with pytest.raises(srt.SrtError):
        srt.load_scene_file(3, tmp_path / "missing.scene")
    assert len(srt.host_scene(3)[0]) == 2                      # failed loads leave the slot alone
    srt.reset_scene(3)
    assert len(srt.host_scene(3)[0]) == 1024


def test_checkpoint_round_trip_and_rejects_damage(srt, tmp_path):
    rng = np.random.default_rng(5)
    acc = rng.random((7, 13, 4), dtype=np.float32)
    acc[0, 0, 0] = np.float32(np.inf)
    path = tmp_path / "a.ckpt"
    srt.save_checkpoint(path, acc, samples=11, scene_index=2, stream_base=(1 << 40) + 3, lane_width=4, fma_dot=0)
    a2, h = srt.load_checkpoint(path)
    assert a2.tobytes() == acc.tobytes()
    assert h == {"width": 13, "height": 7, "samples": 11, "scene_index": 2, "lane_width": 4, "fma_dot": 0, "stream_base": (1 << 40) + 3}
    raw = path.read_bytes()
    # "SRTCKPT2" | 4 x u32 | u64 stream base | 2 x u32 | u64 scene hash | payload
    assert raw[:8] == b"SRTCKPT2" and len(raw) == 8 + 40 + acc.nbytes
    assert int.from_bytes(raw[40:48], "little") == srt.host_scene_hash(2) != 0
    huge = bytearray(raw)
    huge[8:16] = (40000).to_bytes(4, "little") + (40000).to_bytes(4, "little")    # header asks for 25.6 GB: rejected by file size, before any allocation
    for damaged in (raw[:-4], raw + b"\0", b"XRTCKPT2" + raw[8:], b"SRTCKPT1" + raw[8:], raw[:20], bytes(huge)):
        bad = tmp_path / "bad.ckpt"
        bad.write_bytes(damaged)
        with pytest.raises(srt.SrtError):
            srt.load_checkpoint(bad)


def test_scene_hash_follows_the_sphere_table(srt, tmp_path):
    """The checkpoint's scene hash (FNV-1a of the sphere records + sky flag) changes when the slot's scene does."""
    h3 = srt.host_scene_hash(3)
    assert h3 == srt.host_scene_hash(3) and h3 != srt.host_scene_hash(2)
    path = tmp_path / "s.scene"
    srt.save_scene_file(3, path)
    text = path.read_text().splitlines()
    text[-1] = text[-1].replace("sphere ", "sphere  ", 1)       # same values: same hash
    path.write_text("\n".join(text) + "\n")
    srt.load_scene_file(3, path)
    assert srt.host_scene_hash(3) == h3
    path.write_text("\n".join(text[:-1]) + "\n")               # one sphere fewer
    srt.load_scene_file(3, path)
    assert srt.host_scene_hash(3) != h3
    srt.reset_scene(3)
    assert srt.host_scene_hash(3) == h3


def test_magic_division_is_exact():
    """Work-item decode of the pool kernel (csrc/srt_device.cuh magic_div, host side make_magic in srt_backend.cu):
    x // d == (x * m) >> (30 + l) with m = ceil(2^(30+l) / d), 2^l >= d, for every x < 2^30.  Checked here on the
    arithmetic itself (uint64 numpy) for the divisors the launches use (blocks per row, band sizes, band x chunks)
    and random ones, at the multiples of d, their neighbours and random x."""
    rng = np.random.default_rng(11)
    ds = set(range(1, 700)) | {480 * 32 * 2, 480 * 32 * 68 * 4, 240 * 32 * 270, 15360, 30720, 983040, (1 << 26) - 1, 1 << 26}
    ds |= set(int(v) for v in rng.integers(1, 1 << 27, 300))
    for d in sorted(ds):
        l = 0
        while (1 << l) < d:
            l += 1
        m = ((1 << (30 + l)) + d - 1) // d
        assert m < (1 << 32), d
        kmax = ((1 << 30) - 1) // d
        ks = np.unique(np.clip(np.concatenate([rng.integers(0, kmax + 1, 200), [0, 1, kmax]]), 0, kmax)).astype(np.uint64)
        xs = np.concatenate([ks * np.uint64(d), ks * np.uint64(d) + np.uint64(d - 1), np.maximum(ks * np.uint64(d), np.uint64(1)) - np.uint64(1),
                             rng.integers(0, 1 << 30, 400).astype(np.uint64), [np.uint64((1 << 30) - 1)]])
        xs = xs[xs < np.uint64(1 << 30)]
        # (x * m) < 2^62 fits uint64
        q = (xs * np.uint64(m)) >> np.uint64(30 + l)
        assert np.array_equal(q, xs // np.uint64(d)), d


def _read_png(path):
    """Minimal independent PNG reader (8-bit RGB, filter 0 only): checks signature, chunk CRCs, zlib stream."""
    import struct
    import zlib
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    at, chunks = 8, []
    while at < len(data):
        (n,) = struct.unpack(">I", data[at:at + 4])
        typ, body = data[at + 4:at + 8], data[at + 8:at + 8 + n]
        (crc,) = struct.unpack(">I", data[at + 8 + n:at + 12 + n])
        assert zlib.crc32(typ + body) == crc, typ
        chunks.append((typ, body))
        at += 12 + n
    assert [c[0] for c in chunks] == [b"IHDR", b"IDAT", b"IEND"]
    w, h, depth, colour, comp, flt, lace = struct.unpack(">IIBBBBB", chunks[0][1])
    assert (depth, colour, comp, flt, lace) == (8, 2, 0, 0, 0)
    raw = np.frombuffer(zlib.decompress(chunks[1][1]), np.uint8).reshape(h, 1 + 3 * w)   # checks Adler-32 too
    assert not raw[:, 0].any()
    return raw[:, 1:].reshape(h, w, 3)


@pytest.mark.parametrize("w,h", [(1, 1), (7, 5), (300, 90), (257, 256)])   # 300x90 and 257x256 span several 65,535-byte stored blocks
def test_image_writers(srt, tmp_path, w, h):
    rng = np.random.default_rng(w * 1000 + h)
    rgba = rng.integers(0, 2**32, size=(h, w), dtype=np.uint32)
    want = np.stack([(rgba >> s) & 0xFF for s in (0, 8, 16)], axis=-1).astype(np.uint8)[::-1]   # file rows are top-down
    png, ppm = tmp_path / "a.png", tmp_path / "a.ppm"
    srt.write_image(png, rgba)
    srt.write_image(ppm, rgba)
    assert np.array_equal(_read_png(png), want)
    blob = open(ppm, "rb").read()
    head = b"P6\n%d %d\n255\n" % (w, h)
    assert blob.startswith(head) and len(blob) == len(head) + 3 * w * h
    assert np.array_equal(np.frombuffer(blob[len(head):], np.uint8).reshape(h, w, 3), want)
    srt.write_image(tmp_path / "b.ppm", rgba, kind=2)                      # explicit kind beats the extension
    assert np.array_equal(_read_png(tmp_path / "b.ppm"), want)
    acc = rng.standard_normal((h, w, 4)).astype(np.float32)
    srt.write_pfm(tmp_path / "a.pfm", acc)
    blob = open(tmp_path / "a.pfm", "rb").read()
    head = b"PF\n%d %d\n-1.0\n" % (w, h)
    assert blob.startswith(head)
    assert np.array_equal(np.frombuffer(blob[len(head):], np.float32).reshape(h, w, 3), acc[:, :, :3])
    with pytest.raises(srt.SrtError):
        srt.write_image(tmp_path / "no_such_dir" / "a.png", rgba)
