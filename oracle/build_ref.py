#!/usr/bin/env python3
"""Build recipe for oracle/_ref: the reference's OWN CPU path, compiled with g++.

TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is imported, linked or executed by
the product (simd-ray-tracer_b200/, include/); only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs may use it.

What this does
--------------
Reads the reference sources where they lie (/root/reference/{base.h,x64_math.h,
main.cpp}), applies the mechanical patch list P1..P10 of SURVEY.md section 8(c) *in
memory*, writes the patched text to a throw-away temp dir, compiles it unity-style
with oracle/ref_platform.cpp (a headless Linux platform layer: pthread work queue,
arena, timer, IsDown stub, C entry points) and leaves ONLY shared libraries in
oracle/_ref/ (git-ignored, but shipped to the GPU box by gpurun):

    oracle/_ref/libsrt_ref_w8.so   profile F: SIMD_WIDTH=8, -mavx2 -mfma (FMA dot)
    oracle/_ref/libsrt_ref_w4.so   profile I: SIMD_WIDTH=4, -msse4.2     (IEEE, no FMA)

No reference source is copied into the repository; the temp dir is removed.

Why patches are needed at all: HEAD only builds with emscripten,
and the x64 backend is written in a clang/MSVC dialect and lacks three members
that only wasm_math.h defines.  Every patch is either dialect (P1-P5), a missing
declaration (P6), the reference's own buffer overflow (P7), pinning unspecified
argument evaluation order to clang's (P8), or an opt-in run-time switch that
defaults to as-shipped behaviour (P9 IEEE NormalizeFast, P10 per-pixel RNG stream,
debug taps for primary-ray visibility).
"""
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("SRT_REFERENCE_DIR", "/root/reference")
OUT = os.path.join(HERE, "_ref")

# (file, old, new, expected_count)
PATCHES = [
    # ---- P3: g++ has no implicit __m128 <-> __m128i/__m128d conversions -------------
    ("x64_math.h",
     "Register = _mm_load_sd((f64 *)&Value); };",
     "Register = _mm_castpd_ps(_mm_load_sd((f64 *)&Value)); };", 1),
    ("x64_math.h",
     "_mm_store_sd((f64 *)&Result, Register);",
     "_mm_store_sd((f64 *)&Result, _mm_castps_pd(Register));", 1),
    ("x64_math.h",
     "    inline xmm(const __m128 &XMM) : Register(XMM) { };\n",
     "    inline xmm(const __m128 &XMM) : Register(XMM) { };\n"
     "    inline xmm(const __m128i &XMM) : IntRegister(XMM) { };\n"
     "    operator __m128i() const { return IntRegister; }\n", 1),
    ("x64_math.h",
     "    inline ymm(const __m256 &YMM) : Register(YMM) { };\n",
     "    inline ymm(const __m256 &YMM) : Register(YMM) { };\n"
     "    inline ymm(const __m256i &YMM) : IntRegister(YMM) { };\n"
     "    operator __m256i() const { return IntRegister; }\n", 1),
    # ---- P5: MS-style x87 inline asm -> libm, exactly as wasm_math.h:472-478 does ----
    ("x64_math.h",
     "\t__asm {\n\t\tfld float ptr [radians]\n\t\tfcos\n\t\tfstp float ptr [result]\n\t}\n",
     "\tresult = cosf(radians);\n", 1),
    ("x64_math.h",
     "\t__asm {\n\t\tfld float ptr [radians]\n\t\tfsin\n\t\tfstp float ptr [result]\n\t}\n",
     "\tresult = sinf(radians);\n", 1),
    # ---- P9 (opt-in at run time): IEEE NormalizeFast, as wasm_math.h:168-170 ---------
    ("x64_math.h",
     "inline v3 v3::NormalizeFast(const v3 &Value) {\n",
     "inline v3 v3::NormalizeFast(const v3 &Value) {\n"
     "    if (SrtRefIEEENormalizeFast) return v3::Normalize(Value);\n", 1),
    # ---- P6: members main.cpp uses but base.h never declares on the 8-wide types ----
    ("base.h",
     "    static inline u32 HorizontalMinIndex(const f32x &A);\n",
     "    static inline u32 HorizontalMinIndex(const f32x &A);\n"
     "    static inline u32 FindFirstIndex(const f32x8 &A, f32 Value);\n"
     "    inline f32 Extract(u32 Index);\n", 1),
    ("base.h",
     "    static inline void ConditionalMove(u32x8 *A, const u32x8 &B, const u32x8 &MoveMask);\n",
     "    static inline void ConditionalMove(u32x8 *A, const u32x8 &B, const u32x8 &MoveMask);\n"
     "    inline u32 Extract(u32 Index);\n", 1),
    ("main.cpp", "f32x(u32x4(-1))", "f32x(u32x(-1))", 1),
    # ---- P7: RTWeekendSpheres[482] is filled with 488 spheres (F4). Give it storage,
    #          keep Count == 482 and every array_len() derived from it unchanged. -------
    ("main.cpp",
     "static scalar_sphere RTWeekendSpheres[482];\n",
     "static scalar_sphere RTWeekendSpheresStorage[488];\n"
     "static scalar_sphere (&RTWeekendSpheres)[482] = *(scalar_sphere (*)[482])RTWeekendSpheresStorage;\n", 1),
    # ---- P8: pin v3(Rand(),Rand(),Rand()) to clang's order (1st draw -> x) -----------
    ("main.cpp",
     "v3(RandomState.RandomFloat(), RandomState.RandomFloat(), RandomState.RandomFloat())",
     "SrtRefRandV3(RandomState, -1.0f, 1.0f)", 2),
    ("main.cpp",
     "v3(\n\t\t\t\t\tRandomState.RandomFloat(0.0, 1.0),\n\t\t\t\t\tRandomState.RandomFloat(0.0, 1.0),\n"
     "\t\t\t\t\tRandomState.RandomFloat(0.0, 1.0)\n\t\t\t\t)",
     "SrtRefRandV3(RandomState, 0.0, 1.0)", 2),
    # ---- P10 (opt-in at run time): per-(sample,pixel) RNG stream ---------------------
    ("main.cpp",
     "        for (u32 x = TileLeft; x < TileRight; ++x) {\n\n            v3 Attenuation = 1.0f;\n",
     "        for (u32 x = TileLeft; x < TileRight; ++x) {\n"
     "            SRT_REF_PIXEL_PROLOGUE(x, y);\n            v3 Attenuation = 1.0f;\n", 2),
    # ---- debug taps: primary-ray t / primitive id (no effect unless buffers are set) --
    ("main.cpp",
     "                f32 MinValue = f32x::HorizontalMin(MinT);\n",
     "                f32 MinValue = f32x::HorizontalMin(MinT);\n"
     "                SRT_REF_TAP_T(i, x, y, MinValue);\n", 1),
    ("main.cpp",
     "                const u32 AbsoluteIndex = MaterialIndex.Extract(Index) * SIMD_WIDTH + Index;\n",
     "                const u32 AbsoluteIndex = MaterialIndex.Extract(Index) * SIMD_WIDTH + Index;\n"
     "                SRT_REF_TAP_PRIM(i, x, y, AbsoluteIndex);\n", 1),
    ("main.cpp",
     "                if (MinT == F32Max) {\n",
     "                SRT_REF_TAP_T(i, x, y, MinT);\n                if (MinT == F32Max) {\n", 1),
    ("main.cpp",
     "                const material &Material = Scene.ScalarSpheres[SphereIndex].Material;\n",
     "                const material &Material = Scene.ScalarSpheres[SphereIndex].Material;\n"
     "                SRT_REF_TAP_PRIM(i, x, y, SphereIndex);\n", 1),
]

COMMON = ["-std=gnu++17", "-O3", "-ffp-contract=off", "-fno-strict-aliasing", "-fPIC", "-shared",
          "-pthread", "-w"]
VARIANTS = {
    "w8": ["-mavx2", "-mfma", "-mbmi", "-DSIMD_WIDTH=8"],
    "w4": ["-msse4.2", "-mbmi", "-DSIMD_WIDTH=4"],
}


def patched_sources():
    texts = {}
    for name in ("base.h", "x64_math.h", "main.cpp"):
        with open(os.path.join(REF, name), "r", encoding="utf-8") as f:
            texts[name] = f.read()
    for name, old, new, count in PATCHES:
        got = texts[name].count(old)
        if got != count:
            raise SystemExit(f"patch anchor mismatch in {name}: expected {count}, found {got}: {old[:60]!r}")
        texts[name] = texts[name].replace(old, new)
    return texts


def build(verbose=True):
    if not os.path.isdir(REF):
        if verbose:
            print(f"[build_ref] {REF} absent: keeping prebuilt oracle/_ref as is")
        return False
    os.makedirs(OUT, exist_ok=True)
    tmp = tempfile.mkdtemp(prefix=".patched_", dir=OUT)   # inside oracle/_ref (git-ignored), removed below: nothing is written outside the repository
    try:
        for name, text in patched_sources().items():
            with open(os.path.join(tmp, name), "w", encoding="utf-8") as f:
                f.write(text)
        for tag, flags in VARIANTS.items():
            out = os.path.join(OUT, f"libsrt_ref_{tag}.so")
            cmd = ["g++", *COMMON, *flags, "-I", tmp, "-I", os.path.join(HERE, "shim"),
                   os.path.join(HERE, "ref_platform.cpp"), "-o", out, "-lm"]
            if verbose:
                print("[build_ref]", " ".join(cmd))
            subprocess.check_call(cmd)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return True


if __name__ == "__main__":
    sys.exit(0 if build() or True else 1)
