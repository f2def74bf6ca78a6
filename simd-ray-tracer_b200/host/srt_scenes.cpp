// srt_scenes.cpp -- scene tables for the host mirror.
//
// Scenes 0-2 reproduce the reference's built-in scenes bit for bit (tests compare the
// 80-byte sphere records with goldens dumped from the reference's own compiled code):
//   0 "RGB Glass"        main.cpp:171-191      1 "Floating Spheres"  main.cpp:96-167
//   2 "Ray Tracing in One Weekend" main.cpp:196-268 (482 of the 488 generated)
// Scenes 3-4 are the synthetic brute-force stress scenes of BASELINE.json configs 3 and 5
// (no counterpart in the reference).  All draws go through the reference's PCG stream
// (base.h:954-989) with first-draw-to-x ordering.
#include "srt_app.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

namespace {

constexpr f32 kWorldScale = 1.0f / 16.0f;            // main.cpp:56
constexpr f32 kPi = 3.14159265358979323846f;         // base.h:892

struct pcg_stream {
    u64 State;
    u32 Next() {                                      // base.h:954-963
        const u64 Old = State;
        State = Old * 6364136223846793005ULL + 1442695040888963407ULL;
        const u32 X = (u32)(Old >> 32) ^ (u32)Old;
        const u32 R = (u32)(Old >> 59);
        return (X >> R) | (X << ((32u - R) & 31u));
    }
    f32 Float(f32 Min, f32 Max) {                     // base.h:983-989
        const u32 N = Next();
        const f32 Inv = (f32)((f64)(Max - Min) / (f64)0xFFFFFFFFu);
        const f32 Scaled = (f32)N * Inv;
        return Scaled + Min;
    }
    f32 Signed() { return Float(-1.0f, 1.0f); }
    f32 Unit() { return Float(0.0f, 1.0f); }
};

struct rgb { f32 r, g, b; };

void Put(scalar_sphere *S, f32 X, f32 Y, f32 Z, f32 Radius, rgb Color, f32 Specular, f32 Ior, rgb Emissive, bool Scale) {
    memset(S, 0, sizeof(*S));
    const f32 K = kWorldScale;
    if (Scale) { S->Position = { X * K, Y * K, Z * K, 0.0f }; S->Radius = Radius * K; }   // main.cpp:58-62
    else { S->Position = { X, Y, Z, 0.0f }; S->Radius = Radius; }
    S->Material.Color = { Color.r, Color.g, Color.b, 0.0f };
    S->Material.Emissive = { Emissive.r, Emissive.g, Emissive.b, 0.0f };
    S->Material.Specular = Specular;
    S->Material.IndexOfRefraction = Ior;
}

f32 Length3(f32 X, f32 Y, f32 Z) { return sqrtf((X * X + Y * Y) + Z * Z); }              // x64_math.h:224-232

// ---- scene 0 -------------------------------------------------------------------------------
void BuildRgbGlass(scene *S) {
    static const struct { f32 p[3]; f32 r; rgb c; f32 ior; rgb e; } T[5] = {
        { { 0.0f, -258.0f, -15.0f }, 256.0f, { 0.2f, 0.2f, 0.2f }, 0.0f, { 0, 0, 0 } },
        { { 0.0f, 0.0f, -10.0f }, 2.0f, { 1.0f, 1.0f, 1.0f }, 1.5f, { 0, 0, 0 } },
        { { -4.0f, 1.0f, -15.0f }, 1.5f, { 1.0f, 0.0f, 0.0f }, 0.0f, { 8.0f, 0, 0 } },
        { { 0.0f, 1.0f, -15.0f }, 1.5f, { 1.0f, 0.0f, 0.0f }, 0.0f, { 0, 8.0f, 0 } },
        { { 4.0f, 1.0f, -15.0f }, 1.5f, { 1.0f, 0.0f, 0.0f }, 0.0f, { 0, 0, 8.0f } },
    };
    S->SphereCount = 5;
    S->ScalarSpheres = (scalar_sphere *)calloc(5, sizeof(scalar_sphere));
    for (int i = 0; i < 5; ++i) Put(S->ScalarSpheres + i, T[i].p[0], T[i].p[1], T[i].p[2], T[i].r, T[i].c, 0.0f, T[i].ior, T[i].e, true);
    S->LookAt = S->ScalarSpheres[1].Position;
    S->UseSkyColor = false;
    S->DefaultDistanceFromLookAt = 16.0f * kWorldScale;
    S->DefaultXAngle = (f32)((f64)kPi / 3.0);
    S->DefaultYHeight = 4.0f * kWorldScale;
    S->Name = "RGB Glass";
}

// ---- scene 1 -------------------------------------------------------------------------------
void BuildFloatingSpheres(scene *S) {
    constexpr int N = 256, NM = 28;
    pcg_stream Rng = { 0x29D7A0A514F22432ULL };
    struct { rgb c, e; f32 spec; } M[NM];
    for (int i = 0; i < NM; ++i) {
        rgb C;
        C.r = Rng.Float(0.15f, 1.0f);
        C.g = Rng.Float(0.1f, 0.75f);
        C.b = Rng.Float(0.15f, 1.0f);
        rgb E = { 0, 0, 0 };
        f32 Spec = 0.0f;
        if (Rng.Unit() < 0.125f) {
            const f32 K = Rng.Float(2.0f, 5.0f);
            E = { K * C.r, K * C.g, K * C.b };
        } else if (Rng.Unit() < 0.65f) {
            Spec = 1.0f;
        }
        M[i].c = C; M[i].e = E; M[i].spec = Spec;
    }
    S->SphereCount = N;
    S->ScalarSpheres = (scalar_sphere *)calloc(N, sizeof(scalar_sphere));
    scalar_sphere *O = S->ScalarSpheres;
    const f32 R0 = Rng.Float(2.0f, 8.0f);
    Put(O + 0, 1.0f, 0.0f, 0.0f, R0, M[0].c, M[0].spec, 0.0f, M[0].e, false);
    Put(O + 1, 8.0f, -1.0f, 8.0f, R0, M[0].c, M[0].spec, 0.0f, M[0].e, false);
    Put(O + 2, -20.0f, -4.0f, -20.0f, R0, M[0].c, M[0].spec, 0.0f, M[0].e, false);
    for (int i = 3; i < N; ++i) {
        const f32 Vx = Rng.Signed(), Vy = Rng.Signed(), Vz = Rng.Signed();
        f32 Nx = 0, Ny = 0, Nz = 0;
        const f32 L2 = (Vx * Vx + Vy * Vy) + Vz * Vz;                 // v3::Normalize x64_math.h:234-245
        if (L2 > 1e-4f) { const f32 L = sqrtf(L2); Nx = Vx / L; Ny = Vy / L; Nz = Vz / L; }
        const scalar_sphere &Anchor = O[i - 3];
        const f32 Radius = Rng.Float(1.0f, 4.0f);
        const f32 Reach = Rng.Float(1.0f, 8.0f) + Radius + Anchor.Radius;
        const auto &Mat = M[i % NM];
        Put(O + i, Anchor.Position.x + Nx * Reach, Anchor.Position.y + Ny * Reach, Anchor.Position.z + Nz * Reach,
            Radius, Mat.c, Mat.spec, 0.0f, Mat.e, false);
    }
    for (int i = 0; i < N; ++i) {                                      // main.cpp:156-162
        O[i].Radius *= kWorldScale;
        O[i].Position.x *= kWorldScale; O[i].Position.y *= kWorldScale; O[i].Position.z *= kWorldScale;
    }
    S->LookAt = { 2.0f * kWorldScale, 0.0f * kWorldScale, 2.0f * kWorldScale, 0.0f };
    S->UseSkyColor = false;
    S->DefaultDistanceFromLookAt = 48.0f * kWorldScale;
    S->DefaultXAngle = (f32)((f64)(kPi * 2.65f) / 2.0);
    S->DefaultYHeight = 0.0f;
    S->Name = "Floating Spheres";
}

// ---- scene 2 -------------------------------------------------------------------------------
void BuildOneWeekend(scene *S) {
    constexpr int Generated = 488, Rendered = 482;
    std::vector<scalar_sphere> All(Generated);
    const rgb Black = { 0, 0, 0 };
    Put(&All[0], 0, -1000, 0, 1000, { 0.5f, 0.5f, 0.5f }, 0.0f, 0.0f, Black, true);
    Put(&All[1], 0, 1, 0, 1, { 1.0f, 1.0f, 1.0f }, 0.0f, 1.5f, Black, true);
    Put(&All[2], -4, 1, 0, 1, { 0.4f, 0.2f, 0.1f }, 0.0f, 0.0f, Black, true);
    Put(&All[3], 4, 1, 0, 1, { 0.7f, 0.6f, 0.5f }, 1.0f, 0.0f, Black, true);
    pcg_stream Rng = { 0xCD46749A57ACB371ULL };
    int Index = 4;
    for (s32 i = -11; i < 11; ++i) {
        for (s32 j = -11; j < 11; ++j) {
            const f32 Pick = Rng.Unit();
            f32 Cx, Cz;
            const f32 Cy = 0.2f;
            bool Clear;
            do {                                                        // main.cpp:229-236
                Cx = (f32)i + Rng.Signed();
                Cz = (f32)j + Rng.Signed();
                Clear = (f64)Length3(Cx - 4.0f, Cy - 0.2f, Cz - 0.0f) > 0.9 &&
                        (f64)Length3(Cx - 0.0f, Cy - 0.2f, Cz - 0.0f) > 0.9 &&
                        (f64)Length3(Cx - -4.0f, Cy - 0.2f, Cz - 0.0f) > 0.9;
            } while (!Clear);
            rgb Color = Black;
            f32 Spec = 0.0f, Ior = 0.0f;
            if ((f64)Pick < 0.8) {
                Color.r = Rng.Unit(); Color.g = Rng.Unit(); Color.b = Rng.Unit();
            } else if ((f64)Pick < 0.95) {
                Color.r = Rng.Unit(); Color.g = Rng.Unit(); Color.b = Rng.Unit();
                Spec = Rng.Float(0.5f, 1.0f);
            } else {
                Color = { 1.0f, 1.0f, 1.0f };
                Ior = 1.5f;
            }
            Put(&All[Index++], Cx, Cy, Cz, 0.2f, Color, Spec, Ior, Black, true);
        }
    }
    S->SphereCount = Rendered;
    S->ScalarSpheres = (scalar_sphere *)calloc(Rendered, sizeof(scalar_sphere));
    memcpy(S->ScalarSpheres, All.data(), Rendered * sizeof(scalar_sphere));
    S->LookAt = S->ScalarSpheres[1].Position;
    S->UseSkyColor = true;
    S->DefaultDistanceFromLookAt = 12.0f * kWorldScale;
    S->DefaultXAngle = kPi / 8;
    S->DefaultYHeight = 2.0f * kWorldScale;
    S->Name = "Ray Tracing in One Weekend";
}

// ---- scenes 3/4: synthetic brute-force stress scenes (BASELINE.json configs 3 and 5) ------------
// Ground sphere + (Count-1) non-overlapping random spheres in a box centred on the origin,
// y in [0, BoxY]; materials drawn like main.cpp:110-131 (12.5 % emissive, else 65 % mirror).
// Non-overlap keeps the lane tie-break / sticky-inside quirks dormant, and
// makes the scene independent of them.  Rejection test uses a uniform grid.
void BuildSynthetic(scene *S, u32 Count, f32 BoxX, f32 BoxY, f32 BoxZ, f32 RMin, f32 RMax, u64 Seed, const char *Name) {
    pcg_stream Rng = { Seed };
    S->SphereCount = Count;
    S->ScalarSpheres = (scalar_sphere *)calloc(Count, sizeof(scalar_sphere));
    scalar_sphere *O = S->ScalarSpheres;
    const rgb Black = { 0, 0, 0 };
    const f32 GroundR = 1000.0f * kWorldScale;
    Put(O + 0, 0.0f, -GroundR, 0.0f, GroundR, { 0.5f, 0.5f, 0.5f }, 0.0f, 0.0f, Black, false);

    const f32 Cell = 2.0f * RMax;
    const int Gx = (int)(BoxX / Cell) + 1, Gy = (int)(BoxY / Cell) + 1, Gz = (int)(BoxZ / Cell) + 1;
    std::vector<std::vector<u32>> Grid((size_t)Gx * Gy * Gz);
    auto CellOf = [&](f32 V, f32 Lo, int G) { int c = (int)((V - Lo) / Cell); return c < 0 ? 0 : (c >= G ? G - 1 : c); };
    for (u32 i = 1; i < Count; ++i) {
        for (;;) {
            const f32 R = Rng.Float(RMin, RMax);
            const f32 X = Rng.Float(-0.5f * BoxX, 0.5f * BoxX);
            const f32 Y = Rng.Float(R, BoxY);
            const f32 Z = Rng.Float(-0.5f * BoxZ, 0.5f * BoxZ);
            const int cx = CellOf(X, -0.5f * BoxX, Gx), cy = CellOf(Y, 0.0f, Gy), cz = CellOf(Z, -0.5f * BoxZ, Gz);
            bool Free = true;
            for (int dz = -1; dz <= 1 && Free; ++dz) for (int dy = -1; dy <= 1 && Free; ++dy) for (int dx = -1; dx <= 1 && Free; ++dx) {
                const int nx = cx + dx, ny = cy + dy, nz = cz + dz;
                if (nx < 0 || ny < 0 || nz < 0 || nx >= Gx || ny >= Gy || nz >= Gz) continue;
                for (u32 j : Grid[((size_t)nz * Gy + ny) * Gx + nx]) {
                    const f32 D = Length3(X - O[j].Position.x, Y - O[j].Position.y, Z - O[j].Position.z);
                    if (D < (R + O[j].Radius) * 1.02f) { Free = false; break; }
                }
            }
            if (!Free) continue;
            rgb C = { Rng.Float(0.15f, 1.0f), 0.0f, 0.0f };
            C.g = Rng.Float(0.1f, 0.75f);
            C.b = Rng.Float(0.15f, 1.0f);
            rgb E = Black;
            f32 Spec = 0.0f;
            if (Rng.Unit() < 0.125f) { const f32 K = Rng.Float(2.0f, 5.0f); E = { K * C.r, K * C.g, K * C.b }; }
            else if (Rng.Unit() < 0.65f) Spec = 1.0f;
            Put(O + i, X, Y, Z, R, C, Spec, 0.0f, E, false);
            Grid[((size_t)cz * Gy + cy) * Gx + cx].push_back(i);
            break;
        }
    }
    S->LookAt = { 0.0f, 0.25f * BoxY, 0.0f, 0.0f };
    S->UseSkyColor = true;
    S->DefaultDistanceFromLookAt = 0.8f * BoxX;
    S->DefaultXAngle = kPi / 5;
    S->DefaultYHeight = 0.2f * BoxX;
    S->Name = Name;
}

} // namespace

// ---- scene registry (the reference's static Scenes[], main.cpp:53-54) and camera ------------------
namespace {
scene g_Scenes[SceneCount];
bool g_SceneBuilt[SceneCount];
u32 g_SceneGenerations[SceneCount];

struct vec { f32 x, y, z; };
vec Sub(vec a, vec b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
vec Cross(vec a, vec b) { return { a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x }; }
vec Normalize(vec a) {                                                   // x64_math.h:234-245
    const f32 L2 = (a.x * a.x + a.y * a.y) + a.z * a.z;
    if (!(L2 > 1e-4f)) return { 0, 0, 0 };
    const f32 L = sqrtf(L2);
    return { a.x / L, a.y / L, a.z / L };
}
} // namespace

const scene *GetScene(u32 SceneIndex) {
    if (SceneIndex >= SceneCount) return nullptr;
    if (!g_SceneBuilt[SceneIndex]) {      // the two synthetic scenes are built on first use
        BuildScene(SceneIndex, &g_Scenes[SceneIndex]);
        g_SceneBuilt[SceneIndex] = true;
    }
    return &g_Scenes[SceneIndex];
}

u32 GetSceneGeneration(u32 SceneIndex) { return SceneIndex < SceneCount ? g_SceneGenerations[SceneIndex] : 0u; }

// Replace a scene slot by a caller-provided scene (copied).  A slot that is currently uploaded is
// re-uploaded by the next OnRender (which resets the accumulation, as a scene switch does at main.cpp:718-727).
bool SetScene(u32 SceneIndex, const scene &Scene) {
    if (SceneIndex >= SceneCount || !Scene.ScalarSpheres || Scene.SphereCount == 0) return false;
    scalar_sphere *Copy = (scalar_sphere *)malloc(sizeof(scalar_sphere) * Scene.SphereCount);
    if (!Copy) return false;
    memcpy(Copy, Scene.ScalarSpheres, sizeof(scalar_sphere) * Scene.SphereCount);
    if (g_SceneBuilt[SceneIndex]) free(g_Scenes[SceneIndex].ScalarSpheres);
    g_Scenes[SceneIndex] = Scene;
    g_Scenes[SceneIndex].ScalarSpheres = Copy;
    g_SceneBuilt[SceneIndex] = true;
    g_SceneGenerations[SceneIndex] += 1;
    return true;
}

// Drop a replaced slot: the next GetScene(SceneIndex) rebuilds the built-in / synthetic scene.
void ResetScene(u32 SceneIndex) {
    if (SceneIndex >= SceneCount || !g_SceneBuilt[SceneIndex]) return;
    free(g_Scenes[SceneIndex].ScalarSpheres);
    g_Scenes[SceneIndex].ScalarSpheres = nullptr;
    g_SceneBuilt[SceneIndex] = false;
    g_SceneGenerations[SceneIndex] += 1;
}

// FNV-1a over the sphere table and the sky flag: ties a checkpoint to the scene it was rendered from.
u64 HashScene(const scene &Scene) {
    u64 H = 0xCBF29CE484222325ULL;
    const u8 *P = (const u8 *)Scene.ScalarSpheres;
    for (size_t i = 0; i < sizeof(scalar_sphere) * (size_t)Scene.SphereCount; ++i) { H ^= P[i]; H *= 0x100000001B3ULL; }
    H ^= Scene.UseSkyColor ? 1u : 0u; H *= 0x100000001B3ULL;
    return H ? H : 1;
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

void BuildScene(u32 SceneIndex, scene *Out) {
    memset(Out, 0, sizeof(*Out));
    switch (SceneIndex) {
        case 0: BuildRgbGlass(Out); break;
        case 1: BuildFloatingSpheres(Out); break;
        case 2: BuildOneWeekend(Out); break;
        case SyntheticScene1K: BuildSynthetic(Out, 1024, 24.0f, 6.0f, 24.0f, 0.05f, 0.35f, 0x5EED0001ULL, "Synthetic 1,024 spheres"); break;
        case SyntheticScene64K: BuildSynthetic(Out, 65536, 96.0f, 8.0f, 96.0f, 0.02f, 0.12f, 0x5EED0005ULL, "Synthetic 65,536 spheres"); break;
        default: break;
    }
}
