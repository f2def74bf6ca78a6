// srt_files.cpp -- the data formats either side of the hot path.
//
// The reference has no file I/O at all: scenes are hard-coded (main.cpp:93-268) and the
// accumulator lives in an arena (main.cpp:801-804).  These two formats make long renders
// restartable and let scenes beyond the five built-in ones reach srt_set_scene:
//
//   checkpoint (binary, little endian):  "SRTCKPT2" | u32 width | u32 height | u32 samples |
//       u32 scene_index | u64 stream_base | u32 lane_width | u32 fma_dot | u64 scene_hash | f32 accum[W*H*4]
//       `samples` = PreviousRayCount (main.cpp:7): the next frame renders sample index `samples`.
//       lane_width / fma_dot / scene_hash pin the arithmetic profile and the sphere table the samples were
//       rendered with: a resume under anything else would silently break the bit-identical guarantee.
//
//   scene (text, one record per line, '#' comments; %.9g round-trips every f32 exactly):
//       sky 0|1
//       lookat x y z
//       camera distance x_angle y_height
//       sphere x y z radius  cr cg cb  er eg eb  specular ior
#include "srt_app.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

static const char kMagic[8] = {'S', 'R', 'T', 'C', 'K', 'P', 'T', '2'};

bool SaveCheckpoint(const char *Path, const checkpoint_header &Header, const f32 *Accum) {
    FILE *F = fopen(Path, "wb");
    if (!F) return false;
    bool Ok = fwrite(kMagic, 1, 8, F) == 8 && fwrite(&Header, sizeof(Header), 1, F) == 1;
    const size_t N = (size_t)Header.Width * Header.Height * 4;
    Ok = Ok && fwrite(Accum, sizeof(f32), N, F) == N;
    return fclose(F) == 0 && Ok;
}

bool LoadCheckpoint(const char *Path, checkpoint_header *Header, std::vector<f32> *Accum) {
    FILE *F = fopen(Path, "rb");
    if (!F) return false;
    char Magic[8];
    bool Ok = fread(Magic, 1, 8, F) == 8 && memcmp(Magic, kMagic, 8) == 0 && fread(Header, sizeof(*Header), 1, F) == 1;
    if (Ok && (Header->Width == 0 || Header->Height == 0 || (u64)Header->Width * Header->Height > 0x7FFFFFFFull)) Ok = false;
    if (Ok) {
        // the payload length must match the header BEFORE anything is allocated (a damaged header could ask for 32 GiB)
        const size_t N = (size_t)Header->Width * Header->Height * 4;
        const long Here = ftell(F);
        Ok = Here >= 0 && fseek(F, 0, SEEK_END) == 0;
        const long End = Ok ? ftell(F) : -1;
        Ok = Ok && End >= Here && (u64)(End - Here) == (u64)N * sizeof(f32) && fseek(F, Here, SEEK_SET) == 0;
        if (Ok) {
            Accum->resize(N);
            Ok = fread(Accum->data(), sizeof(f32), N, F) == N;
        }
    }
    fclose(F);
    return Ok;
}

bool SaveSceneFile(const char *Path, const scene &S) {
    FILE *F = fopen(Path, "w");
    if (!F) return false;
    fprintf(F, "# SIMD-Ray-Tracer scene: %u spheres\nsky %d\nlookat %.9g %.9g %.9g\ncamera %.9g %.9g %.9g\n", S.SphereCount,
            S.UseSkyColor ? 1 : 0, S.LookAt.x, S.LookAt.y, S.LookAt.z, S.DefaultDistanceFromLookAt, S.DefaultXAngle, S.DefaultYHeight);
    for (u32 i = 0; i < S.SphereCount; ++i) {
        const scalar_sphere &P = S.ScalarSpheres[i];
        const material &M = P.Material;
        fprintf(F, "sphere %.9g %.9g %.9g %.9g  %.9g %.9g %.9g  %.9g %.9g %.9g  %.9g %.9g\n", P.Position.x, P.Position.y, P.Position.z,
                P.Radius, M.Color.x, M.Color.y, M.Color.z, M.Emissive.x, M.Emissive.y, M.Emissive.z, M.Specular, M.IndexOfRefraction);
    }
    return fclose(F) == 0;
}

// Allocates Out->ScalarSpheres with malloc (like BuildScene); Out->Name points to a static string.
bool LoadSceneFile(const char *Path, scene *Out) {
    FILE *F = fopen(Path, "r");
    if (!F) return false;
    std::vector<scalar_sphere> Spheres;
    memset(Out, 0, sizeof(*Out));
    Out->DefaultDistanceFromLookAt = 1.0f;
    char Line[1024];
    bool Ok = true;
    while (Ok && fgets(Line, sizeof(Line), F)) {
        char *P = Line;
        while (*P == ' ' || *P == '\t') ++P;
        if (*P == '#' || *P == '\n' || *P == '\r' || *P == 0) continue;
        int Sky;
        f32 a[12];
        if (sscanf(P, "sky %d", &Sky) == 1) Out->UseSkyColor = Sky != 0;
        else if (sscanf(P, "lookat %f %f %f", &a[0], &a[1], &a[2]) == 3) { Out->LookAt.x = a[0]; Out->LookAt.y = a[1]; Out->LookAt.z = a[2]; }
        else if (sscanf(P, "camera %f %f %f", &a[0], &a[1], &a[2]) == 3) { Out->DefaultDistanceFromLookAt = a[0]; Out->DefaultXAngle = a[1]; Out->DefaultYHeight = a[2]; }
        else if (sscanf(P, "sphere %f %f %f %f %f %f %f %f %f %f %f %f", &a[0], &a[1], &a[2], &a[3], &a[4], &a[5], &a[6], &a[7], &a[8], &a[9], &a[10], &a[11]) == 12) {
            scalar_sphere S;
            memset(&S, 0, sizeof(S));
            S.Position.x = a[0]; S.Position.y = a[1]; S.Position.z = a[2]; S.Radius = a[3];
            S.Material.Color.x = a[4]; S.Material.Color.y = a[5]; S.Material.Color.z = a[6];
            S.Material.Emissive.x = a[7]; S.Material.Emissive.y = a[8]; S.Material.Emissive.z = a[9];
            S.Material.Specular = a[10]; S.Material.IndexOfRefraction = a[11];
            Spheres.push_back(S);
        } else Ok = false;
    }
    fclose(F);
    if (!Ok || Spheres.empty()) return false;
    Out->ScalarSpheres = (scalar_sphere *)malloc(Spheres.size() * sizeof(scalar_sphere));
    if (!Out->ScalarSpheres) return false;
    memcpy(Out->ScalarSpheres, Spheres.data(), Spheres.size() * sizeof(scalar_sphere));
    Out->SphereCount = (u32)Spheres.size();
    Out->Name = "scene file";
    return true;
}
