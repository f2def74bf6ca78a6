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

// ---- images ----------------------------------------------------------------------------------------
// The reference presents `image` through a GL texture / an HTML canvas (win32.cpp, wasm.cpp:176-219); the
// headless driver writes it to a file instead.  Row 0 of the reference's image is the BOTTOM row (GL
// convention), so every writer below emits rows in reverse order.

bool WritePpm(const char *Path, const u32 *Pixels, u32 W, u32 H) {
    FILE *F = fopen(Path, "wb");
    if (!F) return false;
    fprintf(F, "P6\n%u %u\n255\n", W, H);
    std::vector<u8> Row((size_t)W * 3);
    bool Ok = true;
    for (u32 y = 0; y < H && Ok; ++y) {
        const u32 *Src = Pixels + (size_t)(H - 1 - y) * W;
        for (u32 x = 0; x < W; ++x) { Row[x * 3 + 0] = Src[x] & 0xFF; Row[x * 3 + 1] = (Src[x] >> 8) & 0xFF; Row[x * 3 + 2] = (Src[x] >> 16) & 0xFF; }
        Ok = fwrite(Row.data(), 1, Row.size(), F) == Row.size();
    }
    return (fclose(F) == 0) && Ok;
}

// PFM (portable float map): RGB f32, little endian, rows bottom-up -- the accumulator's own row order.
bool WritePfm(const char *Path, const f32 *Accum, u32 W, u32 H) {
    FILE *F = fopen(Path, "wb");
    if (!F) return false;
    fprintf(F, "PF\n%u %u\n-1.0\n", W, H);
    std::vector<f32> Row((size_t)W * 3);
    bool Ok = true;
    for (u32 y = 0; y < H && Ok; ++y) {
        const f32 *Src = Accum + (size_t)y * W * 4;
        for (u32 x = 0; x < W; ++x) { Row[x * 3 + 0] = Src[x * 4 + 0]; Row[x * 3 + 1] = Src[x * 4 + 1]; Row[x * 3 + 2] = Src[x * 4 + 2]; }
        Ok = fwrite(Row.data(), sizeof(f32), Row.size(), F) == Row.size();
    }
    return (fclose(F) == 0) && Ok;
}

// PNG, 8-bit RGB, filter 0 on every row, zlib stream of STORED deflate blocks (no compressor in this tree and no
// zlib dependency: the file is 0.02 % larger than the PPM and every PNG reader opens it).
namespace {
struct Crc32 {
    u32 Table[256];
    Crc32() {
        for (u32 n = 0; n < 256; ++n) {
            u32 c = n;
            for (int k = 0; k < 8; ++k) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            Table[n] = c;
        }
    }
    u32 Update(u32 Crc, const u8 *P, size_t N) const {
        for (size_t i = 0; i < N; ++i) Crc = Table[(Crc ^ P[i]) & 0xFF] ^ (Crc >> 8);
        return Crc;
    }
};
void PutBe32(std::vector<u8> &V, u32 X) { V.push_back((u8)(X >> 24)); V.push_back((u8)(X >> 16)); V.push_back((u8)(X >> 8)); V.push_back((u8)X); }
bool WriteChunk(FILE *F, const Crc32 &Crc, const char Type[4], const std::vector<u8> &Data) {
    std::vector<u8> Head, Tail;
    PutBe32(Head, (u32)Data.size());
    Head.insert(Head.end(), Type, Type + 4);
    u32 C = Crc.Update(0xFFFFFFFFu, Head.data() + 4, 4);
    C = Crc.Update(C, Data.data(), Data.size()) ^ 0xFFFFFFFFu;
    PutBe32(Tail, C);
    return fwrite(Head.data(), 1, 8, F) == 8 && (Data.empty() || fwrite(Data.data(), 1, Data.size(), F) == Data.size()) &&
           fwrite(Tail.data(), 1, 4, F) == 4;
}
} // namespace

bool WritePng(const char *Path, const u32 *Pixels, u32 W, u32 H) {
    if (!W || !H) return false;
    static const Crc32 Crc;
    const size_t Stride = (size_t)W * 3 + 1, RawSize = Stride * H;
    if (RawSize / 65535u * 5u + RawSize + 64u > 0x7FFFFFFFull) return false;      // one IDAT chunk: length field is 31 bits
    std::vector<u8> Raw(RawSize);
    for (u32 y = 0; y < H; ++y) {
        const u32 *Src = Pixels + (size_t)(H - 1 - y) * W;
        u8 *Dst = Raw.data() + Stride * y;
        *Dst++ = 0;                                                              // filter type None
        for (u32 x = 0; x < W; ++x) { *Dst++ = Src[x] & 0xFF; *Dst++ = (Src[x] >> 8) & 0xFF; *Dst++ = (Src[x] >> 16) & 0xFF; }
    }
    std::vector<u8> Z;
    Z.reserve(RawSize + RawSize / 65535u * 5u + 16u);
    Z.push_back(0x78); Z.push_back(0x01);                                        // zlib header: deflate, 32K window, no preset dictionary
    u32 A = 1, B = 0;                                                            // Adler-32 of the raw bytes
    for (size_t At = 0; At < RawSize;) {
        const size_t N = RawSize - At < 65535u ? RawSize - At : 65535u;
        Z.push_back(At + N == RawSize ? 1 : 0);                                  // BFINAL, BTYPE = 00 (stored)
        Z.push_back((u8)N); Z.push_back((u8)(N >> 8)); Z.push_back((u8)~N); Z.push_back((u8)(~N >> 8));
        Z.insert(Z.end(), Raw.begin() + At, Raw.begin() + At + N);
        for (size_t i = 0; i < N;) {                                             // 5552 bytes keep B below 2^32 between reductions
            const size_t M = N - i < 5552u ? N - i : 5552u;
            for (size_t k = 0; k < M; ++k) { A += Raw[At + i + k]; B += A; }
            A %= 65521u; B %= 65521u;
            i += M;
        }
        At += N;
    }
    PutBe32(Z, (B << 16) | A);
    FILE *F = fopen(Path, "wb");
    if (!F) return false;
    static const u8 Sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    std::vector<u8> Ihdr;
    PutBe32(Ihdr, W); PutBe32(Ihdr, H);
    Ihdr.push_back(8); Ihdr.push_back(2); Ihdr.push_back(0); Ihdr.push_back(0); Ihdr.push_back(0);   // 8 bit, truecolour, deflate, adaptive, no interlace
    const bool Ok = fwrite(Sig, 1, 8, F) == 8 && WriteChunk(F, Crc, "IHDR", Ihdr) && WriteChunk(F, Crc, "IDAT", Z) &&
                    WriteChunk(F, Crc, "IEND", std::vector<u8>());
    return (fclose(F) == 0) && Ok;
}

bool WriteImage(const char *Path, const u32 *Pixels, u32 W, u32 H) {
    const size_t L = strlen(Path);
    const bool Png = L >= 4 && (strcmp(Path + L - 4, ".png") == 0 || strcmp(Path + L - 4, ".PNG") == 0);
    return Png ? WritePng(Path, Pixels, W, H) : WritePpm(Path, Pixels, W, H);
}
