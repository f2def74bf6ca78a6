// srt_host_capi.cpp -- C entry points over the host mirror's scene/camera logic, so the
// tests can compare it with goldens without a GPU (no backend call in here).
#include "srt_app.h"
#include <stdlib.h>
#include <string.h>

extern "C" {

// Copies scene `index` (0-2 built-in, 3-4 synthetic) into `out` (80-byte records).
// header = LookAt.xyz, DefaultDistanceFromLookAt, DefaultXAngle, DefaultYHeight.
int srt_host_scene(unsigned index, void *out, unsigned capacity, int *use_sky, float *header) {
    const scene *S = GetScene(index);
    if (!S) return -1;
    if (use_sky) *use_sky = S->UseSkyColor ? 1 : 0;
    if (header) {
        header[0] = S->LookAt.x; header[1] = S->LookAt.y; header[2] = S->LookAt.z;
        header[3] = S->DefaultDistanceFromLookAt; header[4] = S->DefaultXAngle; header[5] = S->DefaultYHeight;
    }
    if (out) memcpy(out, S->ScalarSpheres, sizeof(scalar_sphere) * (S->SphereCount < capacity ? S->SphereCount : capacity));
    return (int)S->SphereCount;
}

void srt_host_camera(unsigned index, float distance, float x_angle, float y_height, unsigned w, unsigned h, srt_camera *out) {
    const scene *S = GetScene(index);
    if (S) ComputeCamera(*S, distance, x_angle, y_height, w, h, out);
}


// Scene text format round trip (srt_files.cpp): write slot `index` / load a file into slot `index`.
int srt_host_save_scene(unsigned index, const char *path) {
    const scene *S = GetScene(index);
    return S && SaveSceneFile(path, *S) ? 0 : -1;
}

int srt_host_load_scene(unsigned index, const char *path) {
    scene Loaded;
    if (!LoadSceneFile(path, &Loaded)) return -1;
    const bool Ok = SetScene(index, Loaded);
    free(Loaded.ScalarSpheres);
    return Ok ? 0 : -1;
}

void srt_host_reset_scene(unsigned index) { ResetScene(index); }

// Checkpoint container round trip.  header = {width, height, samples, scene_index, lane_width, fma_dot};
// the scene hash is taken from slot scene_index as it is now (0 if the slot does not exist).
unsigned long long srt_host_scene_hash(unsigned index) {
    const scene *S = GetScene(index);
    return S ? HashScene(*S) : 0ull;
}
int srt_host_save_checkpoint(const char *path, const unsigned *header, unsigned long long stream_base, const float *accum) {
    checkpoint_header H = { header[0], header[1], header[2], header[3], stream_base, header[4], header[5], srt_host_scene_hash(header[3]) };
    return SaveCheckpoint(path, H, accum) ? 0 : -1;
}

// accum == NULL: header only.  Returns -1 on a malformed file, -2 when capacity (floats) is too small.
int srt_host_load_checkpoint(const char *path, unsigned *header, unsigned long long *stream_base, float *accum, unsigned long long capacity) {
    checkpoint_header H;
    std::vector<f32> A;
    if (!LoadCheckpoint(path, &H, &A)) return -1;
    header[0] = H.Width; header[1] = H.Height; header[2] = H.Samples; header[3] = H.SceneIndex; header[4] = H.LaneWidth; header[5] = H.FmaDot;
    if (stream_base) *stream_base = H.StreamBase;
    if (accum) {
        if (capacity < A.size()) return -2;
        memcpy(accum, A.data(), A.size() * sizeof(f32));
    }
    return 0;
}

// Image writers of the headless driver (srt_files.cpp).  kind: 0 = by extension, 1 = PPM, 2 = PNG.
int srt_host_write_image(const char *path, const unsigned *rgba, unsigned w, unsigned h, int kind) {
    const bool Ok = kind == 1 ? WritePpm(path, rgba, w, h) : (kind == 2 ? WritePng(path, rgba, w, h) : WriteImage(path, rgba, w, h));
    return Ok ? 0 : -1;
}
int srt_host_write_pfm(const char *path, const float *accum, unsigned w, unsigned h) { return WritePfm(path, accum, w, h) ? 0 : -1; }

}
