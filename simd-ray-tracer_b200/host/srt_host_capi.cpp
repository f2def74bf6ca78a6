// srt_host_capi.cpp -- C entry points over the host mirror's scene/camera logic, so the
// tests can compare it with goldens without a GPU (no backend call in here).
#include "srt_app.h"
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

}
