// b200rt_camera_perspective: the matrices a reference PerspectiveCamera holds, computed with the reference's own float
// operations so that the device's camera rays (wf_core.cuh camera_ray_ref) equal the reference's bit for bit.
//
//   Transform = (m, mInv) pairs; operator* multiplies m1.m * m2.m and m2.mInv * m1.mInv (src/core/transform.cpp);
//   Inverse(Transform) swaps the pair; Transform(Matrix4x4) inverts with Gauss-Jordan (transform.cpp:82-136);
//   LookAt (transform.cpp:203-236), Perspective (:303-311), Scale / Translate (:141-153);
//   pbrtCamera: CameraToWorld = Inverse(CTM) (src/core/api.cpp); ProjectiveCamera (src/core/camera.h:87-108):
//   ScreenToRaster = Scale(res.x, res.y, 1) * Scale(1 / (sw.max.x - sw.min.x), 1 / (sw.min.y - sw.max.y), 1) * Translate(-sw.min.x, -sw.max.y, 0),
//   RasterToCamera = Inverse(CameraToScreen) * Inverse(ScreenToRaster); the default screen window is the frame aspect's
//   (cameras/perspective.cpp:238-252).
#include <cmath>
#include <cstring>
#include <utility>
#include "../../include/b200rt_wavefront.h"
#include "wf_core.cuh"

namespace b200rt { int set_error(int code, const char *fmt, ...); }
using namespace b200rt;

namespace {

struct M4 { float m[4][4]; };
struct Xf { M4 m, inv; };

M4 identity() { M4 r; std::memset(&r, 0, sizeof(r)); for (int i = 0; i < 4; ++i) r.m[i][i] = 1.f; return r; }
M4 mul(const M4 &a, const M4 &b) {   // Matrix4x4::Mul (transform.h:86-93)
    M4 r;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j)
            r.m[i][j] = a.m[i][0] * b.m[0][j] + a.m[i][1] * b.m[1][j] + a.m[i][2] * b.m[2][j] + a.m[i][3] * b.m[3][j];
    return r;
}
bool inverse(const M4 &m, M4 *out) {   // Inverse(const Matrix4x4 &) (transform.cpp:82-136): Gauss-Jordan with full pivoting
    int indxc[4], indxr[4], ipiv[4] = {0, 0, 0, 0};
    float minv[4][4];
    std::memcpy(minv, m.m, sizeof(minv));
    for (int i = 0; i < 4; i++) {
        int irow = 0, icol = 0;
        float big = 0.f;
        for (int j = 0; j < 4; j++) {
            if (ipiv[j] != 1) {
                for (int k = 0; k < 4; k++) {
                    if (ipiv[k] == 0) {
                        if (std::abs(minv[j][k]) >= big) { big = std::abs(minv[j][k]); irow = j; icol = k; }
                    } else if (ipiv[k] > 1) return false;
                }
            }
        }
        ++ipiv[icol];
        if (irow != icol)
            for (int k = 0; k < 4; ++k) std::swap(minv[irow][k], minv[icol][k]);
        indxr[i] = irow; indxc[i] = icol;
        if (minv[icol][icol] == 0.f) return false;
        const float pivinv = (float)(1. / (double)minv[icol][icol]);   // `Float pivinv = 1. / minv[icol][icol];`
        minv[icol][icol] = 1.f;
        for (int j = 0; j < 4; j++) minv[icol][j] *= pivinv;
        for (int j = 0; j < 4; j++) {
            if (j != icol) {
                const float save = minv[j][icol];
                minv[j][icol] = 0;
                for (int k = 0; k < 4; k++) minv[j][k] -= minv[icol][k] * save;
            }
        }
    }
    for (int j = 3; j >= 0; j--)
        if (indxr[j] != indxc[j])
            for (int k = 0; k < 4; k++) std::swap(minv[k][indxr[j]], minv[k][indxc[j]]);
    std::memcpy(out->m, minv, sizeof(minv));
    return true;
}
Xf xmul(const Xf &a, const Xf &b) { Xf r; r.m = mul(a.m, b.m); r.inv = mul(b.inv, a.inv); return r; }
Xf xinverse(const Xf &a) { Xf r; r.m = a.inv; r.inv = a.m; return r; }
Xf scale(float x, float y, float z) {
    Xf r; r.m = identity(); r.inv = identity();
    r.m.m[0][0] = x; r.m.m[1][1] = y; r.m.m[2][2] = z;
    r.inv.m[0][0] = 1 / x; r.inv.m[1][1] = 1 / y; r.inv.m[2][2] = 1 / z;
    return r;
}
Xf translate(float x, float y, float z) {
    Xf r; r.m = identity(); r.inv = identity();
    r.m.m[0][3] = x; r.m.m[1][3] = y; r.m.m[2][3] = z;
    r.inv.m[0][3] = -x; r.inv.m[1][3] = -y; r.inv.m[2][3] = -z;
    return r;
}
bool look_at(const float *e, Xf *out) {   // LookAt (transform.cpp:203-236) -> Transform(Inverse(cameraToWorld), cameraToWorld)
    M4 c2w = identity();
    c2w.m[0][3] = e[0]; c2w.m[1][3] = e[1]; c2w.m[2][3] = e[2]; c2w.m[3][3] = 1;
    const V3 pos = v3(e[0], e[1], e[2]), look = v3(e[3], e[4], e[5]), up = v3(e[6], e[7], e[8]);
    const V3 dir = vnormalize(vsub(look, pos));
    if (vlen(vcross(vnormalize(up), dir)) == 0) return false;
    const V3 right = vnormalize(vcross(vnormalize(up), dir));
    const V3 newUp = vcross(dir, right);
    c2w.m[0][0] = right.x; c2w.m[1][0] = right.y; c2w.m[2][0] = right.z; c2w.m[3][0] = 0.f;
    c2w.m[0][1] = newUp.x; c2w.m[1][1] = newUp.y; c2w.m[2][1] = newUp.z; c2w.m[3][1] = 0.f;
    c2w.m[0][2] = dir.x; c2w.m[1][2] = dir.y; c2w.m[2][2] = dir.z; c2w.m[3][2] = 0.f;
    if (!inverse(c2w, &out->m)) return false;
    out->inv = c2w;
    return true;
}

}  // namespace

extern "C" int b200rt_camera_perspective(const float *lookat9, const float *w2c16, float fov, uint32_t xres, uint32_t yres,
                                         b200rt_render_desc *desc) {
    if (!desc || (!lookat9 && !w2c16) || xres == 0 || yres == 0 || !(fov > 0.f && fov < 180.f))
        return set_error(B200RT_EINVAL, "b200rt_camera_perspective: bad argument");
    // the current transformation matrix at `Camera`: Identity * LookAt (api.cpp pbrtLookAt: ctm = ctm * LookAt) or the
    // matrix of `Transform [...]` (api.cpp pbrtTransform: Transform(Matrix4x4) = (m, Inverse(m)))
    Xf ctm;
    if (w2c16) {
        std::memcpy(ctm.m.m, w2c16, 64);
        if (!inverse(ctm.m, &ctm.inv)) return set_error(B200RT_EINVAL, "b200rt_camera_perspective: singular camera matrix");
    } else {
        Xf la;
        if (!look_at(lookat9, &la)) return set_error(B200RT_EINVAL, "b200rt_camera_perspective: \"up\" and the viewing direction are parallel");
        Xf id; id.m = identity(); id.inv = identity();
        ctm = xmul(id, la);
    }
    const Xf c2w = xinverse(ctm);
    // Perspective(fov, 1e-2f, 1000.f) (transform.cpp:303-311)
    const float n = 1e-2f, f = 1000.f;
    Xf persp;
    persp.m = identity();
    persp.m.m[2][2] = f / (f - n); persp.m.m[2][3] = -f * n / (f - n);
    persp.m.m[3][2] = 1.f; persp.m.m[3][3] = 0.f;
    if (!inverse(persp.m, &persp.inv)) return set_error(B200RT_EINVAL, "b200rt_camera_perspective: singular projection");
    const float invTanAng = 1.f / std::tan(((3.14159265358979323846f / 180.f) * fov) / 2.f);
    const Xf c2s = xmul(scale(invTanAng, invTanAng, 1.f), persp);
    // screen window (cameras/perspective.cpp:238-252)
    const float frame = (float)xres / (float)yres;
    float sw[4];   // min.x, max.x, min.y, max.y
    if (frame > 1.f) { sw[0] = -frame; sw[1] = frame; sw[2] = -1.f; sw[3] = 1.f; }
    else { sw[0] = -1.f; sw[1] = 1.f; sw[2] = -1.f / frame; sw[3] = 1.f / frame; }
    const Xf s2r = xmul(xmul(scale((float)xres, (float)yres, 1.f), scale(1.f / (sw[1] - sw[0]), 1.f / (sw[2] - sw[3]), 1.f)),
                        translate(-sw[0], -sw[3], 0.f));
    const Xf r2c = xmul(xinverse(c2s), xinverse(s2r));
    std::memcpy(desc->raster_to_camera, r2c.m.m, 64);
    std::memcpy(desc->camera_to_world, c2w.m.m, 64);
    desc->camera_exact = 1;
    // the legacy frame (eye, dir, right, up) of the same camera, for the callers that still read it
    b200rt_camera &c = desc->camera;
    c.xres = xres; c.yres = yres;
    const float t = std::tan(((3.14159265358979323846f / 180.f) * fov) / 2.f);
    const float hx = frame > 1.f ? t * frame : t, hy = frame > 1.f ? t : t / frame;
    for (int i = 0; i < 3; ++i) {
        c.eye[i] = c2w.m.m[i][3];
        c.dir[i] = c2w.m.m[i][2];
        c.right[i] = c2w.m.m[i][0] * hx;
        c.up[i] = c2w.m.m[i][1] * hy;
    }
    return B200RT_OK;
}
