// scene_pack.cpp -- see scene_pack.h.  Pure host C++ (double precision), no CUDA.  The per-primitive
// functions are the shared host/device implementations of scene_pack_impl.h.
#include "scene_pack.h"
#include "scene_pack_impl.h"

#include <cmath>
#include <cstring>

namespace rtb {

// RayTrace/Cube.cpp:10-45 pushes these 36 corners of the unit cube (12 triangles, 3 each).
const signed char kCubeVertexSigns[36][3] = {
    {-1, -1, -1}, {-1, -1, +1}, {-1, +1, +1},   {+1, +1, -1}, {-1, -1, -1}, {-1, +1, -1},
    {+1, -1, +1}, {-1, -1, -1}, {+1, -1, -1},   {+1, +1, -1}, {+1, -1, -1}, {-1, -1, -1},
    {-1, -1, -1}, {-1, +1, +1}, {-1, +1, -1},   {+1, -1, +1}, {-1, -1, +1}, {-1, -1, -1},
    {-1, +1, +1}, {-1, -1, +1}, {+1, -1, +1},   {+1, +1, +1}, {+1, -1, -1}, {+1, +1, -1},
    {+1, -1, -1}, {+1, +1, +1}, {+1, -1, +1},   {+1, +1, +1}, {+1, +1, -1}, {-1, +1, -1},
    {+1, +1, +1}, {-1, +1, -1}, {-1, +1, +1},   {+1, +1, +1}, {-1, +1, +1}, {+1, -1, +1},
};

namespace {
using pk::cross3;
using pk::dot3;
using pk::dot4;
const double kPow40 = 1099511627776.0;  // 2^40
const double kPow30 = 1073741824.0;     // 2^30
}  // namespace

bool box_from_vertices(const float* v, Box* out) { return pk::box_from_vertices(v, out); }

bool camera_prepare(const float o0[4], const float du[4], const float dv[4], const float dir[4],
                    CameraFrame* cf) {
    for (int k = 0; k < 4; ++k) {
        cf->o0[k] = o0[k]; cf->du[k] = du[k]; cf->dv[k] = dv[k]; cf->dir[k] = dir[k];
        if (!std::isfinite(cf->o0[k]) || !std::isfinite(cf->du[k]) || !std::isfinite(cf->dv[k]) ||
            !std::isfinite(cf->dir[k]))
            return false;
    }
    if (dot3(cf->dir, cf->dir) <= 0.0) return false;
    cf->tau_x = dot4(cf->du, cf->dir);
    cf->tau_y = dot4(cf->dv, cf->dir);
    cf->quu = dot4(cf->du, cf->du) - cf->tau_x * cf->tau_x;
    cf->quv = dot4(cf->du, cf->dv) - cf->tau_x * cf->tau_y;
    cf->qvv = dot4(cf->dv, cf->dv) - cf->tau_y * cf->tau_y;
    const double scale = dot4(cf->du, cf->du) * dot4(cf->dv, cf->dv);
    const double det = cf->quu * cf->qvv - cf->quv * cf->quv;
    if (!(cf->quu > 0.0) || !(det > 1e-12 * scale)) return false;
    cf->l11 = std::sqrt(cf->quu);
    cf->l12 = cf->quv / cf->l11;
    cf->l22 = std::sqrt(cf->qvv - cf->l12 * cf->l12);
    if (!(cf->l22 > 0.0)) return false;
    std::memset(&cf->cam, 0, sizeof cf->cam);
    cf->cam.l11 = (float)cf->l11;
    cf->cam.l12 = (float)cf->l12;
    cf->cam.l22 = (float)cf->l22;
    // wpix = x*tau_x + y*tau_y rewritten in (u, v)
    const double tu = cf->tau_x / cf->l11;
    const double tv = cf->tau_y / cf->l22 - cf->tau_x * cf->l12 / (cf->l11 * cf->l22);
    cf->cam.tau_u = (float)tu;
    cf->cam.tau_v = (float)tv;
    return true;
}

void pack_sphere(const CameraFrame& cf, const float centre[4], float radius, RtbSphere* out) {
    pk::pack_sphere(cf, centre, radius, out);
}

void pack_box(const CameraFrame& cf, const Box& box, RtbSlab out[3]) { pk::pack_box(cf, box, out); }

// ---- conservative conic bounds -------------------------------------------------------------
FrameExtent frame_extent(const CameraFrame& cf, int width, int height) {
    // the kernel's own float constants: u = l11*x + l12*y, v = l22*y
    const double l11 = cf.cam.l11, l12 = cf.cam.l12, l22 = cf.cam.l22;
    const double xs[2] = {0.0, (double)(width > 0 ? width - 1 : 0)};
    const double ys[2] = {0.0, (double)(height > 0 ? height - 1 : 0)};
    FrameExtent fe;
    fe.umin = fe.vmin = 1e300;
    fe.umax = fe.vmax = -1e300;
    for (int i = 0; i < 2; ++i)
        for (int j = 0; j < 2; ++j) {
            const double u = l11 * xs[i] + l12 * ys[j], v = l22 * ys[j];
            fe.umin = std::fmin(fe.umin, u); fe.umax = std::fmax(fe.umax, u);
            fe.vmin = std::fmin(fe.vmin, v); fe.vmax = std::fmax(fe.vmax, v);
        }
    fe.pitch = std::fmin(std::fabs(l11), std::fabs(l22));
    const double grow = fe.pitch + 1e-5 * (std::fabs(fe.umin) + std::fabs(fe.umax) + std::fabs(fe.vmin) +
                                           std::fabs(fe.vmax));
    fe.umin -= grow; fe.umax += grow; fe.vmin -= grow; fe.vmax += grow;
    fe.uabs = std::fmax(std::fabs(fe.umin), std::fabs(fe.umax));
    fe.vabs = std::fmax(std::fabs(fe.vmin), std::fabs(fe.vmax));
    return fe;
}

void pack_bound_box(const FrameExtent& fe, const RtbSlab slabs[3], RtbBound* out) { pk::pack_bound_box(fe, slabs, out); }

void pack_bound_sphere(const FrameExtent& fe, const RtbSphere& sph, RtbBound* out) {
    pk::pack_bound_sphere(fe, sph, out);
}

void pack_row_entry(const RtbBound& b, float v, double uabs, float out[2]) { pk::pack_row_entry(b, v, uabs, out); }

void pack_box_world(const double dir3[3], const Box& box, float out[16]) {
    if (box.degenerate) {
        const float miss[16] = {0, 0, 0, 1e30f, 0, 0, 0, -1e30f, 0, 0, 0, -1e30f, 0, 0, 0, 0};
        std::memcpy(out, miss, sizeof miss);
        return;
    }
    double cr[3][3];
    cross3(box.a[1], box.a[2], cr[0]);
    cross3(box.a[2], box.a[0], cr[1]);
    cross3(box.a[0], box.a[1], cr[2]);
    const double det = dot3(box.a[0], cr[0]);
    for (int i = 0; i < 3; ++i) {
        double m[3];
        for (int k = 0; k < 3; ++k) m[k] = cr[i][k] / det;
        const double mc = dot3(m, box.c);  // q_i(O) = m.O - mc
        const double dl = dot3(m, dir3);
        double n[3], alpha, width;
        if (std::fabs(4.0 * det * dl) < 0.000001) {
            const double minf = std::fmax(std::fabs(m[0]), std::fmax(std::fabs(m[1]), std::fabs(m[2])));
            const double g = 1.0 / minf;
            for (int k = 0; k < 3; ++k) n[k] = -kPow40 * (g * m[k]);
            alpha = kPow40 * (-g + g * mc) - kPow30;
            width = 2.0 * kPow40 * g + 2.0 * kPow30;
        } else {
            const double inv = 1.0 / dl;
            for (int k = 0; k < 3; ++k) n[k] = -m[k] * inv;
            alpha = mc * inv - std::fabs(inv);
            width = 2.0 * std::fabs(inv);
        }
        out[4 * i + 0] = (float)n[0];
        out[4 * i + 1] = (float)n[1];
        out[4 * i + 2] = (float)n[2];
        out[4 * i + 3] = (float)alpha;
        out[12 + i] = (float)width;
    }
    out[15] = 0.0f;
}

}  // namespace rtb
