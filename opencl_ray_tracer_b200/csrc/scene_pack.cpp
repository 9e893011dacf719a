// scene_pack.cpp -- see scene_pack.h.  Pure host C++ (double precision), no CUDA.
#include "scene_pack.h"

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
inline double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
inline double dot4(const double* a, const double* b) {
    return a[0] * b[0] + a[1] * b[1] + a[2] * b[2] + a[3] * b[3];
}
inline void cross3(const double* a, const double* b, double* o) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}
const double kPow40 = 1099511627776.0;  // 2^40
const double kPow30 = 1073741824.0;     // 2^30
}  // namespace

bool box_from_vertices(const float* v, Box* out) {
    // One representative vertex per corner (sign triple), then least-squares centre / half-axes.
    double corner[8][3];
    bool seen[8] = {false, false, false, false, false, false, false, false};
    for (int j = 0; j < 36; ++j) {
        const signed char* s = kCubeVertexSigns[j];
        const int idx = (s[0] > 0 ? 1 : 0) | (s[1] > 0 ? 2 : 0) | (s[2] > 0 ? 4 : 0);
        if (!seen[idx]) {
            seen[idx] = true;
            for (int k = 0; k < 3; ++k) corner[idx][k] = (double)v[4 * j + k];
        }
    }
    double maxabs = 0.0;
    for (int k = 0; k < 3; ++k) {
        out->c[k] = 0.0;
        for (int a = 0; a < 3; ++a) out->a[a][k] = 0.0;
    }
    for (int idx = 0; idx < 8; ++idx) {
        const double sx = (idx & 1) ? 1.0 : -1.0, sy = (idx & 2) ? 1.0 : -1.0, sz = (idx & 4) ? 1.0 : -1.0;
        for (int k = 0; k < 3; ++k) {
            const double p = corner[idx][k];
            out->c[k] += p * 0.125;
            out->a[0][k] += sx * p * 0.125;
            out->a[1][k] += sy * p * 0.125;
            out->a[2][k] += sz * p * 0.125;
            if (std::fabs(p) > maxabs) maxabs = std::fabs(p);
        }
    }
    double ext = 0.0;
    for (int a = 0; a < 3; ++a) ext += std::sqrt(dot3(out->a[a], out->a[a]));
    if (!std::isfinite(maxabs) || !std::isfinite(ext)) return false;
    // every one of the 36 vertices must sit on its corner (float rounding of the transforms only)
    const double tol = 4e-6 * maxabs + 1e-5 * ext + 1e-30;
    for (int j = 0; j < 36; ++j) {
        const signed char* s = kCubeVertexSigns[j];
        for (int k = 0; k < 3; ++k) {
            const double want = out->c[k] + s[0] * out->a[0][k] + s[1] * out->a[1][k] + s[2] * out->a[2][k];
            if (std::fabs((double)v[4 * j + k] - want) > tol) return false;
        }
    }
    double cr[3];
    cross3(out->a[1], out->a[2], cr);
    const double det = dot3(out->a[0], cr);
    const double n0 = std::sqrt(dot3(out->a[0], out->a[0])), n1 = std::sqrt(dot3(out->a[1], out->a[1])),
                 n2 = std::sqrt(dot3(out->a[2], out->a[2]));
    out->degenerate = !(std::fabs(det) > 1e-12 * n0 * n1 * n2) || !(std::fabs(det) > 1e-300);
    return true;
}

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
    double A[4];
    for (int k = 0; k < 4; ++k) A[k] = (double)centre[k] - cf.o0[k];
    const double tau0 = dot4(A, cf.dir);
    const double bx = dot4(A, cf.du) - tau0 * cf.tau_x;
    const double by = dot4(A, cf.dv) - tau0 * cf.tau_y;
    const double r2 = (double)(radius * radius);  // float-rounded like MainState.cpp:312
    const double c0 = r2 - dot4(A, A) + tau0 * tau0;
    const double det = cf.quu * cf.qvv - cf.quv * cf.quv;
    const double px = (cf.qvv * bx - cf.quv * by) / det;
    const double py = (cf.quu * by - cf.quv * bx) / det;
    const double rho = c0 + bx * px + by * py;
    out->uc = (float)(cf.l11 * px + cf.l12 * py);
    out->vc = (float)(cf.l22 * py);
    out->rho = (float)rho;
    out->tau0 = (float)tau0;
}

void pack_box(const CameraFrame& cf, const Box& box, RtbSlab out[3]) {
    if (box.degenerate) {  // never hit: tn = +1e30 > tf = -1e30
        out[0] = {1e30f, 0.f, 0.f, 0.f};
        out[1] = {-1e30f, 0.f, 0.f, 0.f};
        out[2] = {-1e30f, 0.f, 0.f, 0.f};
        return;
    }
    double cr[3][3];
    cross3(box.a[1], box.a[2], cr[0]);
    cross3(box.a[2], box.a[0], cr[1]);
    cross3(box.a[0], box.a[1], cr[2]);
    const double det = dot3(box.a[0], cr[0]);
    double rel[3];
    for (int k = 0; k < 3; ++k) rel[k] = cf.o0[k] - box.c[k];
    for (int i = 0; i < 3; ++i) {
        double m[3];
        for (int k = 0; k < 3; ++k) m[k] = cr[i][k] / det;  // row i of [a0 a1 a2]^-1
        const double q0 = dot3(m, rel), qx = dot3(m, cf.du), qy = dot3(m, cf.dv);
        const double dl = dot3(m, cf.dir);
        double alpha, bx, by, width;
        // The reference skips a triangle when |det_tri| < 1e-6 (MainState.cpp:276-277); for the
        // two faces of slab i that determinant is +-4*det*dl.
        if (std::fabs(4.0 * det * dl) < 0.000001) {
            // Ray parallel to the slab: it constrains only the pixel footprint, |q_i(x,y)| <= 1.
            // Expressed as near/far planes scaled by powers of two so that integer-aligned
            // edges stay exact in float (inclusive edges, MainState.cpp:282-291).
            const double minf = std::fmax(std::fabs(m[0]), std::fmax(std::fabs(m[1]), std::fabs(m[2])));
            const double g = 1.0 / minf;
            alpha = kPow40 * (-g - g * q0) - kPow30;
            bx = -kPow40 * (g * qx);
            by = -kPow40 * (g * qy);
            width = 2.0 * kPow40 * g + 2.0 * kPow30;
        } else {
            const double inv = 1.0 / dl;
            alpha = -q0 * inv - std::fabs(inv);
            bx = -qx * inv;
            by = -qy * inv;
            width = 2.0 * std::fabs(inv);
        }
        // (x, y) -> (u, v):  x = (u - l12*v/l22)/l11,  y = v/l22
        out[i].alpha = (float)alpha;
        out[i].beta = (float)(bx / cf.l11);
        out[i].gamma = (float)(by / cf.l22 - bx * cf.l12 / (cf.l11 * cf.l22));
        out[i].width = (float)width;
    }
}

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
