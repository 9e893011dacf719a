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

namespace {

struct P2 { double u, v; };

const RtbBound kBoundAlways = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};  // q = 0: every ray passes
const RtbBound kBoundNever = {0.f, 0.f, 2.f, 0.f, 0.f, 0.f, 0.f, 0.f};   // w = 2, q = 4: no ray passes

// Keeps the part of the convex polygon `poly` (n vertices) with hu*u + hv*v + hc <= 0.
int clip_halfplane(const P2* poly, int n, double hu, double hv, double hc, P2* out) {
    int m = 0;
    for (int i = 0; i < n; ++i) {
        const P2 a = poly[i], b = poly[(i + 1) % n];
        const double fa = hu * a.u + hv * a.v + hc, fb = hu * b.u + hv * b.v + hc;
        if (fa <= 0.0) out[m++] = a;
        if ((fa <= 0.0) != (fb <= 0.0)) {
            const double t = fa / (fa - fb);
            out[m++] = {a.u + t * (b.u - a.u), a.v + t * (b.v - a.v)};
        }
    }
    return m;
}

// Ellipse  a*(u-cu)^2 + 2b*(u-cu)(v-cv) + d*(v-cv)^2 <= 1  ->  RtbBound.  Completing the square,
//   Q = a*[(u-cu) + (b/a)(v-cv)]^2 + (d - b^2/a)*(v-cv)^2,
// and q = s*Q with the largest s for which the fp32 evaluation (rtb_bound_q) of any point with
// Q <= 1 stays below 2 (error budget below).
void emit_conic(const FrameExtent& fe, double cu, double cv, double a, double b, double d, RtbBound* out) {
    if (!(a > 0.0) || !std::isfinite(a) || !std::isfinite(b) || !std::isfinite(d) || !std::isfinite(cu) ||
        !std::isfinite(cv)) {
        *out = kBoundAlways;
        return;
    }
    const double k = b / a;
    const double e0 = std::fmax(d - b * b / a, 0.0);
    const double U = fe.uabs, V = fe.vabs;
    const double eps22 = 1.0 / 4194304.0;
    double s = 2.0 * (1.0 - 1.0 / 512.0);
    for (int it = 0; it < 400; ++it, s *= 0.9) {
        const double sA = std::sqrt(s * a), k1 = sA * k, k0 = -sA * (cu + k * cv), e = s * e0;
        // |w_fp32 - w| <= ew: roundings of sA, k1, k0 (double -> float) and of the two fma
        const double ew = eps22 * (sA * U + std::fabs(k1) * V + std::fabs(k0));
        // |rowC_fp32 - rowC| <= ec: rounding of cv inside dv (rowC <= 2 there, so e*|dv| <= sqrt(2e)),
        // plus three relative roundings of the products
        const double ec = eps22 * (std::sqrt(2.0 * e) * (std::fabs(cv) + V) + 2.0);
        const double wmax = std::sqrt(s) + ew;
        const double worst = s * (1.0 + 1.0 / 1048576.0) + 2.0 * wmax * ew + ew * ew + ec;
        if (worst < 2.0 * (1.0 - 1.0 / 4096.0)) {
            const RtbBound r = {(float)sA, (float)k1, (float)k0, (float)e, (float)cv, 0.f, 0.f, 0.f};
            const bool ok = std::isfinite(r.sA) && std::isfinite(r.k1) && std::isfinite(r.k0) &&
                            std::isfinite(r.e) && std::isfinite(r.cv);
            *out = ok ? r : kBoundAlways;
            return;
        }
    }
    *out = kBoundAlways;
}

}  // namespace

void pack_bound_box(const FrameExtent& fe, const RtbSlab slabs[3], RtbBound* out) {
    double al[3], be[3], ga[3], wi[3], E[3];
    for (int k = 0; k < 3; ++k) {
        al[k] = slabs[k].alpha; be[k] = slabs[k].beta; ga[k] = slabs[k].gamma; wi[k] = slabs[k].width;
        if (!std::isfinite(al[k]) || !std::isfinite(be[k]) || !std::isfinite(ga[k]) || !std::isfinite(wi[k])) {
            *out = kBoundAlways;
            return;
        }
        // fp32 error of near_k / far_k as the kernel computes them (2 fma + 1 add), doubled
        E[k] = (std::fabs(al[k]) + std::fabs(be[k]) * fe.uabs + std::fabs(ga[k]) * fe.vabs + std::fabs(wi[k])) /
               4194304.0;
    }
    // A hit needs near_i <= far_j for every pair: three strips in (u,v); clip the frame box by them.
    P2 bufa[16], bufb[16];
    P2* poly = bufa;
    P2* next = bufb;
    int n = 4;
    poly[0] = {fe.umin, fe.vmin}; poly[1] = {fe.umax, fe.vmin};
    poly[2] = {fe.umax, fe.vmax}; poly[3] = {fe.umin, fe.vmax};
    for (int i = 0; i < 3 && n > 0; ++i)
        for (int j = 0; j < 3 && n > 0; ++j) {
            if (i == j) continue;
            double hu = be[i] - be[j], hv = ga[i] - ga[j], hc = al[i] - al[j] - wi[j] - (E[i] + E[j]);
            const double m = std::fmax(std::fabs(hu), std::fabs(hv));
            if (!(m * (fe.uabs + fe.vabs) > 1e-9 * std::fabs(hc)) || m == 0.0) {
                if (hc > 0.0) n = 0;  // constant and violated: the cube is never hit
                continue;
            }
            hu /= m; hv /= m; hc /= m;
            n = clip_halfplane(poly, n, hu, hv, hc, next);
            P2* t = poly; poly = next; next = t;
        }
    if (n <= 0) {
        *out = kBoundNever;
        return;
    }
    double cu = 0.0, cv = 0.0;
    for (int i = 0; i < n; ++i) { cu += poly[i].u; cv += poly[i].v; }
    cu /= n; cv /= n;
    // four symmetric helper points keep the fit non-degenerate (a clipped sliver, a single vertex)
    const double r0 = 0.05 * fe.pitch;
    poly[n] = {cu + r0, cv}; poly[n + 1] = {cu - r0, cv}; poly[n + 2] = {cu, cv + r0}; poly[n + 3] = {cu, cv - r0};
    const int nt = n + 4;
    double suu = 0.0, suv = 0.0, svv = 0.0;
    for (int i = 0; i < nt; ++i) {
        const double x = poly[i].u - cu, y = poly[i].v - cv;
        suu += x * x; suv += x * y; svv += y * y;
    }
    suu /= nt; suv /= nt; svv /= nt;
    const double det = suu * svv - suv * suv;
    if (!(det > 0.0) || !std::isfinite(det)) {
        *out = kBoundAlways;
        return;
    }
    const double iuu = svv / det, iuv = -suv / det, ivv = suu / det;
    double T = 0.0;
    for (int i = 0; i < nt; ++i) {
        const double x = poly[i].u - cu, y = poly[i].v - cv;
        T = std::fmax(T, iuu * x * x + 2.0 * iuv * x * y + ivv * y * y);
    }
    T *= 1.0 + 1e-6;
    emit_conic(fe, cu, cv, iuu / T, iuv / T, ivv / T, out);
}

void pack_bound_sphere(const FrameExtent& fe, const RtbSphere& sph, RtbBound* out) {
    const double rho = sph.rho, uc = sph.uc, vc = sph.vc;
    if (!std::isfinite(rho) || !std::isfinite(uc) || !std::isfinite(vc)) {
        *out = kBoundAlways;
        return;
    }
    if (rho < 0.0) {  // rtb_sphere_rad is then negative on every ray (two monotone roundings)
        *out = kBoundNever;
        return;
    }
    // rad >= 0 in fp32  =>  (uc-u)^2 + (vc-v)^2 <= rho*(1 + 3*2^-24 + ...)
    const double r0 = 0.05 * fe.pitch;
    const double T = rho * (1.0 + 1.0 / 1048576.0) + r0 * r0;
    emit_conic(fe, uc, vc, 1.0 / T, 0.0, 1.0 / T, out);
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
