// scene_pack_impl.h -- the per-primitive part of the camera-space precompute (see scene_pack.h),
// written once for BOTH sides: the host functions of scene_pack.cpp (CPU tests, debug entries,
// rtb_render_rays) and the device pack kernels of scene_pack_device.cu (the per-frame path of
// rtb_render*: one thread per primitive, double precision).  Only +, -, *, /, sqrt, fabs, fmax,
// fmin on doubles and double->float conversions are used, all IEEE-exact on both sides (g++
// -ffp-contract=off, nvcc -fmad=false), so host- and device-packed records are bit-identical
// (tests/test_gpu_device_pack.py) and the CPU conservativeness / emulator tests speak for the
// records the kernels really read.
#ifndef RTB_SCENE_PACK_IMPL_H_
#define RTB_SCENE_PACK_IMPL_H_

#include <math.h>

#include "scene_pack.h"

#if defined(__CUDACC__)
#define RTB_PK __host__ __device__ inline
#else
#define RTB_PK inline
#endif

namespace rtb {
namespace pk {

// Bit j of kSignBits[k] = coordinate k of vertex j is +1 in RayTrace/Cube.cpp:10-45's vertex list
// (the same table as kCubeVertexSigns, scene_pack.cpp, from which host/Cube.cpp builds its vertices: a
// mismatch makes every cube of every CPU packer test fail box_from_vertices).
#define RTB_SIGN_BITS_X 0xa5ff08748ull
#define RTB_SIGN_BITS_Y 0x7faa4622cull
#define RTB_SIGN_BITS_Z 0xf4e3da046ull

RTB_PK bool finite_d(double x) { return fabs(x) <= 1.7976931348623157e308; }  // false for NaN / Inf
RTB_PK double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
RTB_PK double dot4(const double* a, const double* b) {
    return a[0] * b[0] + a[1] * b[1] + a[2] * b[2] + a[3] * b[3];
}
RTB_PK void cross3(const double* a, const double* b, double* o) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}

RTB_PK RtbBound bound_always() { RtbBound b = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}; return b; }  // q = 0: every ray passes
RTB_PK RtbBound bound_never() { RtbBound b = {0.f, 0.f, 2.f, 0.f, 0.f, 0.f, 0.f, 0.f}; return b; }   // w = 2, q = 4: no ray passes

RTB_PK bool box_from_vertices(const float* v, Box* out) {
    // One representative vertex per corner (sign triple), then least-squares centre / half-axes.
    double corner[8][3];
    bool seen[8] = {false, false, false, false, false, false, false, false};
    for (int j = 0; j < 36; ++j) {
        const int idx = (int)((RTB_SIGN_BITS_X >> j) & 1ull) | ((int)((RTB_SIGN_BITS_Y >> j) & 1ull) << 1) |
                        ((int)((RTB_SIGN_BITS_Z >> j) & 1ull) << 2);
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
            if (fabs(p) > maxabs) maxabs = fabs(p);
        }
    }
    double ext = 0.0;
    for (int a = 0; a < 3; ++a) ext += sqrt(dot3(out->a[a], out->a[a]));
    if (!finite_d(maxabs) || !finite_d(ext)) return false;
    // every one of the 36 vertices must sit on its corner (float rounding of the transforms only)
    const double tol = 4e-6 * maxabs + 1e-5 * ext + 1e-30;
    for (int j = 0; j < 36; ++j) {
        const double s0 = ((RTB_SIGN_BITS_X >> j) & 1ull) ? 1.0 : -1.0, s1 = ((RTB_SIGN_BITS_Y >> j) & 1ull) ? 1.0 : -1.0,
                     s2 = ((RTB_SIGN_BITS_Z >> j) & 1ull) ? 1.0 : -1.0;
        for (int k = 0; k < 3; ++k) {
            const double want = out->c[k] + s0 * out->a[0][k] + s1 * out->a[1][k] + s2 * out->a[2][k];
            if (fabs((double)v[4 * j + k] - want) > tol) return false;
        }
    }
    double cr[3];
    cross3(out->a[1], out->a[2], cr);
    const double det = dot3(out->a[0], cr);
    const double n0 = sqrt(dot3(out->a[0], out->a[0])), n1 = sqrt(dot3(out->a[1], out->a[1])),
                 n2 = sqrt(dot3(out->a[2], out->a[2]));
    out->degenerate = !(fabs(det) > 1e-12 * n0 * n1 * n2) || !(fabs(det) > 1e-300);
    return true;
}

RTB_PK void pack_sphere(const CameraFrame& cf, const float centre[4], float radius, RtbSphere* out) {
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

RTB_PK void pack_box(const CameraFrame& cf, const Box& box, RtbSlab out[3]) {
    const double kPow40 = 1099511627776.0;  // 2^40
    const double kPow30 = 1073741824.0;     // 2^30
    if (box.degenerate) {  // never hit: tn = +1e30 > tf = -1e30
        out[0].alpha = 1e30f; out[0].beta = 0.f; out[0].gamma = 0.f; out[0].width = 0.f;
        out[1].alpha = -1e30f; out[1].beta = 0.f; out[1].gamma = 0.f; out[1].width = 0.f;
        out[2].alpha = -1e30f; out[2].beta = 0.f; out[2].gamma = 0.f; out[2].width = 0.f;
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
        if (fabs(4.0 * det * dl) < 0.000001) {
            // Ray parallel to the slab: it constrains only the pixel footprint, |q_i(x,y)| <= 1.
            // Expressed as near/far planes scaled by powers of two so that integer-aligned
            // edges stay exact in float (inclusive edges, MainState.cpp:282-291).
            const double minf = fmax(fabs(m[0]), fmax(fabs(m[1]), fabs(m[2])));
            const double g = 1.0 / minf;
            alpha = kPow40 * (-g - g * q0) - kPow30;
            bx = -kPow40 * (g * qx);
            by = -kPow40 * (g * qy);
            width = 2.0 * kPow40 * g + 2.0 * kPow30;
        } else {
            const double inv = 1.0 / dl;
            alpha = -q0 * inv - fabs(inv);
            bx = -qx * inv;
            by = -qy * inv;
            width = 2.0 * fabs(inv);
        }
        // (x, y) -> (u, v):  x = (u - l12*v/l22)/l11,  y = v/l22
        out[i].alpha = (float)alpha;
        out[i].beta = (float)(bx / cf.l11);
        out[i].gamma = (float)(by / cf.l22 - bx * cf.l12 / (cf.l11 * cf.l22));
        out[i].width = (float)width;
    }
}

// ---- conservative conic bounds -------------------------------------------------------------
struct P2 { double u, v; };

// Keeps the part of the convex polygon `poly` (n vertices) with hu*u + hv*v + hc <= 0.
RTB_PK int clip_halfplane(const P2* poly, int n, double hu, double hv, double hc, P2* out) {
    int m = 0;
    for (int i = 0; i < n; ++i) {
        const P2 a = poly[i], b = poly[(i + 1) % n];
        const double fa = hu * a.u + hv * a.v + hc, fb = hu * b.u + hv * b.v + hc;
        if (fa <= 0.0) out[m++] = a;
        if ((fa <= 0.0) != (fb <= 0.0)) {
            const double t = fa / (fa - fb);
            P2 c;
            c.u = a.u + t * (b.u - a.u);
            c.v = a.v + t * (b.v - a.v);
            out[m++] = c;
        }
    }
    return m;
}

// Ellipse  a*(u-cu)^2 + 2b*(u-cu)(v-cv) + d*(v-cv)^2 <= 1  ->  RtbBound.  Completing the square,
//   Q = a*[(u-cu) + (b/a)(v-cv)]^2 + (d - b^2/a)*(v-cv)^2,
// and q = s*Q with the largest s for which the fp32 evaluation (rtb_bound_q) of any point with
// Q <= 1 stays below 2 (error budget below).
RTB_PK void emit_conic(const FrameExtent& fe, double cu, double cv, double a, double b, double d, RtbBound* out) {
    if (!(a > 0.0) || !finite_d(a) || !finite_d(b) || !finite_d(d) || !finite_d(cu) || !finite_d(cv)) {
        *out = bound_always();
        return;
    }
    const double k = b / a;
    const double e0 = fmax(d - b * b / a, 0.0);
    const double U = fe.uabs, V = fe.vabs;
    const double eps22 = 1.0 / 4194304.0;
    double s = 2.0 * (1.0 - 1.0 / 512.0);
    for (int it = 0; it < 400; ++it, s *= 0.9) {
        const double sA = sqrt(s * a), k1 = sA * k, k0 = -sA * (cu + k * cv), e = s * e0;
        // |w_fp32 - w| <= ew: roundings of sA, k1, k0 (double -> float) and of the two fma
        const double ew = eps22 * (sA * U + fabs(k1) * V + fabs(k0));
        // |rowC_fp32 - rowC| <= ec: rounding of cv inside dv (rowC <= 2 there, so e*|dv| <= sqrt(2e)),
        // plus three relative roundings of the products
        const double ec = eps22 * (sqrt(2.0 * e) * (fabs(cv) + V) + 2.0);
        const double wmax = sqrt(s) + ew;
        const double worst = s * (1.0 + 1.0 / 1048576.0) + 2.0 * wmax * ew + ew * ew + ec;
        if (worst < 2.0 * (1.0 - 1.0 / 4096.0)) {
            RtbBound r;
            r.sA = (float)sA; r.k1 = (float)k1; r.k0 = (float)k0; r.e = (float)e; r.cv = (float)cv;
            r.pad0 = 0.f; r.pad1 = 0.f; r.pad2 = 0.f;
            const bool ok = finite_d((double)r.sA) && finite_d((double)r.k1) && finite_d((double)r.k0) &&
                            finite_d((double)r.e) && finite_d((double)r.cv);
            *out = ok ? r : bound_always();
            return;
        }
    }
    *out = bound_always();
}

RTB_PK void pack_bound_box(const FrameExtent& fe, const RtbSlab slabs[3], RtbBound* out) {
    double al[3], be[3], ga[3], wi[3], E[3];
    for (int k = 0; k < 3; ++k) {
        al[k] = slabs[k].alpha; be[k] = slabs[k].beta; ga[k] = slabs[k].gamma; wi[k] = slabs[k].width;
        if (!finite_d(al[k]) || !finite_d(be[k]) || !finite_d(ga[k]) || !finite_d(wi[k])) {
            *out = bound_always();
            return;
        }
        // fp32 error of near_k / far_k as the kernel computes them (2 fma + 1 add), doubled
        E[k] = (fabs(al[k]) + fabs(be[k]) * fe.uabs + fabs(ga[k]) * fe.vabs + fabs(wi[k])) / 4194304.0;
    }
    // A hit needs near_i <= far_j for every pair: three strips in (u,v); clip the frame box by them.
    P2 bufa[16], bufb[16];
    P2* poly = bufa;
    P2* next = bufb;
    int n = 4;
    poly[0].u = fe.umin; poly[0].v = fe.vmin; poly[1].u = fe.umax; poly[1].v = fe.vmin;
    poly[2].u = fe.umax; poly[2].v = fe.vmax; poly[3].u = fe.umin; poly[3].v = fe.vmax;
    for (int i = 0; i < 3 && n > 0; ++i)
        for (int j = 0; j < 3 && n > 0; ++j) {
            if (i == j) continue;
            double hu = be[i] - be[j], hv = ga[i] - ga[j], hc = al[i] - al[j] - wi[j] - (E[i] + E[j]);
            const double m = fmax(fabs(hu), fabs(hv));
            if (!(m * (fe.uabs + fe.vabs) > 1e-9 * fabs(hc)) || m == 0.0) {
                if (hc > 0.0) n = 0;  // constant and violated: the cube is never hit
                continue;
            }
            hu /= m; hv /= m; hc /= m;
            n = clip_halfplane(poly, n, hu, hv, hc, next);
            P2* t = poly; poly = next; next = t;
        }
    if (n <= 0) {
        *out = bound_never();
        return;
    }
    double cu = 0.0, cv = 0.0;
    for (int i = 0; i < n; ++i) { cu += poly[i].u; cv += poly[i].v; }
    cu /= n; cv /= n;
    // four symmetric helper points keep the fit non-degenerate (a clipped sliver, a single vertex)
    const double r0 = 0.05 * fe.pitch;
    poly[n].u = cu + r0; poly[n].v = cv; poly[n + 1].u = cu - r0; poly[n + 1].v = cv;
    poly[n + 2].u = cu; poly[n + 2].v = cv + r0; poly[n + 3].u = cu; poly[n + 3].v = cv - r0;
    const int nt = n + 4;
    double suu = 0.0, suv = 0.0, svv = 0.0;
    for (int i = 0; i < nt; ++i) {
        const double x = poly[i].u - cu, y = poly[i].v - cv;
        suu += x * x; suv += x * y; svv += y * y;
    }
    suu /= nt; suv /= nt; svv /= nt;
    const double det = suu * svv - suv * suv;
    if (!(det > 0.0) || !finite_d(det)) {
        *out = bound_always();
        return;
    }
    const double iuu = svv / det, iuv = -suv / det, ivv = suu / det;
    double T = 0.0;
    for (int i = 0; i < nt; ++i) {
        const double x = poly[i].u - cu, y = poly[i].v - cv;
        T = fmax(T, iuu * x * x + 2.0 * iuv * x * y + ivv * y * y);
    }
    T *= 1.0 + 1e-6;
    emit_conic(fe, cu, cv, iuu / T, iuv / T, ivv / T, out);
}

RTB_PK void pack_bound_sphere(const FrameExtent& fe, const RtbSphere& sph, RtbBound* out) {
    const double rho = sph.rho, uc = sph.uc, vc = sph.vc;
    if (!finite_d(rho) || !finite_d(uc) || !finite_d(vc)) {
        *out = bound_always();
        return;
    }
    if (rho < 0.0) {  // rtb_sphere_rad is then negative on every ray (two monotone roundings)
        *out = bound_never();
        return;
    }
    // rad >= 0 in fp32  =>  (uc-u)^2 + (vc-v)^2 <= rho*(1 + 3*2^-24 + ...)
    const double r0 = 0.05 * fe.pitch;
    const double T = rho * (1.0 + 1.0 / 1048576.0) + r0 * r0;
    emit_conic(fe, uc, vc, 1.0 / T, 0.0, 1.0 / T, out);
}

// ---- row table (the dense grid kernel's per-(primitive, row) entries) ------------------------
// Entry of bound `b` on the pixel row whose camera coordinate is the FLOAT v the kernel computes
// (rtb_cam_v).  Claim (raycast_math.h, RtbRowEntry): for every float u with |u| <= uabs,
//     fmaf(w, w, rowC) < 2,  w = fmaf(u, sA, rowW)    =>    |fmaf(u, sAn, rWn)| < 1.
// Proof.  Let t = u*sA + rowW (real), so w = RN(t).  The premise gives w*w + rowC < 2 (RN(x) >= 2 for x >= 2),
// i.e. |w| < h = sqrt(2 - rowC), hence |t| < h*(1 + 2^-22) =: hh.  With sAn = RN(sA*g), rWn = RN(rowW*g):
// |u*sAn + rWn| <= g*|t| + 2^-24*g*(|u||sA| + |rowW|)(1 + tiny) < g*(hh + 2^-23*E), E = uabs*|sA| + |rowW| (+ a
// subnormal allowance), and the fp32 fma adds one more relative 2^-24.  g = (1 - 2^-20) / (hh + 2^-22*E) makes
// the result < (1 - 2^-20)(1 + 2^-24) < 1.  rowC >= 2 or NaN: the premise is never true -> the "never" entry
// (0, 2).  Non-finite sA / rowW: w is Inf or NaN on every ray, the premise is never true -> "never".
// Only IEEE-exact double operations: host- and device-built tables are bit-identical.
RTB_PK void pack_row_entry(const RtbBound& b, float v, double uabs, float out[2]) {
    const float rowW = rtb_bound_row_w(b, v);
    const float rowC = rtb_bound_row_c(b, v);
    const double sA = (double)b.sA, rW = (double)rowW, rC = (double)rowC;
    if (!(rC < 2.0) || !finite_d(sA) || !finite_d(rW) || !finite_d(uabs)) {
        out[0] = 0.0f;
        out[1] = 2.0f;
        return;
    }
    const double hh = sqrt(2.0 - rC) * (1.0 + 1.0 / 4194304.0);
    const double E = uabs * fabs(sA) + fabs(rW) + 1e-30;
    const double g = (1.0 - 1.0 / 1048576.0) / (hh + E / 4194304.0);
    const float sAn = (float)(sA * g), rWn = (float)(rW * g);
    if (!finite_d((double)sAn) || !finite_d((double)rWn)) {  // cannot happen for finite inputs (g <= 2^22 / E); be safe: always pass
        out[0] = 0.0f;
        out[1] = 0.0f;
        return;
    }
    out[0] = sAn;
    out[1] = rWn;
}

}  // namespace pk
}  // namespace rtb

#endif  // RTB_SCENE_PACK_IMPL_H_
