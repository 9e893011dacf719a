// raycast_math.h -- per-pixel arithmetic of the B200 ray-cast path, shared verbatim between the
// CUDA kernels (raycast_kernels.cu) and the test-only scalar emulator (tests/emu/).  Everything
// here is written with explicit fmaf() and is compiled with FMA contraction OFF on both sides
// (nvcc -fmad=false, g++ -ffp-contract=off), with IEEE divide and sqrt, so the emulator and the
// kernel produce bit-identical pixels.
//
// Formulation.  All rays of a frame are parallel (one shared rayDir,
// RayTrace/states/MainState.cpp:37-39) and their origins form an affine pixel grid
// (:41-50), so every per-primitive quantity the reference evaluates per ray is a polynomial in
// the pixel coordinates.  The host (scene_pack.cpp) changes variables to camera coordinates
//   u = l11*x + l12*y,  v = l22*y
// in which
//   sphere i (MainState.cpp:300-327):  r^2 - d^2 = rho_i - (uc_i-u)^2 - (vc_i-v)^2,
//                                      tca = tau0_i - (tau_u*u + tau_v*v)
//   cube c   (MainState.cpp:347-377):  min over its 12 triangles of the line parameter t
//                                      = max_k near_k,  hit iff max_k near_k <= min_k far_k,
//                                      near_k = alpha_k + beta_k*u + gamma_k*v, far_k = near_k + w_k
//                                      (slab test on the parallelepiped spanned by the cube's
//                                      vertices; edges inclusive, t may be negative, as in
//                                      intersectTri which has no t>0 rejection).
#ifndef RTB_RAYCAST_MATH_H_
#define RTB_RAYCAST_MATH_H_

#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define RTB_HD __host__ __device__ __forceinline__
#else
#define RTB_HD inline
#endif

#define RTB_MISS_T 300000.0f  // `closest` sentinel, MainState.cpp:345

// Camera constants for one frame (8 floats).
struct RtbCam {
    float l11, l12, l22;  // u = l11*x + l12*y, v = l22*y
    float tau_u, tau_v;   // ray-origin offset along rayDir: wpix = tau_u*u + tau_v*v
    float pad0, pad1, pad2;
};

// One cube = 3 slab records; one sphere = 1 record.  16 bytes each (one LDS.128).
struct RtbSlab { float alpha, beta, gamma, width; };
struct RtbSphere { float uc, vc, rho, tau0; };

RTB_HD float rtb_cam_u(const RtbCam& c, float x, float y) { return fmaf(c.l11, x, c.l12 * y); }
RTB_HD float rtb_cam_v(const RtbCam& c, float y) { return c.l22 * y; }

// slab: base = alpha + gamma*v (shared by all pixels of a row), near = base + beta*u
RTB_HD float rtb_slab_base(const RtbSlab& s, float v) { return fmaf(s.gamma, v, s.alpha); }
RTB_HD float rtb_slab_near(const RtbSlab& s, float base, float u) { return fmaf(s.beta, u, base); }

// sphere: radv = rho - (vc-v)^2 (shared by a row), rad = radv - (uc-u)^2 = r^2 - d^2
RTB_HD float rtb_sphere_radv(const RtbSphere& s, float v) {
    const float dv = s.vc - v;
    return fmaf(-dv, dv, s.rho);
}
RTB_HD float rtb_sphere_rad(const RtbSphere& s, float radv, float u) {
    const float du = s.uc - u;
    return fmaf(-du, du, radv);
}

// Nearest-hit update for one cube on one pixel: strict '<' so the lowest index wins ties
// (MainState.cpp:370-374); the slab bounds are inclusive like the reference's u,v tests (:282-291).
RTB_HD void rtb_cube_update(float tn, float tf, int id, float& best, int& best_id) {
    if (tn <= tf && tn < best) { best = tn; best_id = id; }
}

// Sphere update, MainState.cpp:300-327 and :382-393: needs rad >= 0 (d^2 <= r^2), tca >= 0,
// t0 != 0 (0.0f is the reference's miss value), t0 < closest.  t0 < 0 (origin inside) is kept.
RTB_HD void rtb_sphere_update(const RtbSphere& s, float rad, float wpix, int id, float& best,
                              int& best_id) {
    if (rad >= 0.0f) {
        const float tca = s.tau0 - wpix;
        if (tca >= 0.0f) {
            const float t0 = tca - sqrtf(rad);
            if (t0 != 0.0f && t0 < best) { best = t0; best_id = id; }
        }
    }
}

// The reference's float sphere test in the reference's operation order (MainState.cpp:300-327,
// glm::dot order (x+y)+(z+w), glm/detail/func_geometric.inl:75-82); 0.0f = miss.  Used by the
// explicit-origin kernel, where no per-frame precompute is possible.  Needs contraction OFF.
RTB_HD float rtb_sphere_t_reference(const float o[4], const float d[4], float radius, const float c[4]) {
    const float lx = c[0] - o[0], ly = c[1] - o[1], lz = c[2] - o[2], lw = c[3] - o[3];
    const float tca = (lx * d[0] + ly * d[1]) + (lz * d[2] + lw * d[3]);
    if (tca < 0.0f) return 0.0f;
    const float d2 = ((lx * lx + ly * ly) + (lz * lz + lw * lw)) - tca * tca;
    const float r2 = radius * radius;
    if (d2 > r2) return 0.0f;
    return tca - sqrtf(r2 - d2);
}

// ---- conservative screen-space bound (the dense per-ray test of the grid kernel) -------------
// Every primitive's footprint under the frame's parallel projection is convex (a cube projects
// to a hexagon, a sphere to a disc), so the host encloses it in one ellipse, written with the
// square completed in u:
//   q(u,v) = w^2 + e*(v - cv)^2,   w = sA*u + (k1*v + k0),      footprint  =>  q < 2
// (scaled so that fp32 evaluation error of w = fmaf(u, sA, rowW), q = fmaf(w, w, rowC) cannot push a
// footprint ray to q >= 2; scene_pack_impl.h).  The kernels evaluate w per ray (ONE FMA) but q only
// once per thread and primitive: fmaf(w, w, rowC) depends on w through |w| alone and is
// non-decreasing in |w| (rounding is monotone), so over the rays of a thread
//     min_rays fmaf(w, w, rowC)  ==  fmaf(m, m, rowC)   with   m = min_rays |w|,
// bit for bit.  A thread therefore keeps the running minimum of |w| (a 3-input FMNMX per ray pair)
// and asks "fmaf(m, m, rowC) < 2" once: exactly "does some ray of mine have q < 2".  A NaN w drops
// out of fminf (its q would be NaN: fail), Inf gives q = Inf: fail.  Only rays that pass reach the
// reference-exact slab / sphere arithmetic above, therefore the bound never changes a pixel:
// results are those of rtb_cube_update / rtb_sphere_update alone.
struct RtbBound { float sA, k1, k0, e, cv, pad0, pad1, pad2; };  // 32 B: LDS.128 + LDS.32

// per row: rowW = k1*v + k0, rowC = (e*(v-cv))*(v-cv);  per ray: w = sA*u + rowW;  q = w*w + rowC
RTB_HD float rtb_bound_row_w(const RtbBound& b, float v) { return fmaf(b.k1, v, b.k0); }
RTB_HD float rtb_bound_row_c(const RtbBound& b, float v) {
    const float dv = v - b.cv;
    return (b.e * dv) * dv;
}
RTB_HD float rtb_bound_w(float sA, float rowW, float u) { return fmaf(u, sA, rowW); }
RTB_HD float rtb_bound_q_of_w(float w, float rowC) { return fmaf(w, w, rowC); }
RTB_HD float rtb_bound_q(float sA, float rowW, float rowC, float u) {
    return rtb_bound_q_of_w(rtb_bound_w(sA, rowW, u), rowC);
}
// A thread's rays are adjacent pixels of one row: u is monotone in x, w = fmaf(u, sA, rowW) is monotone in u, so
// every w of the run lies between those of its two end pixels, w_a and w_b.  The distance from 0 to that interval
// is a lower bound of min |w| over the run -- equal to it unless the run straddles w = 0 (then 0) -- hence
// fmaf(m, m, rowC) < 2 with this m accepts every run that holds a ray with q < 2.  Used by the binned render,
// where a few primitives meet every thread: 2 evaluations of w instead of one per ray.
RTB_HD float rtb_bound_span_m(float w_a, float w_b) {
    return fmaxf(fmaxf(fminf(w_a, w_b), fminf(-w_a, -w_b)), 0.0f);
}

// Row-table form of the same bound (the dense grid kernel).  For one (primitive, pixel row) the host / device
// packer (scene_pack_impl.h, pack_row_entry) scales the row's (sA, rowW) by gamma ~ 1 / sqrt(2 - rowC), rounded so
// that for every ray of the frame on that row
//     fmaf(w, w, rowC) < 2  with  w = fmaf(u, sA, rowW)      implies      |fmaf(u, sAn, rWn)| < 1
// (a superset test: it can only let MORE rays through to the reference-exact arithmetic, which is a no-op outside
// a primitive's footprint).  The threshold is the same constant for every primitive and row, so the running
// minimum of |w'| may run over SEVERAL primitives before it is compared once: 1 FFMA2 + 1 FMNMX3 per ray pair
// and primitive and nothing else per primitive.  An entry that no ray can pass is (0, 2): |w'| = 2.
struct RtbRowEntry { float sAn, rWn; };  // 8 B; two primitives per LDS.128
RTB_HD float rtb_row_w(float sAn, float rWn, float u) { return fmaf(u, sAn, rWn); }
RTB_HD bool rtb_row_passes(float m_abs_w) { return m_abs_w < 1.0f; }

// Tile cull (RTB_OPT_TILE_CULL): can NO ray of a pixel tile pass a primitive's bound?  The tile is given by the
// ranges of the camera coordinates its rays take: u in [u_min, u_max], v in [v_min, v_max] (min / max over the
// very values the kernel computes for the tile's pixels).  The answer is a statement about the values
// rtb_bound_q would compute, derived from the monotonicity of each rounded operation, not from real-number
// geometry plus a margin:
//   rowW = fmaf(k1, v, k0) is monotone in v        => rowW of every row lies in [r_lo, r_hi]
//   w = fmaf(u, sA, rowW) is monotone in u and rowW => w of every ray lies in [w_lo, w_hi]
//   dv = v - cv is monotone in v, rowC = (e*dv)*dv is even and non-decreasing in |dv| for e >= 0
//                                                   => rowC of every row >= c_min >= 0
//   q = fmaf(w, w, rowC) is non-decreasing in |w| and rowC
//                                                   => q of every ray >= fmaf(wm, wm, c_min)
// with wm the smallest |w| the interval allows: if this returns true, every ray of the tile has q >= 2, none
// of them lies in the footprint, and dropping the primitive for the tile cannot change a pixel.  A NaN makes
// the comparison false (keep).  tests/emu checks the claim against rtb_bound_q pixel by pixel.
RTB_HD bool rtb_bound_tile_misses(const RtbBound& b, float u_min, float u_max, float v_min, float v_max) {
    const float ra = rtb_bound_row_w(b, v_min), rb = rtb_bound_row_w(b, v_max);
    const float r_lo = fminf(ra, rb), r_hi = fmaxf(ra, rb);
    const bool up = b.sA >= 0.0f;
    const float w_lo = fmaf(up ? u_min : u_max, b.sA, r_lo);
    const float w_hi = fmaf(up ? u_max : u_min, b.sA, r_hi);
    const float wm = (w_lo > 0.0f) ? w_lo : ((w_hi < 0.0f) ? -w_hi : 0.0f);
    const float da = v_min - b.cv, db = v_max - b.cv;
    const float dm = (da > 0.0f) ? da : ((db < 0.0f) ? db : 0.0f);
    const float c_min = (b.e * dm) * dm;
    return b.e >= 0.0f && fmaf(wm, wm, c_min) >= 2.0f;
}

// Coordinate ranges of the rays of the pixel rectangle [x_a, x_b] x [y_a, y_b] (frame coordinates as
// floats): u = fmaf(l11, x, l12*y) is non-decreasing or non-increasing in x for fixed y and in y for
// fixed x (a rounded product by a constant and fmaf in its addend are monotone), v = l22*y is
// monotone in y, so the extremes over the rectangle's pixels are attained at its four corners --
// exactly, for the rounded values.  tests/emu compares with the loop over every row.
RTB_HD void rtb_tile_uv_range(const RtbCam& c, float x_a, float x_b, float y_a, float y_b, float* u_min,
                              float* u_max, float* v_min, float* v_max) {
    const float u0 = rtb_cam_u(c, x_a, y_a), u1 = rtb_cam_u(c, x_b, y_a);
    const float u2 = rtb_cam_u(c, x_a, y_b), u3 = rtb_cam_u(c, x_b, y_b);
    const float va = rtb_cam_v(c, y_a), vb = rtb_cam_v(c, y_b);
    *u_min = fminf(fminf(u0, u1), fminf(u2, u3));
    *u_max = fmaxf(fmaxf(u0, u1), fmaxf(u2, u3));
    *v_min = fminf(va, vb);
    *v_max = fmaxf(va, vb);
}

// Depth-fog shade + (int) truncation, MainState.cpp:397-407, misc/Utility.cpp:109-116,
// MainState.cpp:952-955.  Same operation order as the reference; no contraction.
//
// x / 180.0f, correctly rounded (the reference's IEEE divide), without the divide: with
// c = RN(1/180), q0 = x*c, r = fmaf(-q0, 180, x), q = fmaf(r, c, q0) equals RN(x/180) for EVERY float
// with 1e-30 <= |x| <= FLT_MAX -- checked exhaustively over all bit patterns
// (tests/emu emu_check_div180, tests/test_kernel_math_emulated.py); below 1e-30 (186 415 patterns
// whose quotient is subnormal or nearly so) the real divide is used.  3 FP32 instructions per pixel
// instead of the ~12 of div.rn.f32.
#if defined(__CUDACC__)
#define RTB_COLD static __host__ __device__ __noinline__
#else
#define RTB_COLD static __attribute__((noinline, unused))
#endif
RTB_COLD float rtb_div180_ieee(float x) { return x / 180.0f; }  // out of line: never runs in a frame
RTB_HD float rtb_div180(float x) {
    const float c = 0x1.6c16c2p-8f;                 // RN(1/180)
    const float q0 = x * c;
    const float r = fmaf(-q0, 180.0f, x);
    const float q = fmaf(r, c, q0);
    if (!(fabsf(x) >= 1e-30f)) return rtb_div180_ieee(x);  // tiny, zero or NaN: the real divide
    return q;
}

RTB_HD void rtb_shade(float closest, float cr, float cg, float cb, int32_t out[4]) {
    if (closest == RTB_MISS_T) {
        out[0] = 0; out[1] = 0; out[2] = 0; out[3] = 255;
        return;
    }
    const float norm = rtb_div180(closest - 0.0f);  // (closest - 0) / (180 - 0), misc/Utility.cpp:109-116
    const float k = 255.0f - (norm * 255.0f);
    out[0] = (int32_t)(k * cr);
    out[1] = (int32_t)(k * cg);
    out[2] = (int32_t)(k * cb);
    out[3] = 255;
}

// (uint8_t) narrowing of MainState.cpp:1023-1037, packed little-endian R,G,B,A.
RTB_HD uint32_t rtb_pack_rgba8(const int32_t c[4]) {
    return ((uint32_t)c[0] & 0xFFu) | (((uint32_t)c[1] & 0xFFu) << 8) |
           (((uint32_t)c[2] & 0xFFu) << 16) | (((uint32_t)c[3] & 0xFFu) << 24);
}

#endif  // RTB_RAYCAST_MATH_H_
