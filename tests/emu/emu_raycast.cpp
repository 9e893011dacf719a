// emu_raycast.cpp -- TEST INFRASTRUCTURE.  Scalar CPU walk-through of raycast_grid_kernel's
// arithmetic (opencl_ray_tracer_b200/csrc/raycast_kernels.cu), built from the SAME header
// (raycast_math.h) with contraction off, so it reproduces the kernel bit for bit.  It lets the
// CPU test-suite check the host packer + kernel formulation against the oracle without a GPU.
// It is not linked into, nor reachable from, the product library.
#include <cstdint>
#include <cstddef>

#include "../../opencl_ray_tracer_b200/csrc/raycast_math.h"

extern "C" int emu_render_grid(int width, int height, const float* cam8, int nc, const float* cube_rec,
                               int ns, const float* sph_rec, const float* cube_col, const float* sph_col,
                               int32_t* rgba_i32, int32_t* prim_id, float* closest) {
    const RtbCam cam = {cam8[0], cam8[1], cam8[2], cam8[3], cam8[4], 0.f, 0.f, 0.f};
    const RtbSlab* slabs = reinterpret_cast<const RtbSlab*>(cube_rec);
    const RtbSphere* sph = reinterpret_cast<const RtbSphere*>(sph_rec);
#pragma omp parallel for schedule(dynamic, 4)
    for (int y = 0; y < height; ++y) {
        const float fy = (float)y;
        const float v = rtb_cam_v(cam, fy);
        for (int x = 0; x < width; ++x) {
            const float u = rtb_cam_u(cam, (float)x, fy);
            float best = RTB_MISS_T;
            int id = -1;
            for (int c = 0; c < nc; ++c) {
                const RtbSlab &s0 = slabs[3 * c], &s1 = slabs[3 * c + 1], &s2 = slabs[3 * c + 2];
                const float n0 = rtb_slab_near(s0, rtb_slab_base(s0, v), u);
                const float n1 = rtb_slab_near(s1, rtb_slab_base(s1, v), u);
                const float n2 = rtb_slab_near(s2, rtb_slab_base(s2, v), u);
                const float f0 = n0 + s0.width, f1 = n1 + s1.width, f2 = n2 + s2.width;
                const float tn = fmaxf(fmaxf(n0, n1), n2);
                const float tf = fminf(fminf(f0, f1), f2);
                rtb_cube_update(tn, tf, c, best, id);
            }
            const float wv = cam.tau_v * v;
            for (int s = 0; s < ns; ++s) {
                const float rad = rtb_sphere_rad(sph[s], rtb_sphere_radv(sph[s], v), u);
                const float wpix = fmaf(cam.tau_u, u, wv);
                rtb_sphere_update(sph[s], rad, wpix, nc + s, best, id);
            }
            if (best == RTB_MISS_T) id = -1;
            float cr = 0.f, cg = 0.f, cb = 0.f;
            if (id >= 0) {
                const float* col = (id < nc) ? cube_col + 4 * id : sph_col + 4 * (id - nc);
                cr = col[0]; cg = col[1]; cb = col[2];
            }
            const size_t i = (size_t)y * width + x;
            rtb_shade(best, cr, cg, cb, rgba_i32 + 4 * i);
            if (prim_id) prim_id[i] = id;
            if (closest) closest[i] = best;
        }
    }
    return 0;
}

// Conservativeness of the conic bounds (RtbBound): walks every pixel x primitive exactly like the
// kernel's dense loop and counts (a) pixels where the reference-exact update could fire
// (footprint: slab interval non-empty / sphere rad >= 0) although the bound says "fail" --
// must be 0 -- and (b) how many tests pass the bound vs how many are in the footprint
// (selectivity, informational).  out[0] = violations, out[1] = bound passes, out[2] = footprint
// tests, out[3] = total tests.
extern "C" int emu_check_bounds(int width, int height, const float* cam8, int nc, const float* cube_rec, int ns,
                                const float* sph_rec, const float* bound_rec, long long* out) {
    const RtbCam cam = {cam8[0], cam8[1], cam8[2], cam8[3], cam8[4], 0.f, 0.f, 0.f};
    const RtbSlab* slabs = reinterpret_cast<const RtbSlab*>(cube_rec);
    const RtbSphere* sph = reinterpret_cast<const RtbSphere*>(sph_rec);
    const RtbBound* bnd = reinterpret_cast<const RtbBound*>(bound_rec);
    long long viol = 0, pass = 0, foot = 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : viol, pass, foot)
    for (int y = 0; y < height; ++y) {
        const float fy = (float)y;
        const float v = rtb_cam_v(cam, fy);
        for (int x = 0; x < width; ++x) {
            const float u = rtb_cam_u(cam, (float)x, fy);
            for (int i = 0; i < nc + ns; ++i) {
                const RtbBound& b = bnd[i];
                const float q = rtb_bound_q(b.sA, rtb_bound_row_w(b, v), rtb_bound_row_c(b, v), u);
                const bool bound_pass = q < 2.0f;  // NaN fails
                bool in_foot;
                if (i < nc) {
                    const RtbSlab &s0 = slabs[3 * i], &s1 = slabs[3 * i + 1], &s2 = slabs[3 * i + 2];
                    const float n0 = rtb_slab_near(s0, rtb_slab_base(s0, v), u);
                    const float n1 = rtb_slab_near(s1, rtb_slab_base(s1, v), u);
                    const float n2 = rtb_slab_near(s2, rtb_slab_base(s2, v), u);
                    const float f0 = n0 + s0.width, f1 = n1 + s1.width, f2 = n2 + s2.width;
                    in_foot = fmaxf(fmaxf(n0, n1), n2) <= fminf(fminf(f0, f1), f2);
                } else {
                    const RtbSphere& s = sph[i - nc];
                    in_foot = rtb_sphere_rad(s, rtb_sphere_radv(s, v), u) >= 0.0f;
                }
                pass += bound_pass;
                foot += in_foot;
                viol += (in_foot && !bound_pass);
            }
        }
    }
    out[0] = viol; out[1] = pass; out[2] = foot; out[3] = (long long)width * height * (nc + ns);
    return 0;
}

// Tile cull of raycast_cull_kernel (rtb_bound_tile_misses): for every tile_w x tile_h pixel tile
// and primitive, the cull may say "miss" only if rtb_bound_q is >= 2 (or NaN) on every pixel of
// the tile.  out[0] = violations (culled although some pixel passes), out[1] = culled pairs,
// out[2] = pairs in which no pixel passes (what an ideal cull would drop), out[3] = pairs.
extern "C" int emu_check_tile_cull(int width, int height, const float* cam8, int n_bound, const float* bound_rec,
                                   int tile_w, int tile_h, long long* out) {
    const RtbCam cam = {cam8[0], cam8[1], cam8[2], cam8[3], cam8[4], 0.f, 0.f, 0.f};
    const RtbBound* bnd = reinterpret_cast<const RtbBound*>(bound_rec);
    const int tx = (width + tile_w - 1) / tile_w, ty = (height + tile_h - 1) / tile_h;
    long long viol = 0, culled = 0, empty = 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : viol, culled, empty)
    for (int t = 0; t < tx * ty; ++t) {
        const int x_lo = (t % tx) * tile_w, y_lo = (t / tx) * tile_h;
        // the kernel's ranges: all tile_h rows and both end columns of the full tile, in or out of the frame
        float u_min = 0.f, u_max = 0.f, v_min = 0.f, v_max = 0.f;
        for (int r = 0; r < tile_h; ++r) {
            const float fy = (float)(y_lo + r);
            const float v = rtb_cam_v(cam, fy);
            const float ua = rtb_cam_u(cam, (float)x_lo, fy), ub = rtb_cam_u(cam, (float)(x_lo + tile_w - 1), fy);
            if (r == 0) { u_min = u_max = ua; v_min = v_max = v; }
            u_min = fminf(u_min, fminf(ua, ub)); u_max = fmaxf(u_max, fmaxf(ua, ub));
            v_min = fminf(v_min, v); v_max = fmaxf(v_max, v);
        }
        {   // the four-corner form used by bin_tiles_kernel must give the very same ranges
            float cu0, cu1, cv0, cv1;
            rtb_tile_uv_range(cam, (float)x_lo, (float)(x_lo + tile_w - 1), (float)y_lo, (float)(y_lo + tile_h - 1), &cu0, &cu1, &cv0, &cv1);
            if (cu0 != u_min || cu1 != u_max || cv0 != v_min || cv1 != v_max) viol += n_bound;
        }
        for (int i = 0; i < n_bound; ++i) {
            const RtbBound& b = bnd[i];
            const bool miss = rtb_bound_tile_misses(b, u_min, u_max, v_min, v_max);
            bool any_pass = false;
            for (int r = 0; r < tile_h && !any_pass; ++r) {
                const float fy = (float)(y_lo + r);
                const float v = rtb_cam_v(cam, fy);
                const float rw = rtb_bound_row_w(b, v), rc = rtb_bound_row_c(b, v);
                for (int x = x_lo; x < x_lo + tile_w; ++x) {
                    if (rtb_bound_q(b.sA, rw, rc, rtb_cam_u(cam, (float)x, fy)) < 2.0f) { any_pass = true; break; }
                }
            }
            culled += miss;
            empty += !any_pass;
            viol += (miss && any_pass);
        }
    }
    out[0] = viol; out[1] = culled; out[2] = empty; out[3] = (long long)tx * ty * n_bound;
    return 0;
}

// rtb_div180 (raycast_math.h) against the IEEE divide on every stride-th bit pattern in
// [first_bits, last_bits] (Inf / NaN skipped).  Returns the number of mismatches.
extern "C" long long emu_check_div180(unsigned first_bits, unsigned last_bits, unsigned stride) {
    long long bad = 0;
    const long long n = ((long long)last_bits - (long long)first_bits) / stride + 1;
#pragma omp parallel for schedule(static) reduction(+ : bad)
    for (long long k = 0; k < n; ++k) {
        const uint32_t b = (uint32_t)((long long)first_bits + k * stride);
        float x;
        __builtin_memcpy(&x, &b, 4);
        if (!(fabsf(x) <= 3.4028234663852886e38f)) continue;  // Inf / NaN
        const volatile float d = 180.0f;
        const float ref = x / d, got = rtb_div180(x);
        uint32_t a, c;
        __builtin_memcpy(&a, &ref, 4);
        __builtin_memcpy(&c, &got, 4);
        bad += (a != c);
    }
    return bad;
}

// The kernels' per-thread form of the bound: over a run of `run` horizontally adjacent pixels (a thread's rays)
// fmaf(m, m, rowC) with m = min |w| must equal the minimum of the per-ray q, bit for bit ("some ray has q < 2" is
// then one compare per thread).  Returns the number of runs on which the two differ as a pass / fail decision or
// as a value.
extern "C" long long emu_check_min_form(int width, int height, const float* cam8, int n_bound, const float* bound_rec, int run) {
    const RtbCam cam = {cam8[0], cam8[1], cam8[2], cam8[3], cam8[4], 0.f, 0.f, 0.f};
    const RtbBound* bnd = reinterpret_cast<const RtbBound*>(bound_rec);
    long long bad = 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : bad)
    for (int y = 0; y < height; ++y) {
        const float fy = (float)y;
        const float v = rtb_cam_v(cam, fy);
        for (int i = 0; i < n_bound; ++i) {
            const RtbBound& b = bnd[i];
            const float rw = rtb_bound_row_w(b, v), rc = rtb_bound_row_c(b, v);
            for (int x0 = 0; x0 < width; x0 += run) {
                float m = __builtin_inff(), qmin = __builtin_inff();
                bool any = false;
                for (int x = x0; x < x0 + run; ++x) {  // past the frame's edge like the kernel's threads
                    const float w = rtb_bound_w(b.sA, rw, rtb_cam_u(cam, (float)x, fy));
                    m = fminf(m, fabsf(w));
                    const float q = rtb_bound_q_of_w(w, rc);
                    qmin = fminf(qmin, q);
                    any = any || (q < 2.0f);
                }
                const float qm = rtb_bound_q_of_w(m, rc);
                if ((qm < 2.0f) != any) bad += 1;
                else if (qm == qm && qmin == qmin && qm != qmin) bad += 1;
                // the span form (binned render): never above the true minimum, and it accepts whenever a ray passes
                const float wa = rtb_bound_w(b.sA, rw, rtb_cam_u(cam, (float)x0, fy));
                const float wb = rtb_bound_w(b.sA, rw, rtb_cam_u(cam, (float)(x0 + run - 1), fy));
                const float ms = rtb_bound_span_m(wa, wb);
                if (ms > m) bad += 1;
                if (any && !(rtb_bound_q_of_w(ms, rc) < 2.0f)) bad += 1;
            }
        }
    }
    return bad;
}

// Row-table form of the bound (raycast_math.h RtbRowEntry; scene_pack_impl.h pack_row_entry): the dense kernel
// lets a ray through to the exact arithmetic iff |fmaf(u, sAn, rWn)| < 1.  That must hold on every ray on which
// the per-ray test fmaf(w, w, rowC) < 2 holds (which in turn covers the footprint, emu_check_bounds).
// table[(y * n_bound + i) * 2] = the entries as the packer built them.  out[0] = violations, out[1] = rays passing
// the q test, out[2] = rays passing the row entry, out[3] = total.
extern "C" int emu_check_row_table(int width, int height, const float* cam8, int n_bound, const float* bound_rec,
                                   const float* table, long long* out) {
    const RtbCam cam = {cam8[0], cam8[1], cam8[2], cam8[3], cam8[4], 0.f, 0.f, 0.f};
    const RtbBound* bnd = reinterpret_cast<const RtbBound*>(bound_rec);
    long long viol = 0, pass_q = 0, pass_row = 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : viol, pass_q, pass_row)
    for (int y = 0; y < height; ++y) {
        const float fy = (float)y;
        const float v = rtb_cam_v(cam, fy);
        for (int i = 0; i < n_bound; ++i) {
            const RtbBound& b = bnd[i];
            const float rw = rtb_bound_row_w(b, v), rc = rtb_bound_row_c(b, v);
            const float* e = table + 2 * ((size_t)y * n_bound + i);
            for (int x = 0; x < width; ++x) {
                const float u = rtb_cam_u(cam, (float)x, fy);
                const bool q_ok = rtb_bound_q(b.sA, rw, rc, u) < 2.0f;
                const bool row_ok = rtb_row_passes(fabsf(rtb_row_w(e[0], e[1], u)));
                pass_q += q_ok;
                pass_row += row_ok;
                viol += (q_ok && !row_ok);
            }
            // the dense kernel's cold path (dense_group_slow): behind a vote a thread re-tests its run of 16 rays in the
            // span form -- w' at the run's two end rays; it must accept every run in which a ray passes the row entry
            for (int x0 = 0; x0 < width; x0 += 16) {
                bool any = false;
                for (int x = x0; x < x0 + 16; ++x)  // past the frame's edge like the kernel's threads
                    any = any || rtb_row_passes(fabsf(rtb_row_w(e[0], e[1], rtb_cam_u(cam, (float)x, fy))));
                const float wa = rtb_row_w(e[0], e[1], rtb_cam_u(cam, (float)x0, fy));
                const float wb = rtb_row_w(e[0], e[1], rtb_cam_u(cam, (float)(x0 + 15), fy));
                viol += (any && !rtb_row_passes(rtb_bound_span_m(wa, wb)));
            }
        }
    }
    out[0] = viol; out[1] = pass_q; out[2] = pass_row; out[3] = (long long)width * height * n_bound;
    return 0;
}
