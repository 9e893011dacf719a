// raycast_kernels.cu -- hand-written sm_100a kernels of the ray-cast path.
//
// raycast_dense_kernel replaces the OpenCL kernel `rayTracer`
// (RayTrace/resources/shaders/rayTracer.cl:111-202) and, result-wise, the CPU loop
// MainState::executeRayTracerCPU -> collide (RayTrace/states/MainState.cpp:936-956, :330-408).
//
// Mapping: one warp = an 8x4 lane tile, each lane owns PX horizontally adjacent
// pixels (one primary ray each, held as PX/2 packed FP32 pairs), so a warp covers (8*PX) x 4 pixels
// and a lane's PX RGBA8 results leave as 128-bit stores.  The per-(pixel row, primitive) entries of the
// conservative bound -- the row table, 8 B each, built once per frame by row_table_kernel
// (scene_pack_device.cu) -- are staged global->shared in chunks with 1-D TMA bulk copies (cp.async.bulk +
// mbarrier complete_tx) through a 2-stage ring of 256-primitive chunks with full / empty mbarriers: no CTA-wide
// barrier in the loop.  Per ray pair and primitive: 1 FFMA2 (w' = sAn*u + rWn) + 1 FMNMX3 (running min of |w'| over
// the thread's rays AND over 16 primitives); per 16 primitives one compare (min < 1), one warp vote.  The
// reference-exact slab / sphere update (raycast_math.h) runs only behind the vote.  A CTA is two such warps
// (256 x 4 pixels, 8 CTAs per SM); the last tile rows of a launch run in CTAs of 8 rays per thread on a second
// stream (launch_grid), and a warp's rgba8 pixels leave as whole rows through shared memory (shade_and_store_warp).
//
// Compiled with -fmad=false: every FMA is an explicit fmaf() from raycast_math.h so the scalar
// emulator in tests/emu reproduces the kernel bit for bit.
#include "raycast_kernels.cuh"

#include <cmath>
#include <cstdio>
#include <cstdlib>

namespace rtb {

namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
// 1-D TMA bulk copy global -> shared, completion signalled on the mbarrier (SASS: UBLKCP).
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
    } while (!done);
}

// Packed FP32 pairs (sm_100 FFMA2): one 64-bit register pair = two rays' values.
__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ void unpack2u(uint64_t v, uint32_t& lo, uint32_t& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {  // two IEEE fmaf
    uint64_t d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

template <int PX>
struct PixelState {
    uint64_t u2[PX / 2];  // camera u of rays (2j, 2j+1), packed for FFMA2
    float best[PX];
    // The winning primitive ids live in (dynamic) shared memory, one column per thread: they are read
    // and written only in the rare exact update and in the epilogue, and the 16 registers they would
    // pin for the whole dense loop are what lets the next group's table entries be prefetched.
    int* sid;
    int sid_stride;
    __device__ __forceinline__ int& id(int q) const { return sid[q * sid_stride]; }
    __device__ __forceinline__ float u(int q) const {
        float lo, hi;
        unpack2(u2[q >> 1], lo, hi);
        return (q & 1) ? hi : lo;
    }
};

// Reference-exact nearest-hit update of primitive `prim` on this thread's PX rays (the arithmetic
// of raycast_math.h that tests/emu replays).  Reached only from the rare "some ray of this warp
// is inside the primitive's conic bound" branch, so its records come straight from global/L2.
template <int PX>
__device__ __forceinline__ void exact_update_rec(const GridParams& p, int prim, const float4 r0, const float4 r1,
                                                 const float4 r2, float v, PixelState<PX>& st) {
    if (prim < p.n_cubes) {
        const RtbSlab s0 = {r0.x, r0.y, r0.z, r0.w}, s1 = {r1.x, r1.y, r1.z, r1.w}, s2 = {r2.x, r2.y, r2.z, r2.w};
        const float b0 = rtb_slab_base(s0, v), b1 = rtb_slab_base(s1, v), b2 = rtb_slab_base(s2, v);
#pragma unroll
        for (int q = 0; q < PX; ++q) {
            const float n0 = rtb_slab_near(s0, b0, st.u(q));
            const float n1 = rtb_slab_near(s1, b1, st.u(q));
            const float n2 = rtb_slab_near(s2, b2, st.u(q));
            const float f0 = n0 + s0.width, f1 = n1 + s1.width, f2 = n2 + s2.width;
            const float tn = fmaxf(fmaxf(n0, n1), n2);
            const float tf = fminf(fminf(f0, f1), f2);
            rtb_cube_update(tn, tf, prim, st.best[q], st.id(q));
        }
    } else if (prim < p.n_cubes + p.n_spheres) {
        const RtbSphere sp = {r0.x, r0.y, r0.z, r0.w};
        const float radv = rtb_sphere_radv(sp, v);
        const float wv = p.cam.tau_v * v;
#pragma unroll
        for (int q = 0; q < PX; ++q) {
            const float rad = rtb_sphere_rad(sp, radv, st.u(q));
            const float wpix = fmaf(p.cam.tau_u, st.u(q), wv);
            rtb_sphere_update(sp, rad, wpix, prim, st.best[q], st.id(q));
        }
    }
}

// The same update with the exact records fetched from global / L2 (cube: 3 slab records, sphere: 1).
template <int PX>
__device__ __forceinline__ void exact_update(const GridParams& p, int prim, float v, PixelState<PX>& st) {
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    if (prim < p.n_cubes) {
        exact_update_rec<PX>(p, prim, __ldg(p.cube_rec + 3 * (size_t)prim + 0), __ldg(p.cube_rec + 3 * (size_t)prim + 1),
                             __ldg(p.cube_rec + 3 * (size_t)prim + 2), v, st);
    } else if (prim < p.n_cubes + p.n_spheres) {
        exact_update_rec<PX>(p, prim, __ldg(p.sph_rec + (size_t)(prim - p.n_cubes)), z, z, v, st);
    }
}

// One primitive's bound on this thread's PX rays: t = (sA, rowW, rowC, -); returns m = min |w| over the rays.
// w = sA*u + rowW is ONE packed FFMA2 per ray pair (sm_100's FFMA2 with the pair operand HI_LO and
// scalar-broadcast coefficients), folded into the running minimum by a 3-input FMNMX with |.| on both new
// operands: 8 FFMA2 + 8 FMNMX3 for 16 rays.  "Some ray has q < 2" is then fmaf(m, m, rowC) < 2, ONE FMA and one
// compare per thread and primitive -- bit for bit the per-ray q test (raycast_math.h: q is monotone in |w|).
// NaNs drop out of fminf, Inf gives q = Inf: such rays fail, as their own q would.
// (Measured: keeping KQ of the 8 ray pairs on the per-ray q form -- a second FFMA2 and a LOP3 instead of the
// FMNMX3, to load the idle FMA pipe -- gives C3 0.610 / 0.612 / 0.611 / 0.625 / 0.636 / 0.682 / 0.715 ms for
// KQ = 0 / 1 / 2 / 3 / 4 / 6 / 8: every 2-ray instruction holds the dispatch port for 2 clocks whichever pipe it
// goes to, so the form with the fewest of them wins; profiles/r01_variant_sweep_kq.jsonl.)
template <int PX>
__device__ __forceinline__ float bound_min(const float4 t, const PixelState<PX>& st) {
    const uint64_t S2 = pack2(t.x, t.x), W2 = pack2(t.y, t.y);
    float m = __int_as_float(0x7f800000);
#pragma unroll
    for (int j = 0; j < PX / 2; ++j) {
        const uint64_t w2 = ffma2(st.u2[j], S2, W2);
        float lo, hi;
        unpack2(w2, lo, hi);
        m = fminf(fminf(m, fabsf(lo)), fabsf(hi));
    }
    return m;
}
__device__ __forceinline__ bool bound_passes(float m, float rowC) { return rtb_bound_q_of_w(m, rowC) < 2.0f; }

#ifdef RTB_EXPERIMENTS  // helpers of the round-1 kernel (row table built per CTA), kept for A/B sweeps only
// The vote of a group of G primitives and, behind it, the exact updates of those whose bound some ray of the
// warp passed.  m[g] = min |w| of this thread's rays, c[g] = the row's rowC (kept as values: the hot path is
// G FMA + G compares into one predicate, the per-primitive compares happen only behind the vote).
template <int PX, int G>
__device__ __forceinline__ void resolve_group(const GridParams& p, const float (&m)[G], const float (&c)[G], int first_prim,
                                              float v, PixelState<PX>& st) {
    bool any = false;  // (one compare on min_g q_g instead of G compares was measured slower: C3 0.586 -> 0.610 ms)
#pragma unroll
    for (int g = 0; g < G; ++g) any = any || bound_passes(m[g], c[g]);
    if (__any_sync(0xFFFFFFFFu, any)) {
#pragma unroll 1
        for (int g = 0; g < G; ++g) {
            float mg = m[0], cg = c[0];
#pragma unroll
            for (int k = 1; k < G; ++k) {
                mg = (g == k) ? m[k] : mg;
                cg = (g == k) ? c[k] : cg;
            }
            if (__any_sync(0xFFFFFFFFu, bound_passes(mg, cg))) exact_update<PX>(p, first_prim + g, v, st);
        }
    }
}

// The dense test without a row table: G primitives of a staged chunk against this thread's PX rays, the row
// terms (rowW, rowC) computed per thread and primitive.
template <int PX, int G>
__device__ __forceinline__ void bound_group(const GridParams& p, const float4* __restrict__ rec, int first_prim,
                                            float v, PixelState<PX>& st) {
    float m[G], c[G];
#pragma unroll
    for (int g = 0; g < G; ++g) {
        const float4 r0 = rec[2 * g];                                 // sA, k1, k0, e
        const float r1 = *reinterpret_cast<const float*>(rec + 2 * g + 1);  // cv
        const RtbBound b = {r0.x, r0.y, r0.z, r0.w, r1, 0.f, 0.f, 0.f};
        c[g] = rtb_bound_row_c(b, v);
        m[g] = bound_min<PX>(make_float4(b.sA, rtb_bound_row_w(b, v), c[g], 0.f), st);
    }
    resolve_group<PX, G>(p, m, c, first_prim, v, st);
}

// Same dense test with the per-row terms taken from the CTA's row table (built once per chunk by
// all threads, see the kernel): ONE LDS.128 (sA, rowW, rowC, -) per thread and primitive and no
// scalar FP32 work in the loop.  `tab` points at this thread's row of the first primitive; consecutive
// primitives are ROWS apart.
template <int PX, int G, int ROWS>
__device__ __forceinline__ void bound_group_tab(const GridParams& p, const float4* __restrict__ tab, int first_prim,
                                                float v, PixelState<PX>& st) {
    float m[G], c[G];
#pragma unroll
    for (int g = 0; g < G; ++g) {
        const float4 t = tab[g * ROWS];
        c[g] = t.z;
        m[g] = bound_min<PX>(t, st);
    }
    resolve_group<PX, G>(p, m, c, first_prim, v, st);
}

// The dense loop over one staged chunk, software-pipelined: the table entries of the next two
// primitives are requested (LDS.128 off a running pointer) while the current two are being tested, so no
// warp waits for shared memory at the top of a group.  Groups of 4 primitives per vote; this warp takes
// groups split, split+S, ...  (The loop counter and the clamp of the prefetch index stay integers: the
// compiler keeps them on the uniform datapath; a version that ran on pointers moved them to the ALU pipe
// and cost 2 %: C3 0.603 -> 0.624 ms.)
template <int PX, int ROWS, int S>
__device__ __forceinline__ void dense_chunk_tab(const GridParams& p, const float4* __restrict__ tab, int cnt,
                                                int first_prim, int split, float v, PixelState<PX>& st) {
    constexpr int G = 4;
    int j = split * G;
    if (j >= cnt) return;
    const float4* t = tab + j * ROWS;
    float4 c0 = t[0], c1 = t[ROWS];
#pragma unroll 1
    for (; j < cnt; j += S * G) {
        const float4 c2 = t[2 * ROWS], c3 = t[3 * ROWS];
        float m[G], c[G];
        c[0] = c0.z; c[1] = c1.z; c[2] = c2.z; c[3] = c3.z;
        m[0] = bound_min<PX>(c0, st);
        m[1] = bound_min<PX>(c1, st);
        // next group's first two entries; past the end the last group is read again (stays inside the table)
        const int jn = min(j + S * G, cnt - G);
        const float4* tn = tab + jn * ROWS;
        c0 = tn[0];
        c1 = tn[ROWS];
        m[2] = bound_min<PX>(c2, st);
        m[3] = bound_min<PX>(c3, st);
        resolve_group<PX, G>(p, m, c, first_prim + j, v, st);
        t = tn;
    }
}

#endif  // RTB_EXPERIMENTS

// Nearest-hit shading (rtb_shade) + 128-bit coalesced stores of one thread's PX pixels.  The colour
// loads of four pixels are issued together (predicated, no branch around them), so a thread waits
// for L1/L2 four times, not PX times.
template <int PX>
__device__ __forceinline__ void shade_and_store(const GridParams& p, PixelState<PX>& st, int x0, int lrow, int gy) {
    uint32_t rgba8[PX];
    int32_t rgba[PX][4];
    int ids[PX];
#pragma unroll
    for (int q0 = 0; q0 < PX; q0 += 4) {
        float4 col[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int q = q0 + i;
            ids[q] = (st.best[q] == RTB_MISS_T) ? -1 : st.id(q);
            const float4* src = (ids[q] < p.n_cubes) ? p.cube_col + ids[q] : p.sph_col + (ids[q] - p.n_cubes);
            col[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ids[q] >= 0) col[i] = __ldg(src);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int q = q0 + i;
            rtb_shade(st.best[q], col[i].x, col[i].y, col[i].z, rgba[q]);
            rgba8[q] = rtb_pack_rgba8(rgba[q]);
        }
    }
    // packed band buffer (row = local row) or a whole frame -- e.g. rank 0's, mapped over NVLink
    const size_t row_off = (size_t)(p.out_global_rows ? gy : lrow) * p.width;
    const bool full = p.vec_ok && (x0 + PX <= p.width);
    if (full) {
        if (p.out_rgba8) {
#pragma unroll
            for (int q = 0; q < PX; q += 4)
                *reinterpret_cast<uint4*>(p.out_rgba8 + row_off + x0 + q) =
                    make_uint4(rgba8[q], rgba8[q + 1], rgba8[q + 2], rgba8[q + 3]);
        }
        if (p.out_id) {
#pragma unroll
            for (int q = 0; q < PX; q += 4)
                *reinterpret_cast<int4*>(p.out_id + row_off + x0 + q) =
                    make_int4(ids[q], ids[q + 1], ids[q + 2], ids[q + 3]);
        }
        if (p.out_i32) {
#pragma unroll
            for (int q = 0; q < PX; ++q)
                p.out_i32[row_off + x0 + q] = make_int4(rgba[q][0], rgba[q][1], rgba[q][2], rgba[q][3]);
        }
    } else {
#pragma unroll
        for (int q = 0; q < PX; ++q) {
            if (x0 + q < p.width) {
                if (p.out_rgba8) p.out_rgba8[row_off + x0 + q] = rgba8[q];
                if (p.out_id) p.out_id[row_off + x0 + q] = ids[q];
                if (p.out_i32) {
                    int32_t* o = reinterpret_cast<int32_t*>(p.out_i32) + 4 * (row_off + x0 + q);
                    o[0] = rgba[q][0]; o[1] = rgba[q][1]; o[2] = rgba[q][2]; o[3] = rgba[q][3];
                }
            }
        }
    }
}

// The same for a warp of the 8 x 4 lane layout (dense and cull kernels), called by EVERY lane: when the whole warp lies
// inside the frame the rgba8 pixels leave as whole 512-byte rows.  A thread's 16 pixels are 64 contiguous bytes, so a
// per-thread STG.128 writes 32 separate 16-byte pieces per instruction (half sectors, 64 bytes apart); staged through
// shared memory and read back transposed, one instruction writes 4 rows x 128 contiguous bytes (whole sectors).  For the
// local L2 this halves the write requests; for a frame in ANOTHER GPU's memory (multi-GPU frames, stores over NVLink)
// it is the difference between 16-byte and 128-byte write packets.  `stage` = this warp's 4 x 32 uint4 (2 KB).
template <int PX>
__device__ __forceinline__ void shade_and_store_warp(const GridParams& p, PixelState<PX>& st, int x0, int lrow, int gy, bool inside,
                                                     uint4* __restrict__ stage) {
    static_assert(PX % 4 == 0 && PX <= 16, "the row-staged store handles whole uint4s, at most 4 x 32 of them per warp");
    constexpr int kRowU4 = 8 * PX / 4;   // uint4 per warp row (8 lanes x PX pixels)
    const bool full = p.vec_ok && (x0 + PX <= p.width);
    if (!__all_sync(0xFFFFFFFFu, inside && full && p.out_rgba8 != nullptr && p.out_i32 == nullptr)) {
        if (inside) shade_and_store<PX>(p, st, x0, lrow, gy);
        return;
    }
    const int lane = threadIdx.x & 31, lx = lane & 7, ly = lane >> 3;
    int ids[PX];
#pragma unroll
    for (int q0 = 0; q0 < PX; q0 += 4) {
        float4 col[4];
        uint32_t rgba8[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int q = q0 + i;
            ids[q] = (st.best[q] == RTB_MISS_T) ? -1 : st.id(q);
            const float4* src = (ids[q] < p.n_cubes) ? p.cube_col + ids[q] : p.sph_col + (ids[q] - p.n_cubes);
            col[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ids[q] >= 0) col[i] = __ldg(src);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int32_t rgba[4];
            rtb_shade(st.best[q0 + i], col[i].x, col[i].y, col[i].z, rgba);
            rgba8[i] = rtb_pack_rgba8(rgba);
        }
        stage[ly * kRowU4 + lx * (PX / 4) + (q0 >> 2)] = make_uint4(rgba8[0], rgba8[1], rgba8[2], rgba8[3]);
    }
    __syncwarp();
    // the warp's 4 rows are consecutive rows of the output (band_rows is a multiple of 4 and the warp's first row is one)
    const size_t row0 = (size_t)((p.out_global_rows ? gy : lrow) - ly) * p.width + (size_t)(x0 - lx * PX);
#pragma unroll
    for (int r = 0; r < PX / 4; ++r) {
        const int s = r * 32 + lane, wr = s / kRowU4, c = s - wr * kRowU4;
        *reinterpret_cast<uint4*>(p.out_rgba8 + row0 + (size_t)wr * p.width + c * 4) = stage[s];
    }
    if (p.out_id) {
        const size_t off = (size_t)(p.out_global_rows ? gy : lrow) * p.width + x0;
#pragma unroll
        for (int q = 0; q < PX; q += 4)
            *reinterpret_cast<int4*>(p.out_id + off + q) = make_int4(ids[q], ids[q + 1], ids[q + 2], ids[q + 3]);
    }
}

// ---- completion signals -------------------------------------------------------------------------
// progress: the launch's rows are cut into chunks [progress_bound[c], progress_bound[c+1]); the last CTA of a chunk
//   to finish publishes p.progress_epoch in the chunk's flag word, on which the copy stream waits
//   (cuStreamWaitValue32, rtb_render): finished rows leave for the host while the launch runs.
// done_signal: the last CTA of the whole launch adds 1 to a word that may live in ANOTHER GPU's
//   memory (rank 0's peer frame, over NVLink): rank 0's stream waits until every rank has signalled,
//   which replaces a collective as the "all bands have landed" fence of a multi-GPU frame.
// Counters are reset by the CTA that completes them: one launch per context at a time.
// Called by every thread of the CTA after its stores; tile_h = pixel rows per CTA.
__device__ __forceinline__ void signal_completion(const GridParams& p, int tile_h) {
    if (p.progress == nullptr && p.done_signal == nullptr) return;
    __syncthreads();  // every store of this CTA is ordered before thread 0's fence
    if (threadIdx.x != 0) return;
    // device scope orders the stores for this GPU's copy engine and stream waits (both read
    // through L2); stores into a peer's memory and the peer signal need system scope
    if (p.done_signal != nullptr || p.fence_system) __threadfence_system();
    else __threadfence();
    if (p.progress != nullptr) {
        const int row0 = blockIdx.y * tile_h;
        int chunk = 0;
        while (chunk + 1 < p.progress_chunks && row0 >= p.progress_bound[chunk + 1]) ++chunk;
        const int c0 = p.progress_bound[chunk], c1 = p.progress_bound[chunk + 1];
        const unsigned expected = gridDim.x * (unsigned)((c1 - c0 + tile_h - 1) / tile_h);
        if (atomicAdd(&p.progress[chunk], 1u) + 1u == expected) {
            p.progress[chunk] = 0u;
            if (p.fence_system) __threadfence_system();
            else __threadfence();
            *reinterpret_cast<volatile unsigned*>(&p.progress[kMaxProgressChunks + chunk]) = p.progress_epoch;
        }
    }
    if (p.done_signal != nullptr) {
        if (atomicAdd(p.done_counter, 1u) + 1u == gridDim.x * gridDim.y) {
            *p.done_counter = 0u;
            __threadfence_system();
            atomicAdd_system(p.done_signal, 1u);
        }
    }
}

// ---- the dense kernel -------------------------------------------------------------------------------
// One primitive's row entry (sAn, rWn) against this thread's PX rays, folded into running minima of |w'|: ONE
// packed FFMA2 per ray pair (sm_100's FFMA2 with the pair operand HI_LO and scalar-broadcast coefficients) and ONE
// 3-input FMNMX with |.| on both new operands.  The minima are kept in NCH independent chains (ray pair j feeds
// chain j % NCH): a single chain over a group's 8 primitives x 8 ray pairs is 64 dependent FMNMX3 long and the
// loop becomes latency-bound (measured: 2 chains 0.584 ms at C3, no faster than round 1's kernel although it
// issues 15 % fewer instructions).  NaNs drop out of fminf, Inf stays Inf: such rays fail, as their own test would.
#ifndef RTB_MIN_CHAINS
#define RTB_MIN_CHAINS 8
#endif
constexpr int kMinChains = RTB_MIN_CHAINS;
template <int PX>
__device__ __forceinline__ void row_min(float sAn, float rWn, float (&m)[kMinChains], const PixelState<PX>& st) {
    const uint64_t S2 = pack2(sAn, sAn), W2 = pack2(rWn, rWn);
#pragma unroll
    for (int j = 0; j < PX / 2; ++j) {
        const uint64_t w2 = ffma2(st.u2[j], S2, W2);
        float lo, hi;
        unpack2(w2, lo, hi);
        m[j % kMinChains] = fminf(fminf(m[j % kMinChains], fabsf(lo)), fabsf(hi));
    }
}
__device__ __forceinline__ float min_of_chains(const float (&m)[kMinChains]) {
    float r = m[0];
#pragma unroll
    for (int k = 1; k + 1 < kMinChains; k += 2) r = fminf(fminf(r, m[k]), m[k + 1]);
    if (kMinChains % 2 == 0) r = fminf(r, m[kMinChains - 1]);
    return r;
}

constexpr int kDenseGroup = 8;  // primitives per compare + warp vote

// Behind the vote (a few per mille of the groups): which of the group's primitives was it?  Per primitive the SPAN form
// of the same test (raycast_math.h, rtb_bound_span_m: w' at the thread's two end rays bounds every w' between them, u
// and w' being monotone along the run; the distance from 0 to that interval is min |w'| unless the run straddles 0) --
// one FFMA2 and three FMNMX per primitive instead of 8 + 8: it accepts every run the per-ray test accepts, so the warp
// still runs the reference-exact update wherever it has to, and the exact update is a no-op elsewhere.  Rolled: cold
// code, but it is what makes crowded tiles slow (CTA lifetimes 35 .. 112 us around a median of 71 at C3).
template <int PX>
__device__ __forceinline__ void dense_group_slow(const GridParams& p, const float2* __restrict__ ent, int first_prim, float v,
                                                 PixelState<PX>& st) {
    float u_first, u_last, unused;
    unpack2(st.u2[0], u_first, unused);
    unpack2(st.u2[PX / 2 - 1], unused, u_last);
    const uint64_t u_ends = pack2(u_first, u_last);
#pragma unroll 1
    for (int g = 0; g < kDenseGroup; ++g) {
        const float2 e = ent[g];
        float wa, wb;
        unpack2(ffma2(u_ends, pack2(e.x, e.x), pack2(e.y, e.y)), wa, wb);
        if (__any_sync(0xFFFFFFFFu, rtb_row_passes(rtb_bound_span_m(wa, wb)))) exact_update<PX>(p, first_prim + g, v, st);
    }
}

// The dense loop over one staged chunk: `row` = this thread's pixel row of the chunk in shared memory (cnt
// entries of 8 B, cnt a multiple of 8), read as LDS.128 = two primitives, 4 distinct addresses per warp in 4
// different bank groups.  Groups of 8 primitives share the running minima, one compare and one vote; warp `split`
// of the S warps of a pixel tile takes groups split, split + S, ...  The next group's entries are requested while
// the current group is tested.  (Two groups per trip with the table registers alternating -- no register copies at
// the end of a trip -- was measured SLOWER: C3 0.5574 -> 0.5663 ms, C4 8.11 -> 8.24 ms.)
// Measured (profiles/r02_dense_sweep_v16.jsonl), vote per 16 against vote per 8: C4 8.117 -> 7.927 ms (92 % of the issue
// bound), one rank's C4 rows at world 8 1.055 -> 1.030 ms, C3 0.5619 -> 0.5613 ms (its launch is short enough for the
// doubled cold path to eat the gain), C2 0.0381 -> 0.0388 ms.  -DRTB_DENSE_VOTE16=0 builds the vote-per-8 loop.
#ifndef RTB_DENSE_VOTE16
#define RTB_DENSE_VOTE16 1
#endif
template <int PX, int S>
__device__ __forceinline__ void dense_chunk(const GridParams& p, const float4* __restrict__ row, int cnt, int first_prim,
                                            int split, float v, PixelState<PX>& st) {
    constexpr int G = kDenseGroup;
    int j = split * G;
    if (j >= cnt) return;
    const float4* t = row + (j >> 1);
    float4 a0 = t[0], a1 = t[1], a2 = t[2], a3 = t[3];
#if RTB_DENSE_VOTE16
    // Two groups of 8 per trip share the running minima, ONE compare and ONE vote; the table registers alternate
    // (a -> b -> a), so nothing is copied at the end of a trip.  A trip whose second group lies past the end tests the
    // chunk's last group again (harmless: minima and exact updates are idempotent).
#pragma unroll 1
    for (; j < cnt; j += 2 * S * G) {
        float m[kMinChains];
#pragma unroll
        for (int k = 0; k < kMinChains; ++k) m[k] = __int_as_float(0x7f800000);
        const int jb = min(j + S * G, cnt - G);
        const float4* tb = row + (jb >> 1);
        row_min<PX>(a0.x, a0.y, m, st);
        row_min<PX>(a0.z, a0.w, m, st);
        const float4 b0 = tb[0], b1 = tb[1];
        row_min<PX>(a1.x, a1.y, m, st);
        row_min<PX>(a1.z, a1.w, m, st);
        const float4 b2 = tb[2], b3 = tb[3];
        row_min<PX>(a2.x, a2.y, m, st);
        row_min<PX>(a2.z, a2.w, m, st);
        row_min<PX>(a3.x, a3.y, m, st);
        row_min<PX>(a3.z, a3.w, m, st);
        const int jn = min(j + 2 * S * G, cnt - G);
        const float4* tn = row + (jn >> 1);
        row_min<PX>(b0.x, b0.y, m, st);
        row_min<PX>(b0.z, b0.w, m, st);
        a0 = tn[0]; a1 = tn[1];
        row_min<PX>(b1.x, b1.y, m, st);
        row_min<PX>(b1.z, b1.w, m, st);
        a2 = tn[2]; a3 = tn[3];
        row_min<PX>(b2.x, b2.y, m, st);
        row_min<PX>(b2.z, b2.w, m, st);
        row_min<PX>(b3.x, b3.y, m, st);
        row_min<PX>(b3.z, b3.w, m, st);
        if (__any_sync(0xFFFFFFFFu, rtb_row_passes(min_of_chains(m)))) {
#pragma unroll 1
            for (int hgrp = 0; hgrp < 2; ++hgrp)
                dense_group_slow<PX>(p, reinterpret_cast<const float2*>(hgrp ? tb : t), first_prim + (hgrp ? jb : j), v, st);
        }
        t = tn;
    }
#else
#pragma unroll 1
    for (; j < cnt; j += S * G) {
        float m[kMinChains];
#pragma unroll
        for (int k = 0; k < kMinChains; ++k) m[k] = __int_as_float(0x7f800000);
        // past the end the last group is read again (stays inside the chunk)
        const int jn = min(j + S * G, cnt - G);
        const float4* tn = row + (jn >> 1);
        row_min<PX>(a0.x, a0.y, m, st);
        row_min<PX>(a0.z, a0.w, m, st);
        const float4 b0 = tn[0], b1 = tn[1];
        row_min<PX>(a1.x, a1.y, m, st);
        row_min<PX>(a1.z, a1.w, m, st);
        const float4 b2 = tn[2], b3 = tn[3];
        row_min<PX>(a2.x, a2.y, m, st);
        row_min<PX>(a2.z, a2.w, m, st);
        row_min<PX>(a3.x, a3.y, m, st);
        row_min<PX>(a3.z, a3.w, m, st);
        if (__any_sync(0xFFFFFFFFu, rtb_row_passes(min_of_chains(m))))
            dense_group_slow<PX>(p, reinterpret_cast<const float2*>(t), first_prim + j, v, st);
        a0 = b0; a1 = b1; a2 = b2; a3 = b3;
        t = tn;
    }
#endif
}

// Tile = (WARPS_X * 8 * PX) x (WARPS_Y * 4) pixels.  S > 1: S warps share one 8*PX x 4 warp tile and split the
// primitive list between them (groups of 8, round robin); their nearest hits are merged through shared memory
// before shading.  Used when a launch has too few pixel tiles to load every scheduler evenly (the per-GPU share
// of a frame on 8 GPUs).  Nothing in the loop is CTA-wide: a warp waits on the chunk's `full` mbarrier (TMA
// completion), tests, and arrives on its `empty` mbarrier; thread 0 refills a stage one chunk later, when every
// warp has long released it.
template <int PX, int WARPS_X, int WARPS_Y, int S, int STAGES, int CHUNK, int MIN_CTAS>
__global__ void __launch_bounds__(32 * WARPS_X * WARPS_Y * S, MIN_CTAS)
raycast_dense_kernel(const GridParams p) {
    constexpr int kWarps = WARPS_X * WARPS_Y * S;
    constexpr int kThreads = 32 * kWarps;
    constexpr int kTileW = WARPS_X * 8 * PX;
    constexpr int kTileH = WARPS_Y * 4;
    static_assert(CHUNK % kDenseGroup == 0 && CHUNK % 2 == 0, "chunks hold whole vote groups");
    // a chunk row = CHUNK entries (8 B) + 16 B: consecutive rows start 4 banks apart, so the 4 rows a warp reads
    // with one LDS.128 never collide
    constexpr int kRowF4 = CHUNK / 2 + 1;
    constexpr int kStageF4 = kTileH * kRowF4;
    constexpr int kMergeBytes = (S > 1) ? 8 * WARPS_X * WARPS_Y * (S - 1) * PX * 32 : 0;
    constexpr int kTabBytes = STAGES * kStageF4 * (int)sizeof(float4);
    constexpr int kSmemBytes = kTabBytes > kMergeBytes ? kTabBytes : kMergeBytes;
    __shared__ __align__(128) unsigned char smem_raw[kSmemBytes];
    __shared__ __align__(8) uint64_t full_bar[STAGES];
    __shared__ __align__(8) uint64_t empty_bar[STAGES];
    __shared__ unsigned long long row_src[kTileH];  // entry index of each tile row's first entry in p.row_tab
    __shared__ __align__(16) uint4 store_stage[kWarps][4 * 32];  // shade_and_store_warp: a warp's rgba8 rows
    float4* tab = reinterpret_cast<float4*>(smem_raw);

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int split = warp % S, tile_warp = warp / S;  // the S warps of a tile are neighbours
    const int wx = tile_warp % WARPS_X, wy = tile_warp / WARPS_X;
    const int lx = lane & 7, ly = lane >> 3;
    const int x0 = blockIdx.x * kTileW + (wx * 8 + lx) * PX;
    const int tile_y = (int)blockIdx.y + p.tile_y0;          // tile row within the launch set of this frame
    const int lrow = tile_y * kTileH + wy * 4 + ly;  // row in the packed (local) output
    // interleaved row bands: local band lb -> global band lb*world + rank
    const int lb = lrow / p.band_rows;
    const int gy = (lb * p.band_world + p.band_rank) * p.band_rows + (lrow - lb * p.band_rows);

    const int n_pad = p.n_pad;  // entries per table row: a multiple of 8
    const int n_chunks = (n_pad + CHUNK - 1) / CHUNK;
#ifdef RTB_EXPERIMENTS
    unsigned long long* trace = p.trace ? p.trace + 4 * ((size_t)blockIdx.y * gridDim.x + blockIdx.x) : nullptr;
    if (trace && tid == 0) {
        unsigned long long t; unsigned sm;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
        trace[0] = t; trace[3] = sm;
    }
#endif

    if (tid < kTileH) {  // the table is indexed by FRAME row; rows past the frame reuse its last row (never stored)
        const int trow = tile_y * kTileH + tid;
        const int tlb = trow / p.band_rows;
        int tgy = (tlb * p.band_world + p.band_rank) * p.band_rows + (trow - tlb * p.band_rows);
        tgy = min(tgy, p.height - 1);
        row_src[tid] = (unsigned long long)tgy * (unsigned long long)n_pad;
    }
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], kWarps);
        }
        fence_mbar_init();
    }
    __syncthreads();

    auto issue = [&](int k) {  // called by thread 0 only
        const int s = k % STAGES;
        const int first = k * CHUNK;
        const int cnt = min(CHUNK, n_pad - first);
        const uint32_t bytes = (uint32_t)cnt * 8u;
        mbar_arrive_expect_tx(&full_bar[s], bytes * kTileH);
#pragma unroll
        for (int r = 0; r < kTileH; ++r)
            bulk_g2s(tab + s * kStageF4 + r * kRowF4, p.row_tab + row_src[r] + first, bytes, &full_bar[s]);
    };
    if (tid == 0) {
#pragma unroll 1
        for (int k = 0; k < STAGES && k < n_chunks; ++k) issue(k);
    }

    extern __shared__ int dyn_ids[];  // [PX][kThreads] winning primitive ids (see PixelState)
    PixelState<PX> st;
    st.sid = dyn_ids + tid;
    st.sid_stride = kThreads;
    const float fy = (float)gy;
    const float v = rtb_cam_v(p.cam, fy);
#pragma unroll
    for (int q = 0; q < PX; q += 2) {
        st.u2[q >> 1] = pack2(rtb_cam_u(p.cam, (float)(x0 + q), fy), rtb_cam_u(p.cam, (float)(x0 + q + 1), fy));
        st.best[q] = st.best[q + 1] = RTB_MISS_T;
        st.id(q) = -1;
        st.id(q + 1) = -1;
    }

    const float4* my_row = tab + (wy * 4 + ly) * kRowF4;
#pragma unroll 1
    for (int k = 0; k < n_chunks; ++k) {
        const int s = k % STAGES;
        mbar_wait(&full_bar[s], (uint32_t)((k / STAGES) & 1));
#ifdef RTB_EXPERIMENTS
        if (trace && tid == 0 && k == 0) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); trace[1] = t; }
#endif
        const int first = k * CHUNK;
        dense_chunk<PX, S>(p, my_row + s * kStageF4, min(CHUNK, n_pad - first), first, split, v, st);
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[s]);  // this warp is done with the stage
        // refill the stage of the PREVIOUS chunk: by now every warp has released it (else thread 0 waits here)
        if (tid == 0 && k >= 1 && k - 1 + STAGES < n_chunks) {
            mbar_wait(&empty_bar[(k - 1) % STAGES], (uint32_t)(((k - 1) / STAGES) & 1));
            issue(k - 1 + STAGES);
        }
    }

    // ---- merge of the primitive splits (S > 1) -----------------------------------------------
    const bool owner = (split == 0);  // the warp of a tile that shades and stores
    if constexpr (S > 1) {
        constexpr int kSlots = WARPS_X * WARPS_Y * (S - 1);
        float (*m_best)[PX][32] = reinterpret_cast<float (*)[PX][32]>(smem_raw);
        int (*m_id)[PX][32] = reinterpret_cast<int (*)[PX][32]>(smem_raw + sizeof(float) * kSlots * PX * 32);
        __syncthreads();  // every warp has consumed every chunk (all TMA writes have landed): the ring is free
        if (split > 0) {
#pragma unroll
            for (int q = 0; q < PX; ++q) {
                m_best[tile_warp * (S - 1) + split - 1][q][lane] = st.best[q];
                m_id[tile_warp * (S - 1) + split - 1][q][lane] = st.id(q);
            }
        }
        __syncthreads();
        if (owner) {
#pragma unroll 1
            for (int o = 0; o < S - 1; ++o) {
#pragma unroll
                for (int q = 0; q < PX; ++q) {
                    const float ob = m_best[tile_warp * (S - 1) + o][q][lane];
                    const int oi = m_id[tile_warp * (S - 1) + o][q][lane];
                    // same winner as one ascending pass: nearest t, ties to the lowest primitive index
                    const int mine = st.id(q);
                    if (ob < st.best[q] || (ob == st.best[q] && oi >= 0 && (mine < 0 || oi < mine))) {
                        st.best[q] = ob;
                        st.id(q) = oi;
                    }
                }
            }
        }
    }

    // ---- nearest-hit shading + stores -------------------------------------------------------
    shade_and_store_warp<PX>(p, st, x0, lrow, gy, owner && lrow < p.local_rows && gy < p.height && x0 < p.width, store_stage[warp]);
#ifdef RTB_EXPERIMENTS
    if (trace) {
        __syncthreads();
        if (tid == 0) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); trace[2] = t; }
    }
#endif

    signal_completion(p, kTileH);
}

#ifdef RTB_EXPERIMENTS  // the round-1 dense kernel (bound records staged, row table built by every CTA): A/B sweeps only
// S > 1: S warps share one pixel tile and split the primitive list between them (groups of G,
// round robin); their nearest hits are merged through shared memory before shading.  Used when a
// launch has too few pixel tiles to load every scheduler evenly (the per-GPU share of a frame on 8
// GPUs): the 4 warps of a CTA then sit on the 4 schedulers of an SM with equal work.
//
// TAB: the row terms of the bound (rowW, rowC: 4 FP32 ops per thread and primitive when every
// thread computes them for itself) come from a shared-memory table built once per chunk and CTA:
// entry (primitive, CTA row) = (sA, rowW, rowC, -), kTileH rows x kTabPrims primitives x 16 B per
// stage.  The values are those of rtb_bound_row_w / rtb_bound_row_c, computed once per row
// instead of once per thread (16 threads of a CTA share a row).
template <int PX, int WARPS_X, int WARPS_Y, int G = kBoundGroup, int MIN_CTAS = (PX >= 16 ? 3 : 4), int S = 1,
          bool TAB = false, int TP = kTabPrims, int PIPE = 1>
__global__ void __launch_bounds__(32 * WARPS_X * WARPS_Y * S, MIN_CTAS)
raycast_grid_kernel(const GridParams p) {
    constexpr int kTabPrims = TP;  // primitives per staged chunk of the row-table variants
    constexpr int kThreads = 32 * WARPS_X * WARPS_Y * S;
    constexpr int kTileW = WARPS_X * 8 * PX;
    constexpr int kTileH = WARPS_Y * 4;
    constexpr int kChunk = TAB ? kTabPrims : kChunkPrims;   // primitives per staged chunk
    static_assert(kBoundGroup % G == 0, "the host pads the bound list to kBoundGroup");
    static_assert(!TAB || (kThreads % kTileH == 0 && kTileH <= 8), "row table: a thread always builds the same row");
    // One raw buffer: [bound records, 2 stages][row table, 2 stages]; after the loop the primitive-split
    // merge area (S > 1) reuses it from offset 0.
    constexpr int kStageBytes = (int)sizeof(float4) * 2 * 2 * kChunk;
    constexpr int kTabBytes = TAB ? (int)sizeof(float4) * 2 * kTabPrims * kTileH : 0;
    constexpr int kMergeBytes = (S > 1) ? 8 * WARPS_X * WARPS_Y * (S - 1) * PX * 32 : 0;
    constexpr int kSmemBytes = (kStageBytes + kTabBytes > kMergeBytes) ? kStageBytes + kTabBytes : kMergeBytes;
    __shared__ __align__(128) unsigned char smem_raw[kSmemBytes];
    __shared__ __align__(8) uint64_t full_bar[2];
    float4 (*stage)[2 * kChunk] = reinterpret_cast<float4 (*)[2 * kChunk]>(smem_raw);
    float4 (*rowtab)[kTabPrims * kTileH] = reinterpret_cast<float4 (*)[kTabPrims * kTileH]>(smem_raw + kStageBytes);
    (void)kThreads;

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int split = warp % S, tile_warp = warp / S;  // the S warps of a tile are neighbours
    const int wx = tile_warp % WARPS_X, wy = tile_warp / WARPS_X;
    const int lx = lane & 7, ly = lane >> 3;
    const int x0 = blockIdx.x * kTileW + (wx * 8 + lx) * PX;
    const int lrow = blockIdx.y * kTileH + wy * 4 + ly;  // row in the packed (local) output
    // interleaved row bands: local band lb -> global band lb*world + rank
    const int lb = lrow / p.band_rows;
    const int gy = (lb * p.band_world + p.band_rank) * p.band_rows + (lrow - lb * p.band_rows);

    // bound records: one per primitive (cubes, then spheres), padded by the host to a multiple of G
    const int n_bound = p.n_bound;
    const int n_chunks = (n_bound + kChunk - 1) / kChunk;

    auto issue = [&](int k) {  // called by thread 0 only
        const int b = k & 1;
        const int first = k * kChunk;
        const int cnt = min(kChunk, n_bound - first);
        const uint32_t bytes = (uint32_t)cnt * 32u;
        mbar_arrive_expect_tx(&full_bar[b], bytes);
        bulk_g2s(&stage[b][0], p.bound_rec + 2 * (size_t)first, bytes, &full_bar[b]);
    };

    if (tid == 0) {
        mbar_init(&full_bar[0], 1);
        mbar_init(&full_bar[1], 1);
        fence_mbar_init();
    }
    __syncthreads();
    if (tid == 0) {
        if (n_chunks > 0) issue(0);
        if (n_chunks > 1) issue(1);
    }

    extern __shared__ int dyn_ids[];  // [PX][kThreads] winning primitive ids (see PixelState)
    PixelState<PX> st;
    st.sid = dyn_ids + tid;
    st.sid_stride = kThreads;
    const float fy = (float)gy;
    const float v = rtb_cam_v(p.cam, fy);
#pragma unroll
    for (int q = 0; q < PX; q += 2) {
        st.u2[q >> 1] = pack2(rtb_cam_u(p.cam, (float)(x0 + q), fy), rtb_cam_u(p.cam, (float)(x0 + q + 1), fy));
        st.best[q] = st.best[q + 1] = RTB_MISS_T;
        st.id(q) = -1;
        st.id(q + 1) = -1;
    }

    // row-table build: thread t always fills CTA row t % kTileH (kThreads is a multiple of kTileH)
    float tab_v = 0.f;
    if constexpr (TAB) {
        const int trow = blockIdx.y * kTileH + (tid % kTileH);
        const int tlb = trow / p.band_rows;
        const int tgy = (tlb * p.band_world + p.band_rank) * p.band_rows + (trow - tlb * p.band_rows);
        tab_v = rtb_cam_v(p.cam, (float)tgy);
    }

    for (int k = 0; k < n_chunks; ++k) {
        const int b = k & 1;
        mbar_wait(&full_bar[b], (uint32_t)((k >> 1) & 1));
        const int first = k * kChunk;
        const int cnt = min(kChunk, n_bound - first);
        if constexpr (TAB) {
            for (int i = tid; i < cnt * kTileH; i += kThreads) {
                const int prim = i / kTileH;
                const float4 r0 = stage[b][2 * prim];
                const float r1 = *reinterpret_cast<const float*>(&stage[b][2 * prim + 1]);
                const RtbBound bd = {r0.x, r0.y, r0.z, r0.w, r1, 0.f, 0.f, 0.f};
                rowtab[b][i] = make_float4(bd.sA, rtb_bound_row_w(bd, tab_v), rtb_bound_row_c(bd, tab_v), 0.f);
            }
            // table b is complete; stage[b] has been consumed and may be refilled.  Table b is next
            // rebuilt two chunks on, behind the barrier of chunk k+1, which a warp only reaches
            // after its dense loop over table b.
            __syncthreads();
            if (tid == 0 && k + 2 < n_chunks) issue(k + 2);
            const float4* tab = &rowtab[b][wy * 4 + ly];
            if constexpr (PIPE != 0 && G == 4) {
                dense_chunk_tab<PX, kTileH, S>(p, tab, cnt, first, split, v, st);
            } else {
#pragma unroll 1
                for (int j = split * G; j < cnt; j += S * G)
                    bound_group_tab<PX, G, kTileH>(p, tab + j * kTileH, first + j, v, st);
            }
        } else {
#pragma unroll 1
            for (int j = split * G; j < cnt; j += S * G) bound_group<PX, G>(p, &stage[b][2 * j], first + j, v, st);
            if (k + 2 < n_chunks) {
                __syncthreads();  // every warp is done reading stage[b] before TMA refills it
                if (tid == 0) issue(k + 2);
            }
        }
    }

    // ---- merge of the primitive splits (S > 1) -----------------------------------------------
    const bool owner = (split == 0);  // the warp of a tile that shades and stores
    if constexpr (S > 1) {
        constexpr int kSlots = WARPS_X * WARPS_Y * (S > 1 ? S - 1 : 1);
        float (*m_best)[PX][32] = reinterpret_cast<float (*)[PX][32]>(smem_raw);
        int (*m_id)[PX][32] = reinterpret_cast<int (*)[PX][32]>(smem_raw + sizeof(float) * kSlots * PX * 32);
        __syncthreads();  // every warp has left the dense loop: the staging area is free for the merge
        if (split > 0) {
#pragma unroll
            for (int q = 0; q < PX; ++q) {
                m_best[tile_warp * (S - 1) + split - 1][q][lane] = st.best[q];
                m_id[tile_warp * (S - 1) + split - 1][q][lane] = st.id(q);
            }
        }
        __syncthreads();
        if (owner) {
#pragma unroll
            for (int o = 0; o < S - 1; ++o) {
#pragma unroll
                for (int q = 0; q < PX; ++q) {
                    const float ob = m_best[tile_warp * (S - 1) + o][q][lane];
                    const int oi = m_id[tile_warp * (S - 1) + o][q][lane];
                    // same winner as one ascending pass: nearest t, ties to the lowest primitive index
                    const int mine = st.id(q);
                    if (ob < st.best[q] || (ob == st.best[q] && oi >= 0 && (mine < 0 || oi < mine))) {
                        st.best[q] = ob;
                        st.id(q) = oi;
                    }
                }
            }
        }
    }

    // ---- nearest-hit shading + stores -------------------------------------------------------
    if (owner && lrow < p.local_rows && gy < p.height && x0 < p.width) shade_and_store<PX>(p, st, x0, lrow, gy);

    signal_completion(p, kTileH);
}

#endif  // RTB_EXPERIMENTS

// ---- tile-culled variant (GridParams::tile_cull) ------------------------------------------------
// Same pixels, same arithmetic, far fewer tests: before a CTA's rays see a primitive, the CTA asks
// whether the primitive's conic bound can be passed by ANY ray of its 256 x 8 pixel tile
// (rtb_bound_tile_misses: a conservative statement about the very values the dense test would
// compute).  Every thread tests 4 primitives per batch (records straight from global / L2 into
// registers, the next batch's loads in flight during the tests: each record has exactly one
// reader, so there is nothing to share through shared memory); the survivors are compacted IN ORDER
// (ballots + cross-warp prefix: ties between primitives are broken by index, so the exact updates
// must stay in ascending order) into a list of bound records + primitive index, and only that list
// -- a few primitives per tile in the scenes of SURVEY.md s8d -- goes through the per-ray bound
// (same rtb_bound_row_w/_c/ FFMA2 arithmetic as the dense kernel) and the reference-exact update.
// Work per CTA: n_bound tile tests + survivors x rays ray tests, instead of n_bound x rays.
template <int PX, int WARPS_X, int WARPS_Y, int MIN_CTAS>
__global__ void __launch_bounds__(32 * WARPS_X * WARPS_Y, MIN_CTAS)
raycast_cull_kernel(const GridParams p) {
    constexpr int kWarps = WARPS_X * WARPS_Y;
    constexpr int kThreads = 32 * kWarps;
    constexpr int kTileW = WARPS_X * 8 * PX;
    constexpr int kTileH = WARPS_Y * 4;
    constexpr int kUnr = 4;                  // primitives per thread and batch
    constexpr int kBatch = kThreads * kUnr;
    constexpr int kListCap = 2 * kBatch;     // survivors held before a pass over the rays is forced
    __shared__ __align__(16) float4 list[2 * kListCap];  // (sA, k1, k0, e), (cv, primitive index, -, -)
    __shared__ int wcount[2][kUnr][kWarps];
    __shared__ float row_v[kTileH], row_ua[kTileH], row_ub[kTileH];

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int wx = warp % WARPS_X, wy = warp / WARPS_X;
    const int lx = lane & 7, ly = lane >> 3;
    const int x0 = blockIdx.x * kTileW + (wx * 8 + lx) * PX;
    const int lrow = blockIdx.y * kTileH + wy * 4 + ly;
    const int lb = lrow / p.band_rows;
    const int gy = (lb * p.band_world + p.band_rank) * p.band_rows + (lrow - lb * p.band_rows);
    const int n_bound = p.n_bound;
    const int n_batches = (n_bound + kBatch - 1) / kBatch;

    // first batch of records: in flight while the pixel state is set up
    float4 r0[kUnr];
    float cv[kUnr];
    auto load_batch = [&](int k) {
#pragma unroll
        for (int u = 0; u < kUnr; ++u) {
            const int i = k * kBatch + u * kThreads + tid;
            if (i < n_bound) {
                r0[u] = __ldg(p.bound_rec + 2 * (size_t)i);
                cv[u] = __ldg(reinterpret_cast<const float*>(p.bound_rec + 2 * (size_t)i + 1));
            }
        }
    };
    if (n_batches > 0) load_batch(0);

    if (tid < kTileH) {  // the tile's rows: v, and u of the first and the last pixel column
        const int trow = blockIdx.y * kTileH + tid;
        const int tlb = trow / p.band_rows;
        const float ty = (float)((tlb * p.band_world + p.band_rank) * p.band_rows + (trow - tlb * p.band_rows));
        row_v[tid] = rtb_cam_v(p.cam, ty);
        row_ua[tid] = rtb_cam_u(p.cam, (float)(blockIdx.x * kTileW), ty);
        row_ub[tid] = rtb_cam_u(p.cam, (float)(blockIdx.x * kTileW + kTileW - 1), ty);
    }

    extern __shared__ int dyn_ids[];
    PixelState<PX> st;
    st.sid = dyn_ids + tid;
    st.sid_stride = kThreads;
    const float fy = (float)gy;
    const float v = rtb_cam_v(p.cam, fy);
#pragma unroll
    for (int q = 0; q < PX; q += 2) {
        st.u2[q >> 1] = pack2(rtb_cam_u(p.cam, (float)(x0 + q), fy), rtb_cam_u(p.cam, (float)(x0 + q + 1), fy));
        st.best[q] = st.best[q + 1] = RTB_MISS_T;
        st.id(q) = -1;
        st.id(q + 1) = -1;
    }
    __syncthreads();
    // u is monotone in x on a row, so a row's u values lie between those of its two end pixels
    float u_min = row_ua[0], u_max = row_ua[0], v_min = row_v[0], v_max = row_v[0];
#pragma unroll
    for (int r = 0; r < kTileH; ++r) {
        u_min = fminf(u_min, fminf(row_ua[r], row_ub[r]));
        u_max = fmaxf(u_max, fmaxf(row_ua[r], row_ub[r]));
        v_min = fminf(v_min, row_v[r]);
        v_max = fmaxf(v_max, row_v[r]);
    }

    const uint32_t lanes_below = (1u << lane) - 1u;
    int total = 0;  // survivors in the list (the same value in every thread)
    for (int k = 0; k < n_batches; ++k) {
        const int b = k & 1;
        RtbBound bd[kUnr];
        uint32_t ballot[kUnr];
#pragma unroll
        for (int u = 0; u < kUnr; ++u) {
            bd[u].sA = r0[u].x; bd[u].k1 = r0[u].y; bd[u].k0 = r0[u].z; bd[u].e = r0[u].w; bd[u].cv = cv[u];
        }
        if (k + 1 < n_batches) load_batch(k + 1);
#pragma unroll
        for (int u = 0; u < kUnr; ++u) {
            const bool keep = (k * kBatch + u * kThreads + tid < n_bound) &&
                              !rtb_bound_tile_misses(bd[u], u_min, u_max, v_min, v_max);
            ballot[u] = __ballot_sync(0xFFFFFFFFu, keep);
            if (lane == 0) wcount[b][u][warp] = __popc(ballot[u]);
        }
        // counts visible (wcount[b] is next written two batches on, behind the barrier of batch
        // k+1, which a thread reaches only after reading this batch's counts)
        __syncthreads();
#pragma unroll
        for (int u = 0; u < kUnr; ++u) {
            int slot = total + __popc(ballot[u] & lanes_below);
#pragma unroll
            for (int w = 0; w < kWarps; ++w) {
                const int c = wcount[b][u][w];
                if (w < warp) slot += c;
                total += c;
            }
            if ((ballot[u] >> lane) & 1u) {
                list[2 * slot] = make_float4(bd[u].sA, bd[u].k1, bd[u].k0, bd[u].e);
                list[2 * slot + 1] = make_float4(bd[u].cv, __int_as_float(k * kBatch + u * kThreads + tid), 0.f, 0.f);
            }
        }
        if (total > kListCap - kBatch || k == n_batches - 1) {  // the next batch might not fit / no next batch
            __syncthreads();  // list complete
#pragma unroll 1
            for (int j = 0; j < total; ++j) {
                const float4 a = list[2 * j], c = list[2 * j + 1];
                const RtbBound s = {a.x, a.y, a.z, a.w, c.x, 0.f, 0.f, 0.f};
                const float4 t = make_float4(s.sA, rtb_bound_row_w(s, v), rtb_bound_row_c(s, v), 0.f);
                if (__any_sync(0xFFFFFFFFu, bound_passes(bound_min<PX>(t, st), t.z))) exact_update<PX>(p, __float_as_int(c.y), v, st);
            }
            total = 0;
            if (k != n_batches - 1) __syncthreads();  // every warp has read the list before it is refilled
        }
    }

    if (lrow < p.local_rows && gy < p.height && x0 < p.width) shade_and_store<PX>(p, st, x0, lrow, gy);
    signal_completion(p, kTileH);
}


// ---- binned variant (GridParams::tile_cull == 2) ------------------------------------------------
// The per-CTA scan above makes every one of a frame's CTAs read and test the whole primitive list
// (C3: 4050 CTAs x 1024 tile tests, 32 KB of records each, two batches deep).  Here the same test is
// made once per (tile, primitive) by a separate small launch that is organised the other way round,
// and the render CTAs only read the answer:
//
//   bin_tiles_kernel    warp = (one tile column, 32 vertically consecutive tiles, one word of 32
//                       primitives).  Lane l first tests primitive l of the word against the UNION of
//                       the 32 tiles' coordinate ranges (rtb_bound_tile_misses on a larger range is the
//                       same statement about more rays), then every lane tests the few primitives that
//                       survive (ballot; their records travel by shuffle) against its own tile and writes
//                       one 32-bit word: bins[word][tile], bit j = "primitive 32*word + j may be seen by
//                       a ray of this tile".  No atomics, no clearing pass, the same words on every run.
//   raycast_binned_kernel   CTA = tile.  Reads its words (one per thread), expands the set bits IN
//                       ASCENDING PRIMITIVE ORDER (CTA-wide prefix sum; the strict '<' of the nearest-hit
//                       update breaks ties by index) into a list in shared memory, all threads fetch the
//                       listed primitives' bound + exact records in ONE round of loads, and the warps run
//                       the per-ray bound and the reference-exact update from shared memory.  A warp none
//                       of whose rays passed any bound stores the miss colour without shading.
//
// The decision per (tile, primitive) is the one the per-CTA scan makes, so the frame is again the
// dense kernel's, byte for byte.
template <int TILE_W, int TILE_H>
__device__ __forceinline__ void tile_uv_range(const GridParams& p, int tx, int ty, float& u_min, float& u_max,
                                              float& v_min, float& v_max) {
    // ranges over the very values the render kernel computes for the tile's rays (rtb_tile_uv_range:
    // the extremes sit on the corners); frame rows of the tile's first and last local row
    const int ra = ty * TILE_H, rb = ra + TILE_H - 1;
    const int ba = ra / p.band_rows, bb = rb / p.band_rows;
    const float ya = (float)((ba * p.band_world + p.band_rank) * p.band_rows + (ra - ba * p.band_rows));
    const float yb = (float)((bb * p.band_world + p.band_rank) * p.band_rows + (rb - bb * p.band_rows));
    rtb_tile_uv_range(p.cam, (float)(tx * TILE_W), (float)(tx * TILE_W + TILE_W - 1), ya, yb, &u_min, &u_max, &v_min, &v_max);
}

constexpr int kBinWarps = 8;  // warps (= words) per CTA of the binning launch

template <int TILE_W, int TILE_H>
__global__ void __launch_bounds__(32 * kBinWarps) bin_tiles_kernel(const GridParams p, int tiles_x, int tiles_y) {
    // the render launch that follows may be scheduled now (programmatic dependent launch); it still waits
    // for this whole grid to finish before it reads a bin word
    asm volatile("griddepcontrol.launch_dependents;");
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int sy_count = (tiles_y + 31) / 32;
    const int tx = blockIdx.x / sy_count, sy = blockIdx.x - tx * sy_count;
    const int word = blockIdx.y * kBinWarps + warp;
    if (word >= p.bin_words) return;
    const int ty_raw = sy * 32 + lane;
    const int ty = min(ty_raw, tiles_y - 1);  // lanes past the last tile row repeat it (their word is not stored)
    float u_min, u_max, v_min, v_max;
    tile_uv_range<TILE_W, TILE_H>(p, tx, ty, u_min, u_max, v_min, v_max);
    float su_min = u_min, su_max = u_max, sv_min = v_min, sv_max = v_max;  // union over the warp's tiles
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        su_min = fminf(su_min, __shfl_xor_sync(0xFFFFFFFFu, su_min, d));
        su_max = fmaxf(su_max, __shfl_xor_sync(0xFFFFFFFFu, su_max, d));
        sv_min = fminf(sv_min, __shfl_xor_sync(0xFFFFFFFFu, sv_min, d));
        sv_max = fmaxf(sv_max, __shfl_xor_sync(0xFFFFFFFFu, sv_max, d));
    }
    const int prim = word * 32 + lane;
    RtbBound b = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    bool keep = false;
    if (prim < p.n_bound) {
        const float4 r0 = __ldg(p.bound_rec + 2 * (size_t)prim);
        b.sA = r0.x; b.k1 = r0.y; b.k0 = r0.z; b.e = r0.w;
        b.cv = __ldg(reinterpret_cast<const float*>(p.bound_rec + 2 * (size_t)prim + 1));
        keep = !rtb_bound_tile_misses(b, su_min, su_max, sv_min, sv_max);
    }
    uint32_t cand = __ballot_sync(0xFFFFFFFFu, keep);
    uint32_t out = 0u;
    while (cand) {
        const int j = __ffs(cand) - 1;
        cand &= cand - 1u;
        RtbBound c = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        c.sA = __shfl_sync(0xFFFFFFFFu, b.sA, j);
        c.k1 = __shfl_sync(0xFFFFFFFFu, b.k1, j);
        c.k0 = __shfl_sync(0xFFFFFFFFu, b.k0, j);
        c.e = __shfl_sync(0xFFFFFFFFu, b.e, j);
        c.cv = __shfl_sync(0xFFFFFFFFu, b.cv, j);
        if (!rtb_bound_tile_misses(c, u_min, u_max, v_min, v_max)) out |= 1u << j;
    }
    if (ty_raw < tiles_y) p.bins[(size_t)word * ((size_t)tiles_x * tiles_y) + (size_t)tx * tiles_y + ty_raw] = out;
}

// Shading + stores four pixels at a time, as a ROLLED loop (the nearest-hit registers are rotated down by
// four after each group so that the body always reads best[0..3]: the epilogue is 4x smaller, and this
// kernel's straight-line code has to stay inside the 32 KB instruction cache).  `hit` false = every pixel of
// the warp is a miss and `st` was never touched: constants only.
template <int PX>
__device__ __forceinline__ void store_group(const GridParams& p, size_t px_off, int x, bool full, const uint32_t rgba8[4],
                                            const int32_t rgba[4][4], const int ids[4], bool with_rgba8 = true) {
    if (full) {
        if (p.out_rgba8 && with_rgba8) *reinterpret_cast<uint4*>(p.out_rgba8 + px_off) = make_uint4(rgba8[0], rgba8[1], rgba8[2], rgba8[3]);
        if (p.out_id) *reinterpret_cast<int4*>(p.out_id + px_off) = make_int4(ids[0], ids[1], ids[2], ids[3]);
        if (p.out_i32) {
#pragma unroll
            for (int i = 0; i < 4; ++i) p.out_i32[px_off + i] = make_int4(rgba[i][0], rgba[i][1], rgba[i][2], rgba[i][3]);
        }
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (x + i < p.width) {
                if (p.out_rgba8 && with_rgba8) p.out_rgba8[px_off + i] = rgba8[i];
                if (p.out_id) p.out_id[px_off + i] = ids[i];
                if (p.out_i32) {
                    int32_t* o = reinterpret_cast<int32_t*>(p.out_i32) + 4 * (px_off + i);
                    o[0] = rgba[i][0]; o[1] = rgba[i][1]; o[2] = rgba[i][2]; o[3] = rgba[i][3];
                }
            }
        }
    }
}

// Called by EVERY lane of the warp (the per-group vote needs them all); `inside` false = this thread's pixels
// lie outside the frame or the launch's rows and nothing is stored for it.
// Row-staged rgba8 stores (LX = 2 lanes x 16 rows per warp, `stage` = the warp's 16 x 8 uint4 = 2 KB, nullptr = off):
// a thread's 16 pixels are 64 contiguous bytes, so its STG.128 per group of four pixels writes, warp-wide, 32 separate
// 16-byte pieces; with the whole warp inside the frame the groups go to shared memory instead and leave at the end as
// 4 instructions of 4 rows x 128 contiguous bytes each (whole sectors: half the L2 write requests, and 128-byte
// instead of 16-byte packets when the frame lives in another GPU's memory).
template <int PX, int LX>
__device__ __forceinline__ void shade_store_groups(const GridParams& p, PixelState<PX>& st, bool hit, bool inside, int x0,
                                                   int lrow, int gy, uint4* __restrict__ stage) {
    const size_t row_off = (size_t)(p.out_global_rows ? gy : lrow) * p.width + x0;
    const bool full = p.vec_ok && (x0 + PX <= p.width);
    int32_t miss[4];
    rtb_shade(RTB_MISS_T, 0.f, 0.f, 0.f, miss);
    const uint32_t miss8 = rtb_pack_rgba8(miss);
    const int lane = threadIdx.x & 31, lx = lane % LX, ly = lane / LX;
    const bool rows = PX == 16 && LX == 2 && stage != nullptr && p.out_rgba8 != nullptr &&
                      __all_sync(0xFFFFFFFFu, inside && full);
    // the transposed store of instruction r: this lane writes uint4 (r * 32 + lane) of the warp's 16 rows x 8 uint4
    auto row_store = [&](int r, const uint4 val) {
        const int s = r * 32 + lane, wr = s >> 3, c = s & 7;
        const int l = lrow - ly + wr;                               // local row of warp row wr
        int orow = l;
        if (p.out_global_rows) {
            const int lb = l / p.band_rows;
            orow = (lb * p.band_world + p.band_rank) * p.band_rows + (l - lb * p.band_rows);
        }
        *reinterpret_cast<uint4*>(p.out_rgba8 + (size_t)orow * p.width + (size_t)(x0 - lx * PX) + c * 4) = val;
    };
    if (!hit) {
        const uint32_t rgba8[4] = {miss8, miss8, miss8, miss8};
        const int ids[4] = {-1, -1, -1, -1};
        int32_t rgba[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { rgba[i][0] = miss[0]; rgba[i][1] = miss[1]; rgba[i][2] = miss[2]; rgba[i][3] = miss[3]; }
        if (rows) {
#pragma unroll
            for (int r = 0; r < 4; ++r) row_store(r, make_uint4(miss8, miss8, miss8, miss8));
            if (p.out_id || p.out_i32) {
#pragma unroll 1
                for (int q0 = 0; q0 < PX; q0 += 4) store_group<PX>(p, row_off + q0, x0 + q0, full, rgba8, rgba, ids, false);
            }
        } else if (inside) {
#pragma unroll 1
            for (int q0 = 0; q0 < PX; q0 += 4) store_group<PX>(p, row_off + q0, x0 + q0, full, rgba8, rgba, ids);
        }
        return;
    }
#pragma unroll 1
    for (int q0 = 0; q0 < PX; q0 += 4) {
        uint32_t rgba8[4];
        int32_t rgba[4][4];
        int ids[4];
        float4 col[4];
        const bool mine = inside && (st.best[0] != RTB_MISS_T || st.best[1] != RTB_MISS_T || st.best[2] != RTB_MISS_T ||
                                     st.best[3] != RTB_MISS_T);
        if (!__any_sync(0xFFFFFFFFu, mine)) {  // no lane of the warp has a hit among these four pixels
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                ids[i] = -1; rgba8[i] = miss8;
                rgba[i][0] = miss[0]; rgba[i][1] = miss[1]; rgba[i][2] = miss[2]; rgba[i][3] = miss[3];
            }
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                ids[i] = (st.best[i] == RTB_MISS_T) ? -1 : st.id(q0 + i);
                const float4* src = (ids[i] < p.n_cubes) ? p.cube_col + ids[i] : p.sph_col + (ids[i] - p.n_cubes);
                col[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (ids[i] >= 0) col[i] = __ldg(src);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                rtb_shade(st.best[i], col[i].x, col[i].y, col[i].z, rgba[i]);
                rgba8[i] = rtb_pack_rgba8(rgba[i]);
            }
        }
        if (rows) stage[ly * 8 + lx * 4 + (q0 >> 2)] = make_uint4(rgba8[0], rgba8[1], rgba8[2], rgba8[3]);
        if (inside) store_group<PX>(p, row_off + q0, x0 + q0, full, rgba8, rgba, ids, !rows);  // rows: rgba8 leaves at the end
#pragma unroll
        for (int i = 0; i + 4 < PX; ++i) st.best[i] = st.best[i + 4];
    }
    if (rows) {
        __syncwarp();
#pragma unroll
        for (int r = 0; r < 4; ++r) row_store(r, stage[r * 32 + lane]);
    }
}

// Geometry: a thread holds PX horizontally adjacent rays, a warp is LX lanes wide and 32 / LX rows tall
// (LX = 2, PX = 16: 32 x 16 pixels -- close to square, so that a primitive's exact update and shading run
// in few warps), a CTA = WARPS_X x WARPS_Y warps = one tile of the bins.
template <int PX, int LX, int WARPS_X, int WARPS_Y, int MIN_CTAS>
__global__ void __launch_bounds__(32 * WARPS_X * WARPS_Y, MIN_CTAS)
raycast_binned_kernel(const GridParams p) {
    constexpr int kWarps = WARPS_X * WARPS_Y;
    constexpr int kThreads = 32 * kWarps;
    constexpr int kWarpRows = 32 / LX;
    constexpr int kTileW = WARPS_X * LX * PX;
    constexpr int kTileH = WARPS_Y * kWarpRows;
    constexpr int kCap = 64;  // listed primitives per pass over the rays
    __shared__ __align__(16) float4 s_rec[kCap][5];  // bound record (2) + exact records (cube 3, sphere 1)
    __shared__ int s_prim[kCap];
    __shared__ int s_wsum[2][kWarps];
    __shared__ __align__(16) uint4 store_stage[kWarps][16 * 8];  // shade_store_groups: a warp's rgba8 rows (LX = 2)

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int wx = warp % WARPS_X, wy = warp / WARPS_X;
    const int lx = lane % LX, ly = lane / LX;
    const int x0 = blockIdx.x * kTileW + (wx * LX + lx) * PX;
    const int lrow = blockIdx.y * kTileH + wy * kWarpRows + ly;
    const int lb = lrow / p.band_rows;
    const int gy = (lb * p.band_world + p.band_rank) * p.band_rows + (lrow - lb * p.band_rows);
    const size_t n_tiles = (size_t)gridDim.x * gridDim.y;
    const size_t tile = (size_t)blockIdx.x * gridDim.y + blockIdx.y;
    const int n_words = p.bin_words;

    extern __shared__ int dyn_ids[];
    PixelState<PX> st;
    st.sid = dyn_ids + tid;
    st.sid_stride = kThreads;
    const float fy = (float)gy;
    const float v = rtb_cam_v(p.cam, fy);
    bool warp_hit = false;  // some ray run of this warp passed some bound: the pixel state `st` is set up
    // camera u of the thread's first and last ray, the values rtb_cam_u gives ((float)x0 + 15.0f is exact)
    const uint64_t u_ends = pack2(fmaf(p.cam.l11, (float)x0, p.cam.l12 * fy), fmaf(p.cam.l11, (float)x0 + (float)(PX - 1), p.cam.l12 * fy));

    // Launched as a programmatic dependent of bin_tiles_kernel: everything above overlapped with that
    // grid's tail; from here on its bin words (and everything earlier in the stream) are visible.
    asm volatile("griddepcontrol.wait;" ::: "memory");
    int chunk = 0;
    for (int base = 0; base < n_words; base += kThreads, ++chunk) {
        const int w = base + tid;
        const uint32_t m = (w < n_words) ? __ldcg(p.bins + (size_t)w * n_tiles + tile) : 0u;
        // CTA-wide exclusive prefix of the popcounts: list position of each thread's first primitive.
        // (s_wsum is double buffered: a warp that writes chunk k+2's sums has passed the barrier of chunk
        // k+1, which every thread reaches only after reading chunk k's.)
        const int c = __popc(m);
        int incl = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            if (lane >= d) incl += t;
        }
        if (lane == 31) s_wsum[chunk & 1][warp] = incl;
        __syncthreads();
        int off = incl - c, total = 0;
#pragma unroll
        for (int k = 0; k < kWarps; ++k) {
            const int s = s_wsum[chunk & 1][k];
            if (k < warp) off += s;
            total += s;
        }
        if (total == 0) continue;
        for (int pass = 0; pass < total; pass += kCap) {
            {   // the thread that owns a bit fetches that primitive's records (one round of loads per pass)
                uint32_t mm = m;
                int pos = off - pass;
                while (mm != 0u && pos < kCap) {
                    const int j = __ffs(mm) - 1;
                    mm &= mm - 1u;
                    if (pos >= 0) {
                        const int prim = w * 32 + j;
                        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
                        const float4 b0 = __ldg(p.bound_rec + 2 * (size_t)prim), b1 = __ldg(p.bound_rec + 2 * (size_t)prim + 1);
                        float4 e0 = z, e1 = z, e2 = z;
                        if (prim < p.n_cubes) {
                            e0 = __ldg(p.cube_rec + 3 * (size_t)prim);
                            e1 = __ldg(p.cube_rec + 3 * (size_t)prim + 1);
                            e2 = __ldg(p.cube_rec + 3 * (size_t)prim + 2);
                        } else if (prim < p.n_cubes + p.n_spheres) {
                            e0 = __ldg(p.sph_rec + (size_t)(prim - p.n_cubes));
                        }
                        s_prim[pos] = prim;
                        s_rec[pos][0] = b0; s_rec[pos][1] = b1; s_rec[pos][2] = e0; s_rec[pos][3] = e1; s_rec[pos][4] = e2;
                    }
                    ++pos;
                }
            }
            __syncthreads();
            const int n = min(kCap, total - pass);
#pragma unroll 1
            for (int e = 0; e < n; ++e) {
                const float4 a = s_rec[e][0];
                const RtbBound s = {a.x, a.y, a.z, a.w, s_rec[e][1].x, 0.f, 0.f, 0.f};
                // span form of the bound (rtb_bound_span_m): w at the thread's two end pixels, as one FFMA2
                const float rw = rtb_bound_row_w(s, v);
                float wa, wb;
                unpack2(ffma2(u_ends, pack2(s.sA, s.sA), pack2(rw, rw)), wa, wb);
                if (__any_sync(0xFFFFFFFFu, bound_passes(rtb_bound_span_m(wa, wb), rtb_bound_row_c(s, v)))) {
                    if (!warp_hit) {  // the warp's pixel state is set up when its first primitive turns up
                        const float fx0 = (float)x0, row_term = p.cam.l12 * fy;
#pragma unroll
                        for (int q = 0; q < PX; q += 2) {
                            // fmaf(l11, (float)(x0 + q), l12 * fy) = rtb_cam_u: (float)x0 + (float)q is exact
                            st.u2[q >> 1] = pack2(fmaf(p.cam.l11, fx0 + (float)q, row_term),
                                                  fmaf(p.cam.l11, fx0 + (float)(q + 1), row_term));
                            st.best[q] = st.best[q + 1] = RTB_MISS_T;  // ids are read only where best != miss
                        }
                        warp_hit = true;
                    }
                    exact_update_rec<PX>(p, s_prim[e], s_rec[e][2], s_rec[e][3], s_rec[e][4], v, st);
                }
            }
            if (pass + kCap < total || base + kThreads < n_words) __syncthreads();  // list is refilled
        }
    }

    shade_store_groups<PX, LX>(p, st, warp_hit, lrow < p.local_rows && gy < p.height && x0 < p.width, x0, lrow, gy,
                               LX == 2 ? store_stage[warp] : nullptr);
    signal_completion(p, kTileH);
}

}  // namespace

// ---- explicit-origin kernel (rtb_render_rays) --------------------------------------------------
// Argument-for-argument the OpenCL kernel's rayOrigins[] + rayDir form (rayTracer.cl:111-120).
// One thread per ray; world-space slab records; spheres use the reference's float formula in
// the reference's operation order (bit-identical to MainState::intersectSphere).
namespace {

__global__ void __launch_bounds__(256) raycast_rays_kernel(const RaysParams p) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n_rays) return;
    const float4 o = __ldg(p.origins + i);
    float best = RTB_MISS_T;
    int id = -1;
    for (int c = 0; c < p.n_cubes; ++c) {
        const float4 r0 = __ldg(p.cube_rec + 4 * c + 0), r1 = __ldg(p.cube_rec + 4 * c + 1),
                     r2 = __ldg(p.cube_rec + 4 * c + 2), w = __ldg(p.cube_rec + 4 * c + 3);
        const float n0 = fmaf(r0.x, o.x, fmaf(r0.y, o.y, fmaf(r0.z, o.z, r0.w)));
        const float n1 = fmaf(r1.x, o.x, fmaf(r1.y, o.y, fmaf(r1.z, o.z, r1.w)));
        const float n2 = fmaf(r2.x, o.x, fmaf(r2.y, o.y, fmaf(r2.z, o.z, r2.w)));
        const float tn = fmaxf(fmaxf(n0, n1), n2);
        const float tf = fminf(fminf(n0 + w.x, n1 + w.y), n2 + w.z);
        rtb_cube_update(tn, tf, c, best, id);
    }
    const float o4[4] = {o.x, o.y, o.z, o.w};
    for (int s = 0; s < p.n_spheres; ++s) {
        const float4 c4 = __ldg(p.sph_rec + 2 * s), q = __ldg(p.sph_rec + 2 * s + 1);
        const float cc[4] = {c4.x, c4.y, c4.z, c4.w};
        const float t0 = rtb_sphere_t_reference(o4, p.dir, q.x, cc);
        if (t0 != 0.0f && t0 < best) { best = t0; id = p.n_cubes + s; }
    }
    float cr = 0.f, cg = 0.f, cb = 0.f;
    if (best == RTB_MISS_T) id = -1;
    if (id >= 0) {
        const float4 col = (id < p.n_cubes) ? __ldg(p.cube_col + id) : __ldg(p.sph_col + (id - p.n_cubes));
        cr = col.x; cg = col.y; cb = col.z;
    }
    int32_t rgba[4];
    rtb_shade(best, cr, cg, cb, rgba);
    if (p.out_rgba8) p.out_rgba8[i] = rtb_pack_rgba8(rgba);
    if (p.out_id) p.out_id[i] = id;
    if (p.out_i32) p.out_i32[i] = make_int4(rgba[0], rgba[1], rgba[2], rgba[3]);
}

template <int PX, int WX, int WY>
GridLaunchShape shape_of(int width, int local_rows) {
    GridLaunchShape s;
    s.px = PX; s.wx = WX; s.wy = WY; s.split = 1;
    s.block = dim3(32 * WX * WY);
    const int tw = WX * 8 * PX, th = WY * 4;
    s.grid = dim3((width + tw - 1) / tw, (local_rows + th - 1) / th);
    s.smem = 0;
    return s;
}

}  // namespace

// Tilings of the dense kernel: rays per thread, warp tiles per CTA (x, y), warps per warp tile (primitive split).
// The product uses the first for every launch: 256 x 4 pixels, TWO warps, 8 CTAs per SM.  Measured on C2 / C3 / C4 for one
// rank's rows at world 1 / 2 / 4 / 8 (profiles/r02_dense_sweep_t3.jsonl, r02_dense_sweep_c2_t3.jsonl) it is the fastest or
// within 2 % of the fastest everywhere: the work per warp is that of the 4-warp CTA (same table traffic, same loop), in
// CTAs of half the size, so that a launch of few waves spreads evenly over the SMs (C3 at world 8: 0.0905 ms against
// 0.0921 for 4-warp CTAs and 0.0954 for the primitive-split variant; C4 at world 8: 1.053 against 1.135 ms).  The CTA
// timeline (tools/cta_trace.py, profiles/r02_cta_trace_c3.jsonl) shows what is left: CTA lifetimes spread from 35 to
// 112 us around a median of 71 us (the exact updates behind the votes), so the SMs run dry over the last ~20 us of
// any launch.  The other tilings exist in the experiments build only (RTB_DENSE_TILING = 1 / 2 / 3).
struct Tiling { int px, wx, wy, split; };
static const Tiling kTilings[] = {
    {16, 2, 1, 1},  // 256 x 4 pixels, 2 warps
#ifdef RTB_EXPERIMENTS
    {16, 2, 2, 1},  // 256 x 8 pixels, 4 warps (round 1's and v10's CTA)
    {16, 2, 1, 2},  // 256 x 4 pixels, 4 warps, two per warp tile: nearest-hit merge through shared memory
    {16, 1, 1, 4},  // 128 x 4 pixels, 4 warps on one warp tile
    {8, 2, 1, 1},   // 128 x 4 pixels, 2 warps, 8 rays per thread: half the work per CTA
#endif
};

GridLaunchShape grid_launch_shape(int width, int local_rows, int sm_count) {
    (void)sm_count;
    int pick = 0;
#ifdef RTB_EXPERIMENTS
    static const int forced = [] { const char* e = std::getenv("RTB_DENSE_TILING"); return e ? std::atoi(e) : -1; }();
    if (forced >= 0 && forced < (int)(sizeof kTilings / sizeof kTilings[0])) pick = forced;
#endif
    const Tiling& t = kTilings[pick];
    const int tw = t.wx * 8 * t.px, th = t.wy * 4;
    GridLaunchShape s;
    s.px = t.px; s.wx = t.wx; s.wy = t.wy; s.split = t.split;
    s.block = dim3(32 * t.wx * t.wy * t.split);
    s.grid = dim3((width + tw - 1) / tw, (local_rows + th - 1) / th);
    s.smem = 0;
    return s;
}

// Launch of one instantiation: the id columns are dynamic shared memory (PX x threads ints); the opt-in
// attribute is set once per instantiation and device.
#ifndef RTB_DENSE_STAGES
#define RTB_DENSE_STAGES 2
#endif
#ifndef RTB_DENSE_CHUNK
#define RTB_DENSE_CHUNK 256
#endif
constexpr int kDenseStages = RTB_DENSE_STAGES;  // ring depth of the row-table chunks
constexpr int kDenseChunk = RTB_DENSE_CHUNK;    // primitives per chunk: 2 KB per pixel row and stage (C3: 128 x 4 stages 0.572 ms, 256 x 2 0.561)
template <int PX, int WX, int WY, int S, int STAGES, int CHUNK, int MINB>
cudaError_t launch_dense_instance(const GridParams& p, dim3 grid, cudaStream_t stream) {
    constexpr int kThreads = 32 * WX * WY * S;
    constexpr int kDyn = PX * kThreads * (int)sizeof(int);
    static thread_local int configured_device = -1;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (configured_device != dev) {
        e = cudaFuncSetAttribute(raycast_dense_kernel<PX, WX, WY, S, STAGES, CHUNK, MINB>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, kDyn);
        if (e != cudaSuccess) return e;
        configured_device = dev;
    }
#ifdef RTB_EXPERIMENTS
    // occupancy experiment: RTB_DENSE_EXTRA_SMEM bytes of unused dynamic shared memory per CTA (fewer resident CTAs)
    static const int extra = [] { const char* x = std::getenv("RTB_DENSE_EXTRA_SMEM"); return x ? std::atoi(x) : 0; }();
    if (extra > 0) {
        e = cudaFuncSetAttribute(raycast_dense_kernel<PX, WX, WY, S, STAGES, CHUNK, MINB>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, kDyn + extra);
        if (e != cudaSuccess) return e;
        raycast_dense_kernel<PX, WX, WY, S, STAGES, CHUNK, MINB><<<grid, dim3(kThreads), kDyn + extra, stream>>>(p);
        return cudaGetLastError();
    }
#endif
    raycast_dense_kernel<PX, WX, WY, S, STAGES, CHUNK, MINB><<<grid, dim3(kThreads), kDyn, stream>>>(p);
    return cudaGetLastError();
}

#ifdef RTB_EXPERIMENTS
template <int PX, int WX, int WY, int G, int MINB, int S, bool TAB, int TP, int PIPE>
cudaError_t launch_instance(const GridParams& p, dim3 grid, cudaStream_t stream) {
    constexpr int kThreads = 32 * WX * WY * S;
    constexpr int kDyn = PX * kThreads * (int)sizeof(int);
    static thread_local int configured_device = -1;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (configured_device != dev) {
        e = cudaFuncSetAttribute(raycast_grid_kernel<PX, WX, WY, G, MINB, S, TAB, TP, PIPE>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, kDyn);
        if (e != cudaSuccess) return e;
        configured_device = dev;
    }
    raycast_grid_kernel<PX, WX, WY, G, MINB, S, TAB, TP, PIPE><<<grid, dim3(kThreads), kDyn, stream>>>(p);
    return cudaGetLastError();
}

template <int PX, int WX, int WY, int G, int MINB, int S = 1, bool TAB = false, int TP = kTabPrims, int PIPE = 1>
cudaError_t launch_variant(const GridParams& p, cudaStream_t stream) {
    GridLaunchShape s = shape_of<PX, WX, WY>(p.width, p.local_rows);
    if (s.grid.x == 0 || s.grid.y == 0) return cudaSuccess;
    return launch_instance<PX, WX, WY, G, MINB, S, TAB, TP, PIPE>(p, s.grid, stream);
}
#endif  // RTB_EXPERIMENTS

template <int PX, int WX, int WY, int MINB>
cudaError_t launch_cull_instance(const GridParams& p, cudaStream_t stream) {
    constexpr int kThreads = 32 * WX * WY;
    constexpr int kDyn = PX * kThreads * (int)sizeof(int);
    const GridLaunchShape s = shape_of<PX, WX, WY>(p.width, p.local_rows);
    if (s.grid.x == 0 || s.grid.y == 0) return cudaSuccess;
    static thread_local int configured_device = -1;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (configured_device != dev) {
        e = cudaFuncSetAttribute(raycast_cull_kernel<PX, WX, WY, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, kDyn);
        if (e != cudaSuccess) return e;
        configured_device = dev;
    }
    raycast_cull_kernel<PX, WX, WY, MINB><<<s.grid, dim3(kThreads), kDyn, stream>>>(p);
    return cudaGetLastError();
}

// Tile geometry of the binned render (tile_cull == 2): 64 x 32 pixel tiles, four warps of 32 x 16 pixels.  The
// bins are laid out for exactly the tiling of the render kernel that reads them.  (Measured and dropped,
// profiles/r01_variant_sweep_binned_v3_tilings.jsonl, C3 / C4 ms against 0.0320 / 0.102: 128 x 16 tiles 0.0339 / 0.109;
// 64 x 16 tiles of two 32 x 16 warps 0.0356 / 0.132; of two 64 x 8 warps 0.0376 / 0.136; one warp per 32 x 16 tile
// 0.0392 / 0.190; per-warp lists instead of one per CTA 0.0333 / 0.134.)
constexpr int kBinTileW = 64, kBinTileH = 32;
static bool pdl_enabled() {  // RTB_NO_PDL=1: plain stream order between the two launches (fallback knob)
    static const bool on = [] { const char* e = std::getenv("RTB_NO_PDL"); return !(e && e[0] == '1'); }();
    return on;
}

int binned_tile_rows() { return kBinTileH; }

size_t grid_bin_words(int width, int local_rows, int n_bound) {
    const size_t tiles = (size_t)((width + kBinTileW - 1) / kBinTileW) * (size_t)((local_rows + kBinTileH - 1) / kBinTileH);
    return tiles * (size_t)((n_bound + 31) / 32);
}

template <int PX, int LX, int WX, int WY, int MINB>
static cudaError_t launch_binned(const GridParams& p, cudaStream_t stream) {
    constexpr int kThreads = 32 * WX * WY;
    constexpr int kDyn = PX * kThreads * (int)sizeof(int);
    constexpr int kTileW = WX * LX * PX, kTileH = WY * (32 / LX);
    static_assert(kTileW == kBinTileW && kTileH == kBinTileH, "bins are laid out for this tiling");
    const int tiles_x = (p.width + kTileW - 1) / kTileW, tiles_y = (p.local_rows + kTileH - 1) / kTileH;
    if (tiles_x <= 0 || tiles_y <= 0) return cudaSuccess;
    if (p.bin_words > 0) {
        if (p.bins == nullptr) return cudaErrorInvalidValue;
        dim3 bgrid((unsigned)(tiles_x * ((tiles_y + 31) / 32)), (unsigned)((p.bin_words + kBinWarps - 1) / kBinWarps));
        bin_tiles_kernel<kTileW, kTileH><<<bgrid, 32 * kBinWarps, 0, stream>>>(p, tiles_x, tiles_y);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)tiles_x, (unsigned)tiles_y);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = kDyn;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, raycast_binned_kernel<PX, LX, WX, WY, MINB>, p);
}

#ifdef RTB_EXPERIMENTS
namespace { unsigned long long* g_trace = nullptr; size_t g_trace_cap = 0; }
unsigned long long* cta_trace_buffer(size_t ctas) {
    if (!std::getenv("RTB_CTA_TRACE")) return nullptr;
    if (g_trace_cap < ctas) {
        if (g_trace) cudaFree(g_trace);
        g_trace_cap = ctas;
        if (cudaMalloc(&g_trace, 32 * ctas) != cudaSuccess) { g_trace = nullptr; g_trace_cap = 0; return nullptr; }
    }
    cudaMemset(g_trace, 0, 32 * ctas);
    return g_trace;
}
int cta_trace_dump(size_t ctas) {
    const char* path = std::getenv("RTB_CTA_TRACE");
    if (!path || !g_trace) return 0;
    if (cudaDeviceSynchronize() != cudaSuccess) return 1;
    unsigned long long* h = new unsigned long long[4 * ctas];
    cudaMemcpy(h, g_trace, 32 * ctas, cudaMemcpyDeviceToHost);
    FILE* f = std::fopen(path, "a");
    if (f) {
        std::fprintf(f, "launch %zu\n", ctas);
        for (size_t i = 0; i < ctas; ++i) std::fprintf(f, "%llu %llu %llu %llu\n", h[4 * i], h[4 * i + 1], h[4 * i + 2], h[4 * i + 3]);
        std::fclose(f);
    }
    delete[] h;
    return 0;
}
#endif

cudaError_t launch_grid(const GridParams& p, int sm_count, cudaStream_t stream, const TailLane* tail) {
    if (p.tile_cull == 2) return launch_binned<16, 2, 2, 2, 8>(p, stream);  // 64 registers, 8 CTAs/SM
    if (p.tile_cull) return launch_cull_instance<16, 2, 2, 4>(p, stream);
#ifdef RTB_EXPERIMENTS
    static const int variant = [] {
        const char* e = std::getenv("RTB_GRID_VARIANT");
        return e ? std::atoi(e) : 0;
    }();
    switch (variant) {  // the round-1 kernel: bound records staged, row table built per CTA (tools/variant_sweep.py)
        case 10: return launch_variant<16, 2, 2, 4, 4, 1, true>(p, stream);  // round-1 production tiling at C3
        case 20: return launch_variant<16, 2, 1, 4, 4, 2, true>(p, stream);  // 2 warps per 128x4 tile
        default: break;
    }
#endif
    if (p.row_tab == nullptr) return cudaErrorInvalidValue;
    const GridLaunchShape s = grid_launch_shape(p.width, p.local_rows, sm_count);
    if (s.grid.x == 0 || s.grid.y == 0) return cudaSuccess;
#ifdef RTB_EXPERIMENTS
    GridParams q = p;
    const size_t ctas = (size_t)s.grid.x * s.grid.y;
    const bool traced = std::getenv("RTB_CTA_TRACE") != nullptr;
    q.trace = traced ? cta_trace_buffer(ctas) : nullptr;
    cudaError_t e;
    if (s.px == 8) e = launch_dense_instance<8, 2, 1, 1, kDenseStages, kDenseChunk, 10>(q, s.grid, stream);
    else if (s.split == 1 && s.wy == 1) e = launch_dense_instance<16, 2, 1, 1, kDenseStages, kDenseChunk, 8>(q, s.grid, stream);
    else if (s.split == 1) e = launch_dense_instance<16, 2, 2, 1, kDenseStages, kDenseChunk, 4>(q, s.grid, stream);
    else if (s.split == 2) e = launch_dense_instance<16, 2, 1, 2, kDenseStages, kDenseChunk, 4>(q, s.grid, stream);
    else e = launch_dense_instance<16, 1, 1, 4, kDenseStages, kDenseChunk, 4>(q, s.grid, stream);
    if (traced) cta_trace_dump(ctas);
    return e;
#else
    // Short tail.  The CTAs of a launch live ~70 us each (C3) and run out of phase, so when the grid is exhausted the
    // SM slots go idle one by one over a whole CTA lifetime: ~35 us per slot, 6 % of a C3 frame, 12 % of one rank's
    // half of it.  With a tail lane the last tile rows (one wave's worth of work, at most half of the rows) go to CTAs
    // of 8 rays per thread -- half the lifetime, ~6 % more instructions per test -- launched on a second, low-priority
    // stream at the same time: the block scheduler starts them when the body's grid is exhausted, and the frame's
    // stream joins them.  Measured on one rank's rows of C3 at world 1 / 2 / 4: 0.5558 -> 0.5438, 0.2956 -> 0.2878,
    // 0.1663 -> 0.1581 ms; C4 (4096 primitives: the 8-ray CTAs cost 8 % there and re-read twice the table) 7.948 -> 7.949,
    // 3.991 -> 4.000, 2.040 -> 2.052, 1.031 -> 1.066 ms (profiles/r02_dense_tail_sweep.txt): used up to 2048 primitives.
    // Only for launches without in-kernel progress / completion signals (their counters count the CTAs of one grid).
    // RTB_DENSE_TAIL=0 switches it off.
    static const bool tail_on = [] { const char* e = std::getenv("RTB_DENSE_TAIL"); return !(e && e[0] == '0'); }();
    const int slots = 8 * (sm_count > 0 ? sm_count : 1);
    const int strips = (int)s.grid.y;
    int tail_strips = (tail && tail_on && p.progress == nullptr && p.done_signal == nullptr && p.n_pad <= 2048)
                          ? (slots + (int)s.grid.x - 1) / (int)s.grid.x : 0;
    const long long ctas16 = (long long)strips * s.grid.x;
    if (tail_strips > 0 && ctas16 <= (long long)slots + slots / 4) {
        // A launch of about one wave or less: every CTA starts at once and the SMs hold 6 or 7 of them.  Between half a
        // wave and one, as many tile rows as the free slots allow become 8-ray CTAs (two per tile): every SM then holds
        // 8 CTAs of two sizes and the load evens out (one rank's 272 rows of C3 at world 8: 0.0917 -> 0.0847 ms;
        // profiles/r02_dense_tail_sweep.txt).  Not worth a second launch for short kernels (few primitives).
        tail_strips = (ctas16 >= slots / 2 && ctas16 < slots && p.n_pad >= 512) ? (int)((slots - ctas16) / s.grid.x) : 0;
    }
    if (tail_strips > strips / 2) tail_strips = strips / 2;
    if (tail_strips <= 0 || p.width <= 128)
        return launch_dense_instance<16, 2, 1, 1, kDenseStages, kDenseChunk, 8>(p, s.grid, stream);
    cudaError_t e = cudaEventRecord(tail->fork, stream);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(tail->stream, tail->fork, 0);
    if (e != cudaSuccess) return e;
    GridParams body = p, tl = p;
    body.tile_y0 = 0;
    tl.tile_y0 = strips - tail_strips;
    e = launch_dense_instance<16, 2, 1, 1, kDenseStages, kDenseChunk, 8>(body, dim3(s.grid.x, (unsigned)(strips - tail_strips)), stream);
    if (e != cudaSuccess) return e;
    e = launch_dense_instance<8, 2, 1, 1, kDenseStages, kDenseChunk, 10>(tl, dim3((unsigned)((p.width + 127) / 128), (unsigned)tail_strips), tail->stream);
    if (e != cudaSuccess) return e;
    e = cudaEventRecord(tail->join, tail->stream);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(stream, tail->join, 0);
    return e;
#endif
}

cudaError_t launch_rays(const RaysParams& p, cudaStream_t stream) {
    if (p.n_rays == 0) return cudaSuccess;
    const unsigned blocks = (unsigned)((p.n_rays + 255) / 256);
    raycast_rays_kernel<<<blocks, 256, 0, stream>>>(p);
    return cudaGetLastError();
}

// ---- band de-interleave (multi-GPU gather epilogue) --------------------------------------------
namespace {
__global__ void __launch_bounds__(256) assemble_bands_kernel(const AssembleParams p) {
    const int y = blockIdx.y;
    const int b = y / p.band_rows;
    const int r = b % p.band_world;
    const int lrow = (b / p.band_world) * p.band_rows + (y - b * p.band_rows);
    const uint8_t* s = p.src + (size_t)(p.row_off[r] + lrow) * p.row_bytes;
    uint8_t* d = p.dst + (size_t)y * p.row_bytes;
    const bool vec = ((p.row_bytes & 15) == 0) && ((reinterpret_cast<uintptr_t>(p.src) & 15) == 0) &&
                     ((reinterpret_cast<uintptr_t>(p.dst) & 15) == 0);
    if (vec) {
        const size_t n = p.row_bytes >> 4;
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
            reinterpret_cast<uint4*>(d)[i] = __ldg(reinterpret_cast<const uint4*>(s) + i);
    } else {
        const size_t n = p.row_bytes >> 2;  // elements are 4 or 16 bytes: rows are whole words
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
            reinterpret_cast<uint32_t*>(d)[i] = __ldg(reinterpret_cast<const uint32_t*>(s) + i);
    }
}
}  // namespace

cudaError_t launch_assemble(const AssembleParams& p, cudaStream_t stream) {
    if (p.height <= 0 || p.row_bytes == 0) return cudaSuccess;
    const unsigned per_row = (unsigned)((p.row_bytes / 16 + 255) / 256);
    dim3 grid(per_row < 1 ? 1 : (per_row > 8 ? 8 : per_row), (unsigned)p.height);
    assemble_bands_kernel<<<grid, 256, 0, stream>>>(p);
    return cudaGetLastError();
}

namespace {
__global__ void signal_only_kernel(unsigned int* signal) {
    __threadfence_system();
    atomicAdd_system(signal, 1u);
}
}  // namespace

cudaError_t launch_signal_only(unsigned int* d_signal, cudaStream_t stream) {
    signal_only_kernel<<<1, 1, 0, stream>>>(d_signal);
    return cudaGetLastError();
}

}  // namespace rtb
