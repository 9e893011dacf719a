// scene_pack_device.cu -- see scene_pack_device.cuh.  Replaces, on the per-frame path, the host
// loops the reference runs before every OpenCL launch (flattening + 6 uploads,
// RayTrace/states/MainState.cpp:646-655, :677-839) and this library's own former host packer.
#include "scene_pack_device.cuh"

#include "scene_pack_impl.h"

namespace rtb {

namespace {

__global__ void __launch_bounds__(64) boxes_from_vertices_kernel(const float* __restrict__ cube_v, int nc,
                                                                 Box* __restrict__ boxes, int* bad) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= nc) return;
    Box b;
    if (!pk::box_from_vertices(cube_v + (size_t)c * 144, &b)) atomicMin(bad, c);
    boxes[c] = b;
}

__global__ void __launch_bounds__(64) pack_frame_kernel(const PackFrameParams p) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n_bound) return;
    RtbBound bound;
    if (i < p.nc) {
        RtbSlab slabs[3];
        pk::pack_box(p.cf, p.boxes[i], slabs);
        for (int k = 0; k < 3; ++k)
            p.rec[3 * (size_t)i + k] = make_float4(slabs[k].alpha, slabs[k].beta, slabs[k].gamma, slabs[k].width);
        pk::pack_bound_box(p.fe, slabs, &bound);
    } else if (i < p.nc + p.ns) {
        const int s = i - p.nc;
        const float4 o = p.sph_o[s];
        const float centre[4] = {o.x, o.y, o.z, o.w};
        RtbSphere sp;
        pk::pack_sphere(p.cf, centre, p.sph_r[s], &sp);
        p.rec[3 * (size_t)p.nc + s] = make_float4(sp.uc, sp.vc, sp.rho, sp.tau0);
        pk::pack_bound_sphere(p.fe, sp, &bound);
    } else {
        bound = pk::bound_never();
    }
    float4* b = p.rec + p.bound_off + 2 * (size_t)i;
    b[0] = make_float4(bound.sA, bound.k1, bound.k0, bound.e);
    b[1] = make_float4(bound.cv, 0.f, 0.f, 0.f);
    if (i == 0)  // alignment gap between the exact and the bound records
        for (size_t g = 3 * (size_t)p.nc + (size_t)p.ns; g < p.bound_off; ++g) p.rec[g] = make_float4(0.f, 0.f, 0.f, 0.f);
}

constexpr int kRowsPerThread = 8;
__global__ void __launch_bounds__(128) row_table_kernel(const RowTableParams p) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n_bound) return;
    const float4 r0 = __ldg(p.bound_rec + 2 * (size_t)i);
    const float cv = __ldg(reinterpret_cast<const float*>(p.bound_rec + 2 * (size_t)i + 1));
    const RtbBound b = {r0.x, r0.y, r0.z, r0.w, cv, 0.f, 0.f, 0.f};
    const int l0 = blockIdx.y * kRowsPerThread;
#pragma unroll 1
    for (int l = l0; l < l0 + kRowsPerThread && l < p.local_rows; ++l) {
        const int lb = l / p.band_rows;
        const int gy = (lb * p.band_world + p.band_rank) * p.band_rows + (l - lb * p.band_rows);
        if (gy >= p.height) break;
        float e[2];
        pk::pack_row_entry(b, rtb_cam_v(p.cam, (float)gy), p.uabs, e);
        p.table[(size_t)gy * p.n_bound + i] = make_float2(e[0], e[1]);
    }
}

}  // namespace

cudaError_t launch_row_table(const RowTableParams& p, cudaStream_t stream) {
    if (p.n_bound <= 0 || p.local_rows <= 0) return cudaSuccess;
    dim3 grid((unsigned)((p.n_bound + 127) / 128), (unsigned)((p.local_rows + kRowsPerThread - 1) / kRowsPerThread));
    row_table_kernel<<<grid, 128, 0, stream>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_boxes_from_vertices(const float* d_cube_v, int nc, Box* d_boxes, int* d_bad, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(d_bad, 0x7F, sizeof(int), stream);  // kNoBadCube
    if (e != cudaSuccess || nc <= 0) return e;
    boxes_from_vertices_kernel<<<(nc + 63) / 64, 64, 0, stream>>>(d_cube_v, nc, d_boxes, d_bad);
    return cudaGetLastError();
}

cudaError_t launch_pack_frame(const PackFrameParams& p, cudaStream_t stream) {
    if (p.n_bound <= 0) return cudaSuccess;
    pack_frame_kernel<<<(p.n_bound + 63) / 64, 64, 0, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace rtb
