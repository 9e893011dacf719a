// scene_pack_device.cuh -- device-side camera-space precompute: the per-frame records the grid
// kernel reads are produced ON the GPU (one thread per primitive, double precision, the shared
// code of scene_pack_impl.h), so a frame needs no host pass over the scene and no record upload.
#ifndef RTB_SCENE_PACK_DEVICE_CUH_
#define RTB_SCENE_PACK_DEVICE_CUH_

#include <cuda_runtime.h>

#include "scene_pack.h"

namespace rtb {

// Layout of one frame's records in float4 units (shared with rtb_api.cu):
//   [0, 3nc)              RtbSlab x3 per cube      (reference-exact path, read through L2)
//   [3nc, 3nc+ns)         RtbSphere per sphere
//   [bound_off, +2*n_bound)  RtbBound per primitive (cubes, spheres, "never" padding up to a
//                         multiple of kBoundGroup): the list the grid kernel stages with TMA
struct PackFrameParams {
    CameraFrame cf;
    FrameExtent fe;
    int nc, ns, n_bound;
    const Box* boxes;        // nc, from boxes_from_vertices_kernel
    const float4* sph_o;     // ns sphere centres (x,y,z,w)
    const float* sph_r;      // ns radii
    float4* rec;             // destination records
    unsigned long long bound_off;  // float4 index of the first bound record (128-byte aligned)
};

// Recovers every cube's box from its 36 vertices (Cube::getTriangles() order).  *bad receives the
// lowest index of a cube that is not an affine image of the Cube.cpp vertex list (kNoBadCube = none).
constexpr int kNoBadCube = 0x7F7F7F7F;
cudaError_t launch_boxes_from_vertices(const float* d_cube_v, int nc, Box* d_boxes, int* d_bad, cudaStream_t stream);
cudaError_t launch_pack_frame(const PackFrameParams& p, cudaStream_t stream);

// Row table of the dense kernel: entry (frame row y, primitive i) = pk::pack_row_entry(bound i, v(y)) for the rows
// of `band_rank`'s bands (band_world == 1: every row), 8 bytes each, table[y * n_bound + i].  One thread per
// primitive and 8 rows, double precision.  Reads the bound records pack_frame_kernel has just written.
struct RowTableParams {
    RtbCam cam;
    double uabs;                 // FrameExtent::uabs: max |u| over the frame's rays
    int n_bound;                 // entries per row (bound records incl. the "never" padding)
    int height;
    int band_rows, band_rank, band_world, local_rows;
    const float4* bound_rec;     // 2 float4 per primitive
    float2* table;
};
cudaError_t launch_row_table(const RowTableParams& p, cudaStream_t stream);

}  // namespace rtb

#endif
