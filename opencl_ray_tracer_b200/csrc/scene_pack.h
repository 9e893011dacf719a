// scene_pack.h -- host-side (double precision) camera-space precompute for the ray-cast kernels.
// Turns the reference's flattened scene (rayTracer.cl:111-114 argument layout) plus a frame's
// camera into the 16-byte records of raycast_math.h.  O(primitives) per frame.
#ifndef RTB_SCENE_PACK_H_
#define RTB_SCENE_PACK_H_

#include <vector>

#include "raycast_math.h"

namespace rtb {

// Sign pattern (x,y,z in {-1,+1}) of the 36 vertices Cube::Cube pushes (RayTrace/Cube.cpp:10-45).
extern const signed char kCubeVertexSigns[36][3];

// A cube recovered from its 36 world-space vertices: centre + three half-axis vectors
// (any affine image of the unit cube, i.e. a parallelepiped).
struct Box {
    double c[3];
    double a[3][3];  // a[k] = half-axis k (image of +e_k)
    bool degenerate;  // zero volume: never hit (every triangle has det == 0 in the reference)
};

// Returns false if the vertices are not an affine image of the Cube.cpp vertex list.
bool box_from_vertices(const float* verts36x4, Box* out);

struct CameraFrame {
    double o0[4], du[4], dv[4], dir[4];
    // derived
    double l11, l12, l22;  // Cholesky factor of Q (upper)
    double tau_x, tau_y;   // dU.dir, dV.dir (4-D dots, glm::dot order is irrelevant in double)
    double quu, quv, qvv;
    RtbCam cam;            // float constants handed to the kernel
};

// Returns false for a degenerate camera (Q not positive definite).
bool camera_prepare(const float o0[4], const float du[4], const float dv[4], const float dir[4],
                    CameraFrame* out);

void pack_sphere(const CameraFrame& cf, const float centre[4], float radius, RtbSphere* out);
void pack_box(const CameraFrame& cf, const Box& box, RtbSlab out[3]);

// Extent of a frame's ray grid in camera coordinates (pixels x in [0,W-1], y in [0,H-1]).
struct FrameExtent {
    double umin, umax, vmin, vmax;  // bounding box of the (u,v) image of the pixel grid
    double uabs, vabs;              // max |u|, max |v| over the grid
    double pitch;                   // min(l11, l22): (u,v) length of one pixel step
};
FrameExtent frame_extent(const CameraFrame& cf, int width, int height);

// Conservative conic bounds (RtbBound, raycast_math.h) of the footprints that the float records
// `slabs` / `sph` produce in the kernel's fp32 arithmetic: every pixel of the frame on which
// rtb_cube_update / rtb_sphere_update can fire evaluates to q < 2 in fp32.
void pack_bound_box(const FrameExtent& fe, const RtbSlab slabs[3], RtbBound* out);
void pack_bound_sphere(const FrameExtent& fe, const RtbSphere& sph, RtbBound* out);

// One entry of the dense kernel's row table (raycast_math.h RtbRowEntry) for bound `b` on the pixel row whose camera
// coordinate is the float v; uabs = FrameExtent::uabs.  out = (sAn, rWn).
void pack_row_entry(const RtbBound& b, float v, double uabs, float out[2]);

// World-space records for the explicit-origin kernel: out[k] = (nx, ny, nz, alpha) with
// near_k(O) = alpha + n.O for k = 0..2, out[3] = (w0, w1, w2, 0).  dir3 = rayDir.xyz.
void pack_box_world(const double dir3[3], const Box& box, float out[16]);

}  // namespace rtb

#endif
