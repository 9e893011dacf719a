// raycast_kernels.cuh -- launch interface between the C ABI (rtb_api.cu) and the sm_100a kernels.
#ifndef RTB_RAYCAST_KERNELS_CUH_
#define RTB_RAYCAST_KERNELS_CUH_

#include <cuda_runtime.h>
#include <stdint.h>

#include "raycast_math.h"

namespace rtb {

// Bound records (RtbBound, 32 B) per staged shared-memory chunk: 512 primitives = 16 KB per stage.
constexpr int kChunkPrims = 512;
// Row-table kernels: primitives per staged chunk (4 KB of bound records -> 16 B x rows x 128 table).
constexpr int kTabPrims = 128;
// Row chunks a launch can report as finished while it is still running (rtb_render's pipelined copies).
constexpr int kMaxProgressChunks = 32;
// Primitives per warp vote in the dense loop; the host pads the bound list (and the row table) to a multiple of it.
constexpr int kBoundGroup = 8;

struct GridParams {
    int width, height;                      // full frame
    int band_rows, band_rank, band_world;   // interleaved row bands (band_world == 1: whole frame)
    int local_rows;                         // rows this launch renders = rows of the packed outputs
    int out_global_rows;                    // 1: outputs are whole frames (possibly peer memory), row = frame row
    int n_cubes, n_spheres;
    const float4* cube_rec;                 // 3 RtbSlab per cube
    const float4* sph_rec;                  // 1 RtbSphere per sphere
    const float4* bound_rec;                // 2 float4 (RtbBound) per primitive: cubes, spheres, padding
    int n_bound;                            // n_cubes + n_spheres rounded up to kBoundGroup
    // Row table of the dense kernel (row_table_kernel, scene_pack_device.cu): entry (frame row y, primitive i) at
    // row_tab[y * n_pad + i] = (sAn, rWn), height rows of n_pad = n_bound entries; rows of other ranks' bands
    // are not filled.  Unused by the culled / binned kernels.
    const float2* row_tab;
    int n_pad;
    const float4* cube_col;                 // (r,g,b,255)
    const float4* sph_col;
    uint32_t* out_rgba8;                    // nullable, local_rows*width
    int32_t* out_id;                        // nullable
    int4* out_i32;                          // nullable
    int vec_ok;                             // outputs 16-byte aligned and width % 4 == 0
    // Progress words (nullable): [0, kMaxProgressChunks) CTA counters, [kMaxProgressChunks, 2x) flags.
    // Chunk c = local rows [progress_bound[c], progress_bound[c+1]); the bounds are multiples of 16
    // (except the last, = local_rows), ascending, progress_chunks of them in use.
    unsigned int* progress;
    int progress_chunks;
    int progress_bound[kMaxProgressChunks + 1];
    unsigned int progress_epoch;
    // Launch-completion signal (nullable): done_counter = this context's CTA counter, done_signal = the
    // word the last CTA increments with a system-scope atomic (own or peer memory).
    unsigned int* done_counter;
    unsigned int* done_signal;
    int fence_system;                       // 1: system-scope fences also for the progress flags (debug knob)
    int tile_cull;                          // 1: raycast_cull_kernel (primitives culled per CTA tile; same pixels)
                                            // 2: bin_tiles_kernel + raycast_binned_kernel (the cull made once per frame)
    uint32_t* bins;                         // tile_cull == 2: [bin_words][tiles] words, grid_bin_words() of them
    int bin_words;                          // ceil(n_bound / 32)
    int tile_y0;                            // dense kernel: this launch's first tile row (launch_grid splits a frame's rows
                                            // into a body and a tail of shorter-lived CTAs); 0 otherwise
    RtbCam cam;
#ifdef RTB_EXPERIMENTS
    // CTA timeline of the dense kernel (tools/cta_trace.py): 4 words per CTA = globaltimer at CTA start, after the wait
    // for the first table chunk, at CTA end, and the SM id.  nullptr = off.  Experiments build only.
    unsigned long long* trace;
#endif
};
#ifdef RTB_EXPERIMENTS
unsigned long long* cta_trace_buffer(size_t ctas);  // RTB_CTA_TRACE set: a device buffer of 4 words per CTA, else nullptr
int cta_trace_dump(size_t ctas);                    // after the launch: synchronises and appends the words to $RTB_CTA_TRACE
#endif

// Explicit-origin ("rays") form: world-space records, one thread per ray.
struct RaysParams {
    size_t n_rays;
    const float4* origins;                  // n_rays
    float dir[4];
    int n_cubes, n_spheres;
    const float4* cube_rec;                 // 4 float4 per cube: 3 x (nx,ny,nz,alpha) + (w0,w1,w2,0)
    const float4* sph_rec;                  // 2 float4 per sphere: centre(4) + (r2, 0,0,0)
    const float4* cube_col;
    const float4* sph_col;
    uint32_t* out_rgba8;
    int32_t* out_id;
    int4* out_i32;
};

// De-interleave of gathered row bands (multi-GPU): `src` holds rank 0's packed rows, then rank 1's,
// ...; row_off[r] = first packed row of rank r.  dst = the final frame.  row_bytes = width * element size.
constexpr int kMaxBandWorld = 64;
struct AssembleParams {
    int height, band_rows, band_world;
    size_t row_bytes;
    const uint8_t* src;
    uint8_t* dst;
    int row_off[kMaxBandWorld];
};
cudaError_t launch_assemble(const AssembleParams& p, cudaStream_t stream);

// Tile geometry of the grid kernel variant chosen for a frame.
struct GridLaunchShape { int px, wx, wy, split; dim3 grid, block; size_t smem; };

GridLaunchShape grid_launch_shape(int width, int local_rows, int sm_count);
// 32-bit words of GridParams::bins a launch of this size needs (tile_cull == 2).
size_t grid_bin_words(int width, int local_rows, int n_bound);
// Pixel rows of a tile of the binned render (a multiple of 16); progress chunks must not cut a tile.
int binned_tile_rows();
// One frame's launch(es).  `tail` (nullable): a second stream plus two events of the caller's, which let a dense launch
// of more than one wave put its last tile rows into CTAs of half the lifetime on that stream (see launch_grid).
struct TailLane { cudaStream_t stream; cudaEvent_t fork, join; };
cudaError_t launch_grid(const GridParams& p, int sm_count, cudaStream_t stream, const TailLane* tail = nullptr);
cudaError_t launch_rays(const RaysParams& p, cudaStream_t stream);
cudaError_t launch_signal_only(unsigned int* d_signal, cudaStream_t stream);  // a rank with no rows still signals

}  // namespace rtb

#endif
