// rtb_internal.h -- the context behind the C ABI (include/rtb.h), shared by rtb_api.cu (single-device entry
// points) and rtb_group.cu (multi-GPU frames: several devices in one process, or one process per GPU).
#ifndef RTB_INTERNAL_H_
#define RTB_INTERNAL_H_

#include <cuda.h>  // types of the driver entry points fetched at run time; libcuda is NOT linked
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "../../include/rtb.h"
#include "raycast_kernels.cuh"
#include "scene_pack.h"
#include "scene_pack_device.cuh"

struct EventPair { cudaEvent_t a, b; };

// One frame's device-side precompute: the exact + bound records (pack_frame_kernel) and, for the dense kernel,
// the row table (row_table_kernel), with what they were built for.
struct FrameSet {
    float4* rec = nullptr;         // [0,3nc) slabs, [3nc,3nc+ns) spheres, [bound_off, +2*n_bound) bounds
    size_t rec_cap = 0;            // float4
    float2* tab = nullptr;         // height x n_bound row entries
    size_t tab_cap = 0;            // float2
    bool rec_valid = false;        // rec holds cam16 at w x h of the current scene
    float cam16[16] = {0};
    int w = 0, h = 0;
    bool tab_valid = false;        // tab holds the rows of (band_rows, band_rank, band_world) for the same camera
    int band_rows = 0, band_rank = 0, band_world = 0;
    double uabs = 0.0;             // FrameExtent::uabs of the pack (row_table_kernel needs it)
    cudaEvent_t ready = nullptr;   // slots only: recorded on pack_stream behind the pack (+ table)
};

struct rtb_group;

struct rtb_ctx {
    int device = 0;
    int sm_count = 0;
    int clock_khz = 0;
    char name[128] = {0};
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;  // device->host copies of finished row chunks (rtb_render)
    cudaEvent_t ev_chunk[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    // In-launch progress (rtb_render): the grid kernel publishes finished row chunks in d_progress and
    // copy_stream waits on the flag words with cuStreamWaitValue32 (fetched through the runtime).
    unsigned int* d_progress = nullptr;  // 2 * kMaxProgressChunks words + the launch-done CTA counter
    unsigned int progress_epoch = 0;
    CUresult (*stream_wait_value32)(CUstream, CUdeviceptr, cuuint32_t, unsigned int) = nullptr;
    CUresult (*stream_write_value32)(CUstream, CUdeviceptr, cuuint32_t, unsigned int) = nullptr;
    CUresult (*stream_batch_memop)(CUstream, unsigned int, CUstreamBatchMemOpParams*, unsigned int) = nullptr;
    std::string last_error;

    // scene: resident on the device in the reference's layout + the boxes recovered from the cubes
    bool have_scene = false;
    int ns = 0, nc = 0;
    float* d_cube_v = nullptr;     // 144 floats per cube (rayTracer.cl cubeVertices)
    float4* d_sph_o = nullptr;
    float* d_sph_r = nullptr;
    rtb::Box* d_boxes = nullptr;
    // the five scene arrays are sub-ranges of ONE device block (d_scene), laid out like the pinned staging buffer: one
    // host-to-device copy per rtb_set_scene instead of five; the boxes have their own buffer
    unsigned char* d_scene = nullptr;
    size_t d_scene_cap = 0, d_boxes_cap = 0;
    int* d_bad = nullptr;          // lowest index of a cube that is not a box (boxes_from_vertices_kernel)
    int* h_bad = nullptr;          // pinned
    // rtb_set_scene stages the caller's arrays in pinned memory and returns; ev_scene = staging consumed + verdict
    // copied back.  scene_check_pending: the NOT_A_BOX verdict of the last rtb_set_scene has not been read yet.
    unsigned char* h_scene = nullptr;
    size_t h_scene_cap = 0;
    cudaEvent_t ev_scene = nullptr;
    bool scene_check_pending = false;
    int defer_scene_check = 0;     // RTB_OPT_DEFER_SCENE_CHECK
    // host mirror of boxes / spheres, downloaded on demand (rtb_render_rays packs on the host)
    bool host_scene_valid = false;
    std::vector<float> sph_o, sph_r;
    std::vector<rtb::Box> boxes;

    // device buffers
    float4* d_cube_col = nullptr;
    float4* d_sph_col = nullptr;
    FrameSet cur;                  // the frame set of the last camera rendered without a slot
    float4* h_rec = nullptr;       // pinned staging for host-packed records (rtb_render_rays only)
    size_t h_rec_cap = 0;
    void* d_out[3] = {nullptr, nullptr, nullptr};  // rgba8, id, i32 frame buffers, grow-only
    size_t d_out_cap[3] = {0, 0, 0};
    unsigned long long h2d_bytes = 0, d2h_bytes = 0;  // bytes this context copied over PCIe
    float4* d_origins = nullptr;   // rays path
    size_t d_origins_cap = 0;

    // Per-context device state (cur, d_bins, progress words, slots) is shared by every stream the caller passes:
    // when the stream changes, the new one waits for an event recorded on the old one (order_streams).
    cudaStream_t state_stream = nullptr;
    cudaEvent_t ev_state = nullptr;

    // timing / accounting
    cudaEvent_t ev_total0 = nullptr, ev_total1 = nullptr;
    cudaEvent_t ev_rec = nullptr;  // last upload out of the pinned record staging area
    std::vector<EventPair> ev_free, ev_pending;
    double kernel_ms_sum = 0.0;
    int kernel_ms_n = 0;
    unsigned long long launches = 0;
    int tile_cull = 3;             // RTB_OPT_TILE_CULL (3 = auto)
    void* d_bins = nullptr;        // binned mode: per-tile primitive bit words of the launch in flight, grow-only
    size_t d_bins_cap = 0;
    // rtb_render_sequence_device renders consecutive frames on alternating lanes (lane 0 = the context's stream and
    // d_bins; lane 1 = seq_stream and its own bin words), so that a frame's ragged end / second launch overlaps the
    // next frame's start.
    rtb::TailLane tail_lane = {nullptr, nullptr, nullptr};   // dense launches of more than one wave: see launch_grid
    cudaStream_t seq_stream = nullptr;
    cudaEvent_t ev_seq_fork = nullptr, ev_seq_join = nullptr;
    void* d_bins_seq = nullptr;
    size_t d_bins_seq_cap = 0;
    // Frame sets packed ahead of their frames (rtb_prepack_cameras, rtb_render_sequence_device) on pack_stream.
    std::vector<FrameSet> slots;   // buffers and events are created once and reused
    int slots_used = 0;
    cudaStream_t pack_stream = nullptr;
    cudaEvent_t ev_prepack = nullptr;

    // rtb_init(n_devices > 1): this context fronts a group of per-device contexts in this process (rtb_group.cu)
    rtb_group* multi = nullptr;
};

namespace rtb {

int fail(rtb_ctx* ctx, int status, const std::string& msg);

#define RTB_CUDA(ctx, call)                                                                   \
    do {                                                                                      \
        cudaError_t e__ = (call);                                                             \
        if (e__ != cudaSuccess) {                                                             \
            const int st__ = (e__ == cudaErrorMemoryAllocation) ? RTB_ERR_OUT_OF_MEMORY : RTB_ERR_CUDA; \
            return rtb::fail(ctx, st__, std::string(#call) + ": " + cudaGetErrorString(e__)); \
        }                                                                                     \
    } while (0)

// ---- internals of rtb_api.cu that rtb_group.cu builds on ----------------------------------------
int init_single(rtb_ctx** out, int device);          // one context on one device
int effective_cull(const rtb_ctx* ctx);              // RTB_OPT_TILE_CULL with "auto" resolved for the current scene
int order_streams(rtb_ctx* ctx, cudaStream_t stream);
int collect_scene_verdict(rtb_ctx* ctx);             // blocks for the last rtb_set_scene's NOT_A_BOX verdict
int resolve_pending(rtb_ctx* ctx);
void drain(rtb_ctx* ctx, cudaStream_t extra);        // error path: nothing of this context is in flight afterwards
int ensure_device(rtb_ctx* ctx, void** p, size_t* cap, size_t need_bytes);
int check_frame_args(rtb_ctx* ctx, int width, int height, const float* o0, const float* du, const float* dv,
                     const float* dir);
// Makes this frame's records (and, for the dense kernel, the row table of the given bands) available to `stream`.
int prepare_frame(rtb_ctx* ctx, const CameraFrame& cf, int width, int height, const float* cam16, int band_rows,
                  int band_rank, int band_world, cudaStream_t stream, const FrameSet** fs);
// Enqueues one launch of the ray-cast kernels (with its timing event pair) on `stream`.
int enqueue_grid(rtb_ctx* ctx, const RtbCam& cam, const FrameSet* fs, int width, int height, int band_rows,
                 int band_rank, int band_world, int local_rows, void* d_rgba8, void* d_id, void* d_i32,
                 cudaStream_t stream, int out_global_rows = 0, int progress_chunks = 0,
                 const int* progress_bound = nullptr, void* done_signal = nullptr, int bins_lane = 0);
int bands_host_render(rtb_ctx* ctx, int width, int height, const float* cam16, const CameraFrame& cf, int band_rows,
                      int band_rank, int band_world, void* const host[3], float* kernel_ms);
int enqueue_bands_frame(rtb_ctx* ctx, int width, int height, const float* cam16, const CameraFrame& cf, int band_rows,
                        int band_rank, int band_world, void* d_frame_rgba8, void* d_signal, cudaStream_t stream);
int prepack_cameras(rtb_ctx* ctx, int width, int height, int band_rows, int band_rank, int band_world, int n,
                    const float* cameras, cudaStream_t stream);
int plan_progress_chunks(int rows, size_t row_bytes, int unit, int bound[kMaxProgressChunks + 1]);
int set_scene_single(rtb_ctx* ctx, int numSpheres, const float* sphereOrigins, const float* sphereRadius,
                     const float* sphereColours, int numCubes, const float* cubeVertices, const float* cubeColours);

}  // namespace rtb

#endif
