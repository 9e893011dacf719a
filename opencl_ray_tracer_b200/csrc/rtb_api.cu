// rtb_api.cu -- implementation of the C ABI declared in include/rtb.h (single-device entry points;
// multi-GPU frames: rtb_group.cu).
//
// Host-side replacement for MainState::openCLInit / executeRayTracerOpenCL
// (RayTrace/states/MainState.cpp:1181-1326, :641-934): device pick, resident scene buffers,
// per-frame camera-space pack ON the device (scene_pack_device.cu), kernel launch
// (raycast_kernels.cu), read-back.
// There is NO CPU path in this library.
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "rtb_internal.h"
#include "rtb_group.h"

using namespace rtb;

namespace rtb {

int fail(rtb_ctx* ctx, int status, const std::string& msg) {
    if (ctx) ctx->last_error = msg;
    return status;
}

int ensure_device(rtb_ctx* ctx, void** p, size_t* cap, size_t need_bytes) {
    if (*cap >= need_bytes) return RTB_OK;
    if (*p) RTB_CUDA(ctx, cudaFree(*p));
    *p = nullptr;
    *cap = 0;
    RTB_CUDA(ctx, cudaMalloc(p, need_bytes));
    *cap = need_bytes;
    return RTB_OK;
}

int resolve_pending(rtb_ctx* ctx) {
    for (const EventPair& e : ctx->ev_pending) {
        RTB_CUDA(ctx, cudaEventSynchronize(e.b));
        float ms = 0.f;
        RTB_CUDA(ctx, cudaEventElapsedTime(&ms, e.a, e.b));
        ctx->kernel_ms_sum += ms;
        ctx->kernel_ms_n += 1;
        ctx->ev_free.push_back(e);
    }
    ctx->ev_pending.clear();
    return RTB_OK;
}

// Error path of the blocking entry points: whatever was enqueued before the failure may still be writing into
// the caller's buffers, so nothing of this context may be in flight when the error is returned.
void drain(rtb_ctx* ctx, cudaStream_t extra) {
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
    if (ctx->seq_stream) cudaStreamSynchronize(ctx->seq_stream);
    if (ctx->tail_lane.stream) cudaStreamSynchronize(ctx->tail_lane.stream);
    if (extra) cudaStreamSynchronize(extra);
    for (const EventPair& e : ctx->ev_pending) ctx->ev_free.push_back(e);  // timings of a failed call are dropped
    ctx->ev_pending.clear();
    (void)cudaGetLastError();
}

// The per-context device state (frame sets, bins, progress / done words) has one writer at a time: when a call
// arrives on another stream than the previous one, the new stream first waits for everything the old one was given.
int order_streams(rtb_ctx* ctx, cudaStream_t stream) {
    if (ctx->state_stream == stream) return RTB_OK;
    if (ctx->state_stream != nullptr) {
        if (cudaEventRecord(ctx->ev_state, ctx->state_stream) == cudaSuccess) {
            RTB_CUDA(ctx, cudaStreamWaitEvent(stream, ctx->ev_state, 0));
        } else {  // the old stream is gone (destroyed by the caller): its work still has to finish first
            (void)cudaGetLastError();
            RTB_CUDA(ctx, cudaDeviceSynchronize());
        }
    }
    ctx->state_stream = stream;
    return RTB_OK;
}

int effective_cull(const rtb_ctx* ctx) {
    if (ctx->tile_cull != 3) return ctx->tile_cull;
    return (ctx->nc + ctx->ns >= 64) ? 2 : 0;  // auto: binned from 64 primitives on, else the dense kernel
}

}  // namespace rtb

namespace {

int ensure_frame_set(rtb_ctx* ctx, FrameSet* fs, size_t n_float4, size_t n_tab) {
    if (n_float4 == 0) n_float4 = 1;
    void* p = fs->rec;
    size_t cap = fs->rec_cap * sizeof(float4);
    int rc = ensure_device(ctx, &p, &cap, n_float4 * sizeof(float4));
    fs->rec = static_cast<float4*>(p);
    fs->rec_cap = cap / sizeof(float4);
    if (rc != RTB_OK) return rc;
    if (n_tab) {
        p = fs->tab;
        cap = fs->tab_cap * sizeof(float2);
        rc = ensure_device(ctx, &p, &cap, n_tab * sizeof(float2));
        fs->tab = static_cast<float2*>(p);
        fs->tab_cap = cap / sizeof(float2);
    }
    return rc;
}

int ensure_host_records(rtb_ctx* ctx, size_t n_float4) {
    if (ctx->h_rec_cap < n_float4) {
        if (ctx->h_rec) RTB_CUDA(ctx, cudaFreeHost(ctx->h_rec));
        ctx->h_rec = nullptr;
        RTB_CUDA(ctx, cudaMallocHost(&ctx->h_rec, n_float4 * sizeof(float4)));
        ctx->h_rec_cap = n_float4;
    }
    return RTB_OK;
}

// Event pairs for per-launch kernel timing on whatever stream the launch goes to.
int take_events(rtb_ctx* ctx, EventPair* out) {
    if (ctx->ev_free.empty()) {
        EventPair e;
        RTB_CUDA(ctx, cudaEventCreate(&e.a));
        RTB_CUDA(ctx, cudaEventCreate(&e.b));
        ctx->ev_free.push_back(e);
    }
    *out = ctx->ev_free.back();
    ctx->ev_free.pop_back();
    return RTB_OK;
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

void free_frame_set(FrameSet* fs) {
    cudaFree(fs->rec);
    cudaFree(fs->tab);
    if (fs->ready) cudaEventDestroy(fs->ready);
    *fs = FrameSet();
}

}  // namespace

extern "C" {

const char* rtb_version(void) { return "rtb 0.1 (sm_100a)"; }

const char* rtb_error_string(int status) {
    switch (status) {
        case RTB_OK: return "RTB_OK";
        case RTB_ERR_INVALID_ARG: return "RTB_ERR_INVALID_ARG";
        case RTB_ERR_NO_DEVICE: return "RTB_ERR_NO_DEVICE";
        case RTB_ERR_CUDA: return "RTB_ERR_CUDA";
        case RTB_ERR_NO_SCENE: return "RTB_ERR_NO_SCENE";
        case RTB_ERR_NOT_A_BOX: return "RTB_ERR_NOT_A_BOX";
        case RTB_ERR_CAMERA: return "RTB_ERR_CAMERA";
        case RTB_ERR_OUT_OF_MEMORY: return "RTB_ERR_OUT_OF_MEMORY";
        case RTB_ERR_UNSUPPORTED: return "RTB_ERR_UNSUPPORTED";
        default: return "RTB_ERR_UNKNOWN";
    }
}

const char* rtb_last_error_message(const rtb_ctx* ctx) { return ctx ? ctx->last_error.c_str() : ""; }

int rtb_init(rtb_ctx** out, int n_devices, const int* device_ids) {
    if (!out) return RTB_ERR_INVALID_ARG;
    *out = nullptr;
    if (n_devices < 1 || n_devices > RTB_MAX_GROUP) return RTB_ERR_INVALID_ARG;
    if (n_devices > 1) return group_init_local(out, n_devices, device_ids);  // rtb_group.cu
    return init_single(out, device_ids ? device_ids[0] : 0);
}

void rtb_shutdown(rtb_ctx* ctx) {
    if (!ctx) return;
    if (ctx->multi) group_shutdown_local(ctx);  // the other devices' contexts and the group itself
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
    if (ctx->pack_stream) cudaStreamSynchronize(ctx->pack_stream);
    for (const EventPair& e : ctx->ev_pending) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); }
    for (const EventPair& e : ctx->ev_free) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); }
    if (ctx->ev_total0) cudaEventDestroy(ctx->ev_total0);
    if (ctx->ev_total1) cudaEventDestroy(ctx->ev_total1);
    if (ctx->ev_rec) cudaEventDestroy(ctx->ev_rec);
    if (ctx->ev_state) cudaEventDestroy(ctx->ev_state);
    if (ctx->ev_scene) cudaEventDestroy(ctx->ev_scene);
    free_frame_set(&ctx->cur);
    for (auto& fs : ctx->slots) free_frame_set(&fs);
    cudaFree(ctx->d_scene);
    cudaFree(ctx->d_boxes);
    cudaFree(ctx->d_bad);
    cudaFree(ctx->d_progress);
    cudaFree(ctx->d_bins);
    cudaFree(ctx->d_bins_seq);
    if (ctx->tail_lane.stream) cudaStreamDestroy(ctx->tail_lane.stream);
    if (ctx->tail_lane.fork) cudaEventDestroy(ctx->tail_lane.fork);
    if (ctx->tail_lane.join) cudaEventDestroy(ctx->tail_lane.join);
    if (ctx->seq_stream) cudaStreamDestroy(ctx->seq_stream);
    if (ctx->ev_seq_fork) cudaEventDestroy(ctx->ev_seq_fork);
    if (ctx->ev_seq_join) cudaEventDestroy(ctx->ev_seq_join);
    if (ctx->pack_stream) cudaStreamDestroy(ctx->pack_stream);
    if (ctx->ev_prepack) cudaEventDestroy(ctx->ev_prepack);
    if (ctx->h_bad) cudaFreeHost(ctx->h_bad);
    if (ctx->h_rec) cudaFreeHost(ctx->h_rec);
    if (ctx->h_scene) cudaFreeHost(ctx->h_scene);
    for (int i = 0; i < 3; ++i) cudaFree(ctx->d_out[i]);
    cudaFree(ctx->d_origins);
    for (int i = 0; i < 8; ++i)
        if (ctx->ev_chunk[i]) cudaEventDestroy(ctx->ev_chunk[i]);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int rtb_set_scene(rtb_ctx* ctx, int numSpheres, const float* sphereOrigins, const float* sphereRadius,
                  const float* sphereColours, int numCubes, const float* cubeVertices,
                  const float* cubeColours) {
    if (!ctx) return RTB_ERR_INVALID_ARG;
    if (numSpheres < 0 || numCubes < 0) return fail(ctx, RTB_ERR_INVALID_ARG, "negative primitive count");
    if (numSpheres > 0 && (!sphereOrigins || !sphereRadius || !sphereColours))
        return fail(ctx, RTB_ERR_INVALID_ARG, "null sphere array");
    if (numCubes > 0 && (!cubeVertices || !cubeColours)) return fail(ctx, RTB_ERR_INVALID_ARG, "null cube array");
    if (ctx->multi)
        return group_set_scene_local(ctx, numSpheres, sphereOrigins, sphereRadius, sphereColours, numCubes, cubeVertices,
                                     cubeColours);
    return set_scene_single(ctx, numSpheres, sphereOrigins, sphereRadius, sphereColours, numCubes, cubeVertices, cubeColours);
}

}  // extern "C"

namespace rtb {

int init_single(rtb_ctx** out, int dev) {
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) return RTB_ERR_NO_DEVICE;
    if (dev < 0 || dev >= count) return RTB_ERR_NO_DEVICE;
    if (cudaSetDevice(dev) != cudaSuccess) return RTB_ERR_NO_DEVICE;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess) return RTB_ERR_CUDA;
    if (prop.major < 10) return RTB_ERR_NO_DEVICE;  // sm_100a binary only
    rtb_ctx* ctx = new (std::nothrow) rtb_ctx();
    if (!ctx) return RTB_ERR_OUT_OF_MEMORY;
    ctx->device = dev;
    ctx->sm_count = prop.multiProcessorCount;
    cudaDeviceGetAttribute(&ctx->clock_khz, cudaDevAttrClockRate, dev);
    std::snprintf(ctx->name, sizeof ctx->name, "%.127s", prop.name);
    bool chunk_ok = cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) == cudaSuccess;
    for (int i = 0; i < 8 && chunk_ok; ++i)
        chunk_ok = cudaEventCreateWithFlags(&ctx->ev_chunk[i], cudaEventDisableTiming) == cudaSuccess;
    if (!chunk_ok || cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreate(&ctx->ev_total0) != cudaSuccess || cudaEventCreate(&ctx->ev_total1) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_rec, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_state, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_scene, cudaEventDisableTiming) != cudaSuccess ||
        cudaMalloc(&ctx->d_bad, sizeof(int)) != cudaSuccess ||
        cudaMalloc(&ctx->d_progress, (2 * kMaxProgressChunks + 8) * sizeof(unsigned int)) != cudaSuccess ||
        cudaMemset(ctx->d_progress, 0, (2 * kMaxProgressChunks + 8) * sizeof(unsigned int)) != cudaSuccess ||
        cudaMallocHost(&ctx->h_bad, sizeof(int)) != cudaSuccess) {
        rtb_shutdown(ctx);
        return RTB_ERR_CUDA;
    }
    {   // optional: without them rtb_render falls back to one launch per row chunk, rtb_group_* to RTB_ERR_UNSUPPORTED
        const char* off = std::getenv("RTB_NO_STREAM_WAIT");
        if (!(off && off[0] == '1')) {
            void* fn = nullptr;
            cudaDriverEntryPointQueryResult qr;
            if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &fn, cudaEnableDefault, &qr) == cudaSuccess &&
                qr == cudaDriverEntryPointSuccess && fn)
                ctx->stream_wait_value32 = reinterpret_cast<CUresult (*)(CUstream, CUdeviceptr, cuuint32_t, unsigned int)>(fn);
            else
                (void)cudaGetLastError();
            fn = nullptr;
            if (cudaGetDriverEntryPoint("cuStreamWriteValue32", &fn, cudaEnableDefault, &qr) == cudaSuccess &&
                qr == cudaDriverEntryPointSuccess && fn)
                ctx->stream_write_value32 = reinterpret_cast<CUresult (*)(CUstream, CUdeviceptr, cuuint32_t, unsigned int)>(fn);
            else
                (void)cudaGetLastError();
            fn = nullptr;
            if (cudaGetDriverEntryPoint("cuStreamBatchMemOp", &fn, cudaEnableDefault, &qr) == cudaSuccess &&
                qr == cudaDriverEntryPointSuccess && fn)
                ctx->stream_batch_memop =
                    reinterpret_cast<CUresult (*)(CUstream, unsigned int, CUstreamBatchMemOpParams*, unsigned int)>(fn);
            else
                (void)cudaGetLastError();
        }
    }
    *out = ctx;
    return RTB_OK;
}

// Blocks until the last rtb_set_scene's staging area has been consumed and its NOT_A_BOX verdict is on the host.
int collect_scene_verdict(rtb_ctx* ctx) {
    if (!ctx->scene_check_pending) return RTB_OK;
    RTB_CUDA(ctx, cudaEventSynchronize(ctx->ev_scene));
    ctx->scene_check_pending = false;
    if (ctx->nc && *ctx->h_bad != kNoBadCube) {
        char buf[160];
        std::snprintf(buf, sizeof buf, "cube %d: its 36 vertices are not an affine image of the Cube.cpp vertex list",
                      *ctx->h_bad);
        ctx->have_scene = false;
        return fail(ctx, RTB_ERR_NOT_A_BOX, buf);
    }
    return RTB_OK;
}

int set_scene_single(rtb_ctx* ctx, int numSpheres, const float* sphereOrigins, const float* sphereRadius,
                     const float* sphereColours, int numCubes, const float* cubeVertices, const float* cubeColours) {
    RTB_CUDA(ctx, cudaSetDevice(ctx->device));
    ctx->have_scene = false;
    ctx->cur.rec_valid = ctx->cur.tab_valid = false;
    ctx->slots_used = 0;
    ctx->host_scene_valid = false;
    cudaStream_t st = ctx->stream;
    int rc = order_streams(ctx, st);  // frames of the previous scene may still be reading the buffers
    if (rc != RTB_OK) return rc;
    const size_t nc = (size_t)numCubes, ns = (size_t)numSpheres;
    // The caller may free its arrays on return: they are staged in the context's pinned buffer (one host memcpy each)
    // and go to the device from there in ONE copy, into a device block with the same layout -- the 5 arrays of
    // MainState.cpp:677-839 as the reference lays them out, each 256-byte aligned; the boxes are recovered from the
    // vertices on the device.  Nothing here waits for the GPU unless the staging area is still being read by the
    // previous rtb_set_scene.
    const size_t bytes[5] = {144 * sizeof(float) * nc, sizeof(float4) * nc, sizeof(float4) * ns, sizeof(float) * ns,
                             sizeof(float4) * ns};
    const void* src[5] = {cubeVertices, cubeColours, sphereOrigins, sphereRadius, sphereColours};
    size_t total = 0, off[5];
    for (int k = 0; k < 5; ++k) { off[k] = total; total += (bytes[k] + 255) & ~(size_t)255; }
    if (total == 0) total = 256;
    {
        const size_t need_boxes = sizeof(Box) * (nc ? nc : 1);
        if (ctx->d_scene_cap < total || ctx->d_boxes_cap < need_boxes)
            RTB_CUDA(ctx, cudaDeviceSynchronize());  // cudaFree below: nothing may still read the old buffers
        void* blk = ctx->d_scene;
        int rc0 = ensure_device(ctx, &blk, &ctx->d_scene_cap, total);
        ctx->d_scene = static_cast<unsigned char*>(blk);
        void* bx = ctx->d_boxes;
        if (rc0 == RTB_OK) rc0 = ensure_device(ctx, &bx, &ctx->d_boxes_cap, need_boxes);
        ctx->d_boxes = static_cast<Box*>(bx);
        if (rc0 != RTB_OK) return rc0;
        ctx->d_cube_v = reinterpret_cast<float*>(ctx->d_scene + off[0]);
        ctx->d_cube_col = reinterpret_cast<float4*>(ctx->d_scene + off[1]);
        ctx->d_sph_o = reinterpret_cast<float4*>(ctx->d_scene + off[2]);
        ctx->d_sph_r = reinterpret_cast<float*>(ctx->d_scene + off[3]);
        ctx->d_sph_col = reinterpret_cast<float4*>(ctx->d_scene + off[4]);
    }
    if (ctx->scene_check_pending) RTB_CUDA(ctx, cudaEventSynchronize(ctx->ev_scene));  // staging area and h_bad free again
    ctx->scene_check_pending = false;
    if (ctx->h_scene_cap < total) {
        if (ctx->h_scene) RTB_CUDA(ctx, cudaFreeHost(ctx->h_scene));
        ctx->h_scene = nullptr;
        ctx->h_scene_cap = 0;
        RTB_CUDA(ctx, cudaMallocHost(&ctx->h_scene, total));
        ctx->h_scene_cap = total;
    }
    size_t used = 0;
    for (int k = 0; k < 5; ++k)
        if (bytes[k]) {
            std::memcpy(ctx->h_scene + off[k], src[k], bytes[k]);
            used = off[k] + bytes[k];
        }
    if (used) RTB_CUDA(ctx, cudaMemcpyAsync(ctx->d_scene, ctx->h_scene, used, cudaMemcpyHostToDevice, st));
    RTB_CUDA(ctx, launch_boxes_from_vertices(ctx->d_cube_v, numCubes, ctx->d_boxes, ctx->d_bad, st));
    if (nc) ctx->launches += 1;
    RTB_CUDA(ctx, cudaMemcpyAsync(ctx->h_bad, ctx->d_bad, sizeof(int), cudaMemcpyDeviceToHost, st));
    RTB_CUDA(ctx, cudaEventRecord(ctx->ev_scene, st));
    ctx->scene_check_pending = true;
    ctx->h2d_bytes += (144 * sizeof(float) + sizeof(float4)) * nc + (2 * sizeof(float4) + sizeof(float)) * ns;
    ctx->d2h_bytes += sizeof(int);
    ctx->ns = numSpheres;
    ctx->nc = numCubes;
    ctx->have_scene = true;
    // RTB_OPT_DEFER_SCENE_CHECK = 0 (default): the NOT_A_BOX verdict is this call's return value (one wait for the
    // device); 1: the call returns at once and the next render call reports it.
    if (!ctx->defer_scene_check) return collect_scene_verdict(ctx);
    return RTB_OK;
}

}  // namespace rtb

namespace rtb {
namespace {

// One frame's records (FrameSet::rec, float4 units):
//   [0, 3nc)            RtbSlab x3 per cube      (reference-exact path, read through L2)
//   [3nc, 3nc+ns)       RtbSphere per sphere
//   [bound_offset, +2*n_bound)  RtbBound per primitive (cubes, spheres, "never" padding up to a
//                       multiple of kBoundGroup): what the culled / binned kernels and row_table_kernel read
size_t exact_records(const rtb_ctx* ctx) { return 3 * (size_t)ctx->nc + (size_t)ctx->ns; }
size_t bound_offset(const rtb_ctx* ctx) { return (exact_records(ctx) + 7) & ~(size_t)7; }  // 128-byte aligned
int bound_count(const rtb_ctx* ctx) {
    return (int)(((size_t)ctx->nc + (size_t)ctx->ns + kBoundGroup - 1) / kBoundGroup * kBoundGroup);
}
size_t frame_records(const rtb_ctx* ctx) { return bound_offset(ctx) + 2 * (size_t)bound_count(ctx); }
size_t table_entries(const rtb_ctx* ctx, int height) { return (size_t)height * (size_t)bound_count(ctx); }

// Enqueues the device pack of one frame's records into `fs` (scene_pack_device.cu).
int enqueue_pack(rtb_ctx* ctx, const CameraFrame& cf, int width, int height, const float* cam16, FrameSet* fs,
                 cudaStream_t stream) {
    fs->rec_valid = fs->tab_valid = false;
    int rc = ensure_frame_set(ctx, fs, frame_records(ctx) + 32, 0);
    if (rc != RTB_OK) return rc;
    PackFrameParams pp;
    pp.cf = cf;
    pp.fe = frame_extent(cf, width, height);
    pp.nc = ctx->nc; pp.ns = ctx->ns; pp.n_bound = bound_count(ctx);
    pp.boxes = ctx->d_boxes; pp.sph_o = ctx->d_sph_o; pp.sph_r = ctx->d_sph_r;
    pp.rec = fs->rec;
    pp.bound_off = bound_offset(ctx);
    RTB_CUDA(ctx, launch_pack_frame(pp, stream));
    ctx->launches += 1;
    std::memcpy(fs->cam16, cam16, sizeof fs->cam16);
    fs->w = width; fs->h = height;
    fs->uabs = pp.fe.uabs;
    fs->rec_valid = true;
    return RTB_OK;
}

// Enqueues the row table of the dense kernel for the rows of the given bands, behind the pack on `stream`.
int enqueue_table(rtb_ctx* ctx, const CameraFrame& cf, FrameSet* fs, int band_rows, int band_rank, int band_world,
                  cudaStream_t stream) {
    fs->tab_valid = false;
    const int rc = ensure_frame_set(ctx, fs, frame_records(ctx) + 32, table_entries(ctx, fs->h));
    if (rc != RTB_OK) return rc;
    RowTableParams tp;
    tp.cam = cf.cam;
    tp.uabs = fs->uabs;
    tp.n_bound = bound_count(ctx);
    tp.height = fs->h;
    tp.band_rows = band_rows; tp.band_rank = band_rank; tp.band_world = band_world;
    tp.local_rows = rtb_band_local_rows(fs->h, band_rows, band_rank, band_world);
    tp.bound_rec = fs->rec + bound_offset(ctx);
    tp.table = fs->tab;
    RTB_CUDA(ctx, launch_row_table(tp, stream));
    if (tp.local_rows > 0 && tp.n_bound > 0) ctx->launches += 1;
    fs->band_rows = band_rows; fs->band_rank = band_rank; fs->band_world = band_world;
    fs->tab_valid = true;
    return RTB_OK;
}

bool same_bands(const FrameSet* fs, int band_rows, int band_rank, int band_world) {
    if (fs->band_world == 1 && fs->tab_valid) return true;  // every row is filled: good for any band set
    return fs->band_rows == band_rows && fs->band_rank == band_rank && fs->band_world == band_world;
}

}  // namespace

// Makes this frame's records available to `stream` and returns where they are: a slot packed ahead of time for
// exactly this camera and frame size (rtb_prepack_cameras; the stream then waits for that pack), ctx->cur if it
// already holds them (same 16 floats, same frame size, same scene), else ctx->cur after a pack on `stream`.  With
// the dense kernel selected the set's row table is made to hold the rows of the given bands as well.
int prepare_frame(rtb_ctx* ctx, const CameraFrame& cf, int width, int height, const float* cam16, int band_rows,
                  int band_rank, int band_world, cudaStream_t stream, const FrameSet** out) {
    const bool need_tab = effective_cull(ctx) == 0;
    FrameSet* fs = nullptr;
    for (int s = 0; s < ctx->slots_used && !fs; ++s) {
        FrameSet* c = &ctx->slots[(size_t)s];
        if (c->rec_valid && c->w == width && c->h == height && std::memcmp(cam16, c->cam16, sizeof c->cam16) == 0) {
            RTB_CUDA(ctx, cudaStreamWaitEvent(stream, c->ready, 0));
            fs = c;
        }
    }
    if (!fs) {
        fs = &ctx->cur;
        if (!(fs->rec_valid && fs->w == width && fs->h == height && std::memcmp(cam16, fs->cam16, sizeof fs->cam16) == 0)) {
            const int rc = enqueue_pack(ctx, cf, width, height, cam16, fs, stream);
            if (rc != RTB_OK) return rc;
        }
    }
    if (need_tab && !(fs->tab_valid && same_bands(fs, band_rows, band_rank, band_world))) {
        const int rc = enqueue_table(ctx, cf, fs, band_rows, band_rank, band_world, stream);
        if (rc != RTB_OK) return rc;
    }
    *out = fs;
    return RTB_OK;
}

// Packs the frame sets of n cameras into slots on pack_stream, behind everything already enqueued on `stream`
// (earlier frames may still be reading the slots).  Always packs: nothing is kept from an earlier call.
namespace { constexpr int kMaxCamSlots = 256; }
int prepack_cameras(rtb_ctx* ctx, int width, int height, int band_rows, int band_rank, int band_world, int n,
                    const float* cameras, cudaStream_t stream) {
    ctx->slots_used = 0;
    if (n <= 0) return RTB_OK;
    if (n > kMaxCamSlots) n = kMaxCamSlots;
    const bool with_tab = effective_cull(ctx) == 0;
    const size_t per_slot = frame_records(ctx) * sizeof(float4) + (with_tab ? table_entries(ctx, height) * sizeof(float2) : 0);
    while (n > 1 && per_slot * (size_t)n > ((size_t)4 << 30)) n /= 2;  // at most 4 GB of frame sets
    if (!ctx->pack_stream) {
        int lo = 0, hi = 0;
        RTB_CUDA(ctx, cudaDeviceGetStreamPriorityRange(&lo, &hi));
        RTB_CUDA(ctx, cudaStreamCreateWithPriority(&ctx->pack_stream, cudaStreamNonBlocking, hi));
        RTB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_prepack, cudaEventDisableTiming));
    }
    while ((int)ctx->slots.size() < n) {
        FrameSet fs;
        RTB_CUDA(ctx, cudaEventCreateWithFlags(&fs.ready, cudaEventDisableTiming));
        ctx->slots.push_back(fs);
    }
    bool grow = false;
    for (int s = 0; s < n; ++s)
        grow = grow || ctx->slots[(size_t)s].rec_cap < frame_records(ctx) + 32 ||
               (with_tab && ctx->slots[(size_t)s].tab_cap < table_entries(ctx, height));
    if (grow) RTB_CUDA(ctx, cudaDeviceSynchronize());  // buffers are re-allocated below
    RTB_CUDA(ctx, cudaEventRecord(ctx->ev_prepack, stream));
    RTB_CUDA(ctx, cudaStreamWaitEvent(ctx->pack_stream, ctx->ev_prepack, 0));
    for (int s = 0; s < n; ++s) {
        const float* c = cameras + 16 * (size_t)s;
        CameraFrame cf;
        if (!camera_prepare(c, c + 4, c + 8, c + 12, &cf)) {
            char buf[96];
            std::snprintf(buf, sizeof buf, "degenerate camera %d", s);
            return fail(ctx, RTB_ERR_CAMERA, buf);
        }
        FrameSet* fs = &ctx->slots[(size_t)s];
        int rc = enqueue_pack(ctx, cf, width, height, c, fs, ctx->pack_stream);
        if (rc == RTB_OK && with_tab) rc = enqueue_table(ctx, cf, fs, band_rows, band_rank, band_world, ctx->pack_stream);
        if (rc != RTB_OK) return rc;
        RTB_CUDA(ctx, cudaEventRecord(fs->ready, ctx->pack_stream));
        ctx->slots_used = s + 1;
    }
    return RTB_OK;
}

namespace {

// Host mirror of the device scene (boxes, sphere centres / radii) for the host-packed rays path.
int ensure_host_scene(rtb_ctx* ctx) {
    if (ctx->host_scene_valid) return RTB_OK;
    ctx->boxes.resize((size_t)ctx->nc);
    ctx->sph_o.resize(4 * (size_t)ctx->ns);
    ctx->sph_r.resize((size_t)ctx->ns);
    RTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // the scene upload / box recovery of rtb_set_scene
    if (ctx->nc)
        RTB_CUDA(ctx, cudaMemcpy(ctx->boxes.data(), ctx->d_boxes, sizeof(Box) * (size_t)ctx->nc, cudaMemcpyDeviceToHost));
    if (ctx->ns) {
        RTB_CUDA(ctx, cudaMemcpy(ctx->sph_o.data(), ctx->d_sph_o, sizeof(float4) * (size_t)ctx->ns, cudaMemcpyDeviceToHost));
        RTB_CUDA(ctx, cudaMemcpy(ctx->sph_r.data(), ctx->d_sph_r, sizeof(float) * (size_t)ctx->ns, cudaMemcpyDeviceToHost));
    }
    ctx->d2h_bytes += sizeof(Box) * (size_t)ctx->nc + (sizeof(float4) + sizeof(float)) * (size_t)ctx->ns;
    ctx->host_scene_valid = true;
    return RTB_OK;
}

}  // namespace

// Row chunks in which a launch reports progress (rtb_render, rtb_render_bands_host): 2-4 MB of
// output each, bounds on multiples of `unit` rows (a multiple of 16); returns the number of chunks.
// The copy stream is nearly as busy as the kernel (C3: 0.6 ms of PCIe time under a 0.5 ms launch), so
// the copies have to start early and every stream wait (1-2 us of the copy stream's time) counts:
// measured on C3, device span of rtb_render: 16 x 2 MB 0.83 ms; 64 x 0.5 MB with copies grouped by four
// 0.89 ms; sizes shrinking towards the end (8, 8, 8, 4, 2, 1, 0.5, 0.5 MB) 0.93 ms; 16 x 2 MB with the last
// two cut in four 0.87 ms.
int plan_progress_chunks(int rows, size_t row_bytes, int unit, int bound[kMaxProgressChunks + 1]) {
    // about an eighth of the output per chunk, between 2 and 4 MB (RTB_CHUNK_KB overrides: tools/e2e_breakdown.py).
    // Re-measured with the 0.60 ms kernel of v8, whole C3 frame, e2e: 2 MB 0.930 ms, 3 MB 0.913, 4 MB 0.888, 6 MB 0.938,
    // 8 MB 0.943 (profiles/r01_e2e_chunks_v8.txt): the copy stream spends ~10 us per chunk on the wait + copy launch.
    static const size_t chunk_override = [] {
        const char* e = std::getenv("RTB_CHUNK_KB");
        const long kb = e ? std::atol(e) : 0;
        return kb > 0 ? (size_t)kb << 10 : (size_t)0;
    }();
    size_t chunk_bytes = row_bytes * (size_t)rows / 8;
    if (chunk_bytes < ((size_t)2 << 20)) chunk_bytes = (size_t)2 << 20;
    if (chunk_bytes > ((size_t)4 << 20)) chunk_bytes = (size_t)4 << 20;
    if (chunk_override) chunk_bytes = chunk_override;
    size_t want = (row_bytes * (size_t)rows + chunk_bytes - 1) / chunk_bytes;
    if (want < 1) want = 1;
    if (want > (size_t)kMaxProgressChunks) want = kMaxProgressChunks;
    int chunk_rows = (int)(((size_t)rows + want - 1) / want);
    chunk_rows = (chunk_rows + unit - 1) / unit * unit;
    int n = (rows + chunk_rows - 1) / chunk_rows;
    for (int i = 0; i < n; ++i) bound[i] = i * chunk_rows;
    bound[n] = rows;
    // The copy stream is the longer of the two pipelines (C3: 0.60 ms of PCIe time + ~10 us per chunk against a
    // 0.5 ms launch), so it should start early: the first chunk is cut at a quarter (RTB_CHUNK_RAMP=0: uniform).
    static const bool ramp = [] { const char* e = std::getenv("RTB_CHUNK_RAMP"); return !(e && e[0] == '0'); }();
    const int first = chunk_rows / 4 / unit * unit;
    if (ramp && n >= 4 && n < kMaxProgressChunks && first >= unit) {
        for (int i = n; i >= 1; --i) bound[i + 1] = bound[i];
        bound[1] = first;
        n += 1;
    }
    return n;
}

// Enqueues one launch of the ray-cast kernels (with its timing event pair) on `stream`.
int enqueue_grid(rtb_ctx* ctx, const RtbCam& cam, const FrameSet* fs, int width, int height, int band_rows,
                 int band_rank, int band_world, int local_rows, void* d_rgba8, void* d_id, void* d_i32,
                 cudaStream_t stream, int out_global_rows, int progress_chunks, const int* progress_bound,
                 void* done_signal, int bins_lane) {
    GridParams p;
#ifdef RTB_EXPERIMENTS
    p.trace = nullptr;
#endif
    p.tile_y0 = 0;
    p.out_global_rows = out_global_rows;
    static const int fence_sys = [] { const char* e = std::getenv("RTB_PROGRESS_FENCE_SYSTEM"); return (e && e[0] == '1') ? 1 : 0; }();
    p.fence_system = fence_sys;
    p.tile_cull = effective_cull(ctx);
    p.bins = nullptr;
    p.bin_words = 0;
    if (p.tile_cull == 2) {
        const int n_bound = bound_count(ctx);
        const size_t words = grid_bin_words(width, local_rows, n_bound);
        void** bins = bins_lane ? &ctx->d_bins_seq : &ctx->d_bins;
        size_t* bins_cap = bins_lane ? &ctx->d_bins_seq_cap : &ctx->d_bins_cap;
        if (*bins_cap < (words ? words : 1) * sizeof(uint32_t)) RTB_CUDA(ctx, cudaDeviceSynchronize());  // an earlier launch may read the old buffer
        const int rc = ensure_device(ctx, bins, bins_cap, (words ? words : 1) * sizeof(uint32_t));
        if (rc != RTB_OK) return rc;
        p.bins = static_cast<uint32_t*>(*bins);
        p.bin_words = (n_bound + 31) / 32;
    }
    p.done_counter = ctx->d_progress + 2 * kMaxProgressChunks;
    p.done_signal = static_cast<unsigned int*>(done_signal);
    p.progress = progress_chunks > 0 ? ctx->d_progress : nullptr;
    p.progress_chunks = progress_chunks;
    for (int i = 0; i <= kMaxProgressChunks; ++i) p.progress_bound[i] = (progress_chunks > 0 && i <= progress_chunks) ? progress_bound[i] : 0;
    p.progress_epoch = ctx->progress_epoch;
    p.width = width; p.height = height;
    p.band_rows = band_rows; p.band_rank = band_rank; p.band_world = band_world;
    p.local_rows = local_rows;
    p.n_cubes = ctx->nc; p.n_spheres = ctx->ns;
    p.cube_rec = fs->rec;
    p.sph_rec = fs->rec + 3 * (size_t)ctx->nc;
    p.bound_rec = fs->rec + bound_offset(ctx);
    p.n_bound = bound_count(ctx);
    p.row_tab = (p.tile_cull == 0) ? fs->tab : nullptr;
    p.n_pad = p.n_bound;
    p.cube_col = ctx->d_cube_col;
    p.sph_col = ctx->d_sph_col;
    p.out_rgba8 = static_cast<uint32_t*>(d_rgba8);
    p.out_id = static_cast<int32_t*>(d_id);
    p.out_i32 = static_cast<int4*>(d_i32);
    p.vec_ok = (width % 4 == 0) && aligned16(d_rgba8) && aligned16(d_id) && aligned16(d_i32);
    p.cam = cam;
    EventPair ev;
    const int rc = take_events(ctx, &ev);
    if (rc != RTB_OK) return rc;
    RTB_CUDA(ctx, cudaEventRecord(ev.a, stream));
    if (!ctx->tail_lane.stream) {  // a low-priority stream for the short-lived tail CTAs of a dense launch
        int lo = 0, hi = 0;
        RTB_CUDA(ctx, cudaDeviceGetStreamPriorityRange(&lo, &hi));
        RTB_CUDA(ctx, cudaStreamCreateWithPriority(&ctx->tail_lane.stream, cudaStreamNonBlocking, lo));
        RTB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->tail_lane.fork, cudaEventDisableTiming));
        RTB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->tail_lane.join, cudaEventDisableTiming));
    }
    RTB_CUDA(ctx, launch_grid(p, ctx->sm_count, stream, &ctx->tail_lane));
    RTB_CUDA(ctx, cudaEventRecord(ev.b, stream));
    ctx->ev_pending.push_back(ev);
    ctx->launches += (p.tile_cull == 2 && p.bin_words > 0) ? 2 : 1;  // bin_tiles_kernel + raycast_binned_kernel
    if (ctx->ev_pending.size() > 1024) return resolve_pending(ctx);
    return RTB_OK;
}

int check_frame_args(rtb_ctx* ctx, int width, int height, const float* o0, const float* du, const float* dv,
                     const float* dir) {
    if (!ctx) return RTB_ERR_INVALID_ARG;
    if (width <= 0 || height <= 0 || !o0 || !du || !dv || !dir)
        return fail(ctx, RTB_ERR_INVALID_ARG, "bad frame size or null camera vector");
    if ((long long)width * height > (1LL << 31) - 1) return fail(ctx, RTB_ERR_INVALID_ARG, "frame too large");
    if (!ctx->have_scene) return fail(ctx, RTB_ERR_NO_SCENE, "rtb_set_scene has not been called");
    // A deferred NOT_A_BOX verdict (RTB_OPT_DEFER_SCENE_CHECK) is reported here once the upload has finished, else by
    // the blocking render calls after their own wait (verdict_after_sync): nothing waits for the scene upload alone.
    if (ctx->scene_check_pending && cudaEventQuery(ctx->ev_scene) == cudaSuccess) return collect_scene_verdict(ctx);
    (void)cudaGetLastError();
    return RTB_OK;
}

}  // namespace rtb

namespace {

void cam16_of(const float* o0, const float* du, const float* dv, const float* dir, float out[16]) {
    std::memcpy(out, o0, 16); std::memcpy(out + 4, du, 16);
    std::memcpy(out + 8, dv, 16); std::memcpy(out + 12, dir, 16);
}

}  // namespace

extern "C" {

int rtb_band_local_rows(int height, int band_rows, int band_rank, int band_world) {
    if (height <= 0 || band_rows <= 0 || band_world <= 0 || band_rank < 0 || band_rank >= band_world) return 0;
    const int n_bands = (height + band_rows - 1) / band_rows;
    int rows = 0;
    for (int b = band_rank; b < n_bands; b += band_world) {
        const int y0 = b * band_rows;
        rows += (height - y0 < band_rows) ? (height - y0) : band_rows;
    }
    return rows;
}

#define RTB_SINGLE_ONLY(ctx)                                                                                         \
    do {                                                                                                             \
        if ((ctx) && (ctx)->multi)                                                                                   \
            return fail(ctx, RTB_ERR_UNSUPPORTED, "device-pointer entry points need a single-device context "        \
                                                  "(a multi-device context renders with rtb_render / rtb_render_sequence_device)"); \
    } while (0)

int rtb_render_bands_device(rtb_ctx* ctx, int width, int height, const float origin0[4], const float dU[4],
                            const float dV[4], const float rayDir[4], int band_rows, int band_rank,
                            int band_world, void* d_rgba8, void* d_prim_id, void* d_rgba_i32,
                            void* cuda_stream) {
    RTB_SINGLE_ONLY(ctx);
    int rc = check_frame_args(ctx, width, height, origin0, dU, dV, rayDir);
    if (rc != RTB_OK) return rc;
    if (band_rows <= 0 || band_rows % 4 != 0 || band_world <= 0 || band_rank < 0 || band_rank >= band_world)
        return fail(ctx, RTB_ERR_INVALID_ARG, "band_rows must be a positive multiple of 4; 0 <= rank < world");
    RTB_CUDA(ctx, cudaSetDevice(ctx->device));
    CameraFrame cf;
    if (!camera_prepare(origin0, dU, dV, rayDir, &cf))
        return fail(ctx, RTB_ERR_CAMERA, "degenerate camera (dU, dV, rayDir do not span a ray grid)");
    cudaStream_t stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->stream;
    rc = order_streams(ctx, stream);
    if (rc != RTB_OK) return rc;
    float cam16[16];
    cam16_of(origin0, dU, dV, rayDir, cam16);
    const FrameSet* fs = nullptr;
    rc = prepare_frame(ctx, cf, width, height, cam16, band_rows, band_rank, band_world, stream, &fs);
    if (rc == RTB_OK) {
        const int local_rows = rtb_band_local_rows(height, band_rows, band_rank, band_world);
        rc = enqueue_grid(ctx, cf.cam, fs, width, height, band_rows, band_rank, band_world, local_rows,
                          d_rgba8, d_prim_id, d_rgba_i32, stream);
    }
    if (rc != RTB_OK) { drain(ctx, stream); return rc; }
    if (!cuda_stream) RTB_CUDA(ctx, cudaStreamSynchronize(stream));
    return RTB_OK;
}

int rtb_render_bands_frame(rtb_ctx* ctx, int width, int height, const float origin0[4], const float dU[4],
                           const float dV[4], const float rayDir[4], int band_rows, int band_rank,
                           int band_world, void* d_frame_rgba8, void* cuda_stream) {
    return rtb_render_bands_frame_signal(ctx, width, height, origin0, dU, dV, rayDir, band_rows, band_rank, band_world,
                                         d_frame_rgba8, nullptr, cuda_stream);
}

size_t rtb_peer_frame_signal_offset(size_t bytes) { return (bytes + 255) & ~(size_t)255; }

int rtb_stream_wait_signal(rtb_ctx* ctx, void* d_signal, unsigned int value, void* cuda_stream) {
    if (!ctx || !d_signal) return RTB_ERR_INVALID_ARG;
    if (!ctx->stream_wait_value32) return fail(ctx, RTB_ERR_UNSUPPORTED, "stream memory operations are unavailable");
    RTB_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->stream;
    if (ctx->stream_wait_value32(stream, reinterpret_cast<CUdeviceptr>(d_signal), value, CU_STREAM_WAIT_VALUE_GEQ) !=
        CUDA_SUCCESS)
        return fail(ctx, RTB_ERR_CUDA, "cuStreamWaitValue32 failed");
    if (!cuda_stream) RTB_CUDA(ctx, cudaStreamSynchronize(stream));
    return RTB_OK;
}

int rtb_render_bands_frame_signal(rtb_ctx* ctx, int width, int height, const float origin0[4], const float dU[4],
                                  const float dV[4], const float rayDir[4], int band_rows, int band_rank,
                                  int band_world, void* d_frame_rgba8, void* d_signal, void* cuda_stream) {
    RTB_SINGLE_ONLY(ctx);
    int rc = check_frame_args(ctx, width, height, origin0, dU, dV, rayDir);
    if (rc != RTB_OK) return rc;
    if (band_rows <= 0 || band_rows % 4 != 0 || band_world <= 0 || band_rank < 0 || band_rank >= band_world ||
        !d_frame_rgba8)
        return fail(ctx, RTB_ERR_INVALID_ARG, "band_rows must be a positive multiple of 4; 0 <= rank < world");
    RTB_CUDA(ctx, cudaSetDevice(ctx->device));
    CameraFrame cf;
    if (!camera_prepare(origin0, dU, dV, rayDir, &cf))
        return fail(ctx, RTB_ERR_CAMERA, "degenerate camera (dU, dV, rayDir do not span a ray grid)");
    cudaStream_t stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->stream;
    float cam16[16];
    cam16_of(origin0, dU, dV, rayDir, cam16);
    rc = enqueue_bands_frame(ctx, width, height, cam16, cf, band_rows, band_rank, band_world, d_frame_rgba8, d_signal, stream);
    if (rc != RTB_OK) { drain(ctx, stream); return rc; }
    if (!cuda_stream) RTB_CUDA(ctx, cudaStreamSynchronize(stream));
    return RTB_OK;
}

int rtb_peer_frame_create(rtb_ctx* ctx, size_t bytes, void** d_ptr, unsigned char handle_out[64]) {
    if (!ctx || !d_ptr || !handle_out || bytes == 0) return RTB_ERR_INVALID_ARG;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "rtb.h promises a 64-byte handle");
    RTB_CUDA(ctx, cudaSetDevice(ctx->device));
    *d_ptr = nullptr;
    // a whole allocation (IPC handles name allocations): the frame(s), then 256 bytes of signal words
    const size_t sig = rtb_peer_frame_signal_offset(bytes);
    RTB_CUDA(ctx, cudaMalloc(d_ptr, sig + 256));
    RTB_CUDA(ctx, cudaMemset(static_cast<uint8_t*>(*d_ptr) + sig, 0, 256));
    cudaIpcMemHandle_t h;
    const cudaError_t e = cudaIpcGetMemHandle(&h, *d_ptr);
    if (e != cudaSuccess) {
        cudaFree(*d_ptr);
        *d_ptr = nullptr;
        return fail(ctx, RTB_ERR_CUDA, std::string("cudaIpcGetMemHandle: ") + cudaGetErrorString(e));
    }
    std::memcpy(handle_out, &h, 64);
    return RTB_OK;
}

int rtb_peer_frame_open(rtb_ctx* ctx, const unsigned char handle[64], void** d_ptr) {
    if (!ctx || !d_ptr || !handle) return RTB_ERR_INVALID_ARG;
    RTB_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, 64);
    *d_ptr = nullptr;
    RTB_CUDA(ctx, cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return RTB_OK;
}

int rtb_peer_frame_close(rtb_ctx* ctx, void* d_ptr, int owner) {
    if (!ctx || !d_ptr) return RTB_ERR_INVALID_ARG;
    RTB_CUDA(ctx, cudaSetDevice(ctx->device));
    if (owner) RTB_CUDA(ctx, cudaFree(d_ptr));
    else RTB_CUDA(ctx, cudaIpcCloseMemHandle(d_ptr));
    return RTB_OK;
}

int rtb_assemble_bands_device(rtb_ctx* ctx, int width, int height, int band_rows, int band_world,
                               int elem_bytes, const void* d_packed, void* d_frame, void* cuda_stream) {
    if (!ctx) return RTB_ERR_INVALID_ARG;
    RTB_SINGLE_ONLY(ctx);
    if (width <= 0 || height <= 0 || band_rows <= 0 || band_world <= 0 || band_world > kMaxBandWorld ||
        (elem_bytes != 4 && elem_bytes != 16) || !d_packed || !d_frame)
        return fail(ctx, RTB_ERR_INVALID_ARG, "bad assemble arguments (elem_bytes is 4 or 16; world <= 64)");
    RTB_CUDA(ctx, cudaSetDevice(ctx->device));
    AssembleParams p;
    p.height = height; p.band_rows = band_rows; p.band_world = band_world;
    p.row_bytes = (size_t)width * (size_t)elem_bytes;
    p.src = static_cast<const uint8_t*>(d_packed);
    p.dst = static_cast<uint8_t*>(d_frame);
    int off = 0;
    for (int r = 0; r < kMaxBandWorld; ++r) {
        p.row_off[r] = off;
        if (r < band_world) off += rtb_band_local_rows(height, band_rows, r, band_world);
    }
    cudaStream_t stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->stream;
    RTB_CUDA(ctx, launch_assemble(p, stream));
    ctx->launches += 1;
    if (!cuda_stream) RTB_CUDA(ctx, cudaStreamSynchronize(stream));
    return RTB_OK;
}

}  // extern "C"

namespace {

// Body of rtb_render on a single-device context; any failure after the first enqueue is followed by drain().
int render_single(rtb_ctx* ctx, int width, int height, const float origin0[4], const float dU[4], const float dV[4],
                  const float rayDir[4], const CameraFrame& cf, void* const host[3], float* kernel_ms, float* total_ms) {
    const size_t n = (size_t)width * height;
    const size_t bytes[3] = {4 * n, 4 * n, 16 * n};
    int rc;
    for (int i = 0; i < 3; ++i)
        if (host[i]) {
            if (ctx->d_out_cap[i] < bytes[i]) RTB_CUDA(ctx, cudaDeviceSynchronize());
            rc = ensure_device(ctx, &ctx->d_out[i], &ctx->d_out_cap[i], bytes[i]);
            if (rc != RTB_OK) return rc;
        }
    rc = resolve_pending(ctx);
    if (rc != RTB_OK) return rc;
    const double k_sum0 = ctx->kernel_ms_sum;
    cudaStream_t st = ctx->stream;
    rc = order_streams(ctx, st);
    if (rc != RTB_OK) return rc;
    RTB_CUDA(ctx, cudaEventRecord(ctx->ev_total0, st));
    float cam16[16];
    cam16_of(origin0, dU, dV, rayDir, cam16);
    const FrameSet* fs = nullptr;
    rc = prepare_frame(ctx, cf, width, height, cam16, 16, 0, 1, st, &fs);
    if (rc != RTB_OK) return rc;
    const bool any_host = host[0] || host[1] || host[2];
    const size_t elem[3] = {4, 4, 16};
    size_t row_bytes = 0;
    for (int i = 0; i < 3; ++i)
        if (host[i]) row_bytes += elem[i] * (size_t)width;
    cudaStream_t cs = st;
    const int cull = effective_cull(ctx);
    // With a tile cull the launch is far shorter than the frame's PCIe time (C3: 0.03 ms against 0.6 ms): nothing to
    // overlap, so the frame goes out as ONE copy behind the launch instead of flagged chunks.
    if (any_host && ctx->stream_wait_value32 && height >= 64 && cull == 0) {
        // ONE launch for the whole frame; the kernel publishes finished row chunks (plan_progress_chunks)
        // and copy_stream copies chunk k to the host as soon as its flag shows this frame's epoch, while
        // later rows are still being rendered.  (The reference's OpenCL path does one blocking
        // clEnqueueMapBuffer after the kernel, MainState.cpp:876-886.)
        int bound[kMaxProgressChunks + 1];
        const int n_chunks = plan_progress_chunks(height, row_bytes, 16, bound);
        cs = ctx->copy_stream;
        ctx->progress_epoch += 1;
        if (ctx->progress_epoch == 0) ctx->progress_epoch = 1;
        rc = enqueue_grid(ctx, cf.cam, fs, width, height, 16, 0, 1, height, host[0] ? ctx->d_out[0] : nullptr,
                          host[1] ? ctx->d_out[1] : nullptr, host[2] ? ctx->d_out[2] : nullptr, st, 0, n_chunks, bound);
        if (rc != RTB_OK) return rc;
        for (int k = 0; k < n_chunks; ++k) {
            const int y0 = bound[k], rows = bound[k + 1] - bound[k];
            const CUdeviceptr flag = reinterpret_cast<CUdeviceptr>(ctx->d_progress + kMaxProgressChunks + k);
            if (ctx->stream_wait_value32(cs, flag, ctx->progress_epoch, CU_STREAM_WAIT_VALUE_EQ) != CUDA_SUCCESS)
                return fail(ctx, RTB_ERR_CUDA, "cuStreamWaitValue32 failed");
            for (int i = 0; i < 3; ++i)
                if (host[i]) {
                    const size_t off = elem[i] * (size_t)y0 * width, nb = elem[i] * (size_t)rows * width;
                    RTB_CUDA(ctx, cudaMemcpyAsync(static_cast<uint8_t*>(host[i]) + off,
                                                  static_cast<uint8_t*>(ctx->d_out[i]) + off, nb,
                                                  cudaMemcpyDeviceToHost, cs));
                    ctx->d2h_bytes += nb;
                }
        }
    } else {
        // Culled modes: one launch, one copy.  Dense without stream memory operations: up to 4 launches of whole-row
        // chunks, chunk k travelling to the host on copy_stream while chunk k+1 is rendered.
        int n_chunks = (any_host && height >= 256 && cull == 0) ? 4 : 1;
        int chunk_rows = ((height + n_chunks - 1) / n_chunks + 7) & ~7;
        n_chunks = (height + chunk_rows - 1) / chunk_rows;
        cs = (n_chunks > 1) ? ctx->copy_stream : st;
        for (int k = 0; k < n_chunks; ++k) {
            const int y0 = k * chunk_rows;
            const int rows = (height - y0 < chunk_rows) ? (height - y0) : chunk_rows;
            void* d[3];
            for (int i = 0; i < 3; ++i)
                d[i] = host[i] ? static_cast<uint8_t*>(ctx->d_out[i]) + elem[i] * (size_t)y0 * width : nullptr;
            rc = enqueue_grid(ctx, cf.cam, fs, width, height, chunk_rows, k, n_chunks, rows, d[0], d[1], d[2], st);
            if (rc != RTB_OK) return rc;
            if (n_chunks > 1) {
                RTB_CUDA(ctx, cudaEventRecord(ctx->ev_chunk[k], st));
                RTB_CUDA(ctx, cudaStreamWaitEvent(cs, ctx->ev_chunk[k], 0));
            }
            for (int i = 0; i < 3; ++i)
                if (host[i]) {
                    const size_t nb = elem[i] * (size_t)rows * width;
                    RTB_CUDA(ctx, cudaMemcpyAsync(static_cast<uint8_t*>(host[i]) + elem[i] * (size_t)y0 * width, d[i], nb,
                                                  cudaMemcpyDeviceToHost, cs));
                    ctx->d2h_bytes += nb;
                }
        }
    }
    RTB_CUDA(ctx, cudaEventRecord(ctx->ev_total1, cs));
    RTB_CUDA(ctx, cudaStreamSynchronize(cs));
    RTB_CUDA(ctx, cudaStreamSynchronize(st));
    rc = resolve_pending(ctx);
    if (rc != RTB_OK) return rc;
    if (kernel_ms) *kernel_ms = (float)(ctx->kernel_ms_sum - k_sum0);
    if (total_ms) RTB_CUDA(ctx, cudaEventElapsedTime(total_ms, ctx->ev_total0, ctx->ev_total1));
    return collect_scene_verdict(ctx);  // a deferred NOT_A_BOX verdict: the upload has finished by now
}

}  // namespace

namespace rtb {

// This rank's bands rendered and copied into whole HOST frames (row = frame row; each of host[0..2] = rgba8, prim_id,
// rgba_i32 nullable): body of rtb_render_bands_host and of a multi-device rtb_render.  Blocking.
int bands_host_render(rtb_ctx* ctx, int width, int height, const float* cam16, const CameraFrame& cf, int band_rows,
                      int band_rank, int band_world, void* const host[3], float* kernel_ms) {
    const int local_rows = rtb_band_local_rows(height, band_rows, band_rank, band_world);
    const size_t elem[3] = {4, 4, 16};
    int rc;
    size_t row_bytes_all = 0;
    for (int i = 0; i < 3; ++i)
        if (host[i]) {
            const size_t need = elem[i] * (size_t)width * (size_t)(local_rows > 0 ? local_rows : 1);
            if (ctx->d_out_cap[i] < need) RTB_CUDA(ctx, cudaDeviceSynchronize());
            rc = ensure_device(ctx, &ctx->d_out[i], &ctx->d_out_cap[i], need);
            if (rc != RTB_OK) return rc;
            row_bytes_all += elem[i] * (size_t)width;
        }
    rc = resolve_pending(ctx);
    if (rc != RTB_OK) return rc;
    const double k_sum0 = ctx->kernel_ms_sum;
    cudaStream_t st = ctx->stream;
    rc = order_streams(ctx, st);
    if (rc != RTB_OK) return rc;
    const FrameSet* fs = nullptr;
    rc = prepare_frame(ctx, cf, width, height, cam16, band_rows, band_rank, band_world, st, &fs);
    if (rc != RTB_OK) return rc;
    if (local_rows > 0 && row_bytes_all > 0) {
        // chunk = whole bands (and a multiple of 16 rows, the progress granularity): about 2-4 MB each; with a tile
        // cull the launch is far shorter than the copy, so the rows go out in one piece behind it
        const int cull = effective_cull(ctx);
        int unit = band_rows;
        while (unit % (cull == 2 ? binned_tile_rows() : 16) != 0) unit *= 2;
        int bound[kMaxProgressChunks + 1];
        int n_chunks = plan_progress_chunks(local_rows, row_bytes_all, unit, bound);
        const bool flags = ctx->stream_wait_value32 != nullptr && cull == 0;
        if (!flags) { n_chunks = 1; bound[0] = 0; bound[1] = local_rows; }
        cudaStream_t cs = flags ? ctx->copy_stream : st;
        if (flags) {
            ctx->progress_epoch += 1;
            if (ctx->progress_epoch == 0) ctx->progress_epoch = 1;
        }
        rc = enqueue_grid(ctx, cf.cam, fs, width, height, band_rows, band_rank, band_world, local_rows,
                          host[0] ? ctx->d_out[0] : nullptr, host[1] ? ctx->d_out[1] : nullptr,
                          host[2] ? ctx->d_out[2] : nullptr, st, 0, flags ? n_chunks : 0, bound);
        if (rc != RTB_OK) return rc;
        for (int k = 0; k < n_chunks; ++k) {
            const int l0 = bound[k], rows = bound[k + 1] - bound[k];
            if (flags) {
                const CUdeviceptr flag = reinterpret_cast<CUdeviceptr>(ctx->d_progress + kMaxProgressChunks + k);
                if (ctx->stream_wait_value32(cs, flag, ctx->progress_epoch, CU_STREAM_WAIT_VALUE_EQ) != CUDA_SUCCESS)
                    return fail(ctx, RTB_ERR_CUDA, "cuStreamWaitValue32 failed");
            }
            // local rows [l0, l0+rows) = whole bands lb0.. (the frame's last band may be short): one strided
            // copy moves them to their frame rows (band lb -> frame band lb*world + rank)
            const int lb0 = l0 / band_rows;
            const int full = rows / band_rows, tail = rows - full * band_rows;
            for (int i = 0; i < 3; ++i) {
                if (!host[i]) continue;
                const size_t row_bytes = elem[i] * (size_t)width, band_bytes = row_bytes * (size_t)band_rows;
                const uint8_t* src = static_cast<const uint8_t*>(ctx->d_out[i]) + row_bytes * (size_t)l0;
                uint8_t* dst = static_cast<uint8_t*>(host[i]) + band_bytes * ((size_t)lb0 * band_world + band_rank);
                if (band_world == 1) {  // contiguous
                    RTB_CUDA(ctx, cudaMemcpyAsync(dst, src, row_bytes * (size_t)rows, cudaMemcpyDeviceToHost, cs));
                } else {
                    if (full > 0)
                        RTB_CUDA(ctx, cudaMemcpy2DAsync(dst, band_bytes * (size_t)band_world, src, band_bytes, band_bytes,
                                                        (size_t)full, cudaMemcpyDeviceToHost, cs));
                    if (tail > 0)
                        RTB_CUDA(ctx, cudaMemcpyAsync(dst + band_bytes * (size_t)band_world * full, src + band_bytes * full,
                                                      row_bytes * (size_t)tail, cudaMemcpyDeviceToHost, cs));
                }
                ctx->d2h_bytes += row_bytes * (size_t)rows;
            }
        }
        RTB_CUDA(ctx, cudaStreamSynchronize(cs));
    }
    RTB_CUDA(ctx, cudaStreamSynchronize(st));
    rc = resolve_pending(ctx);
    if (rc != RTB_OK) return rc;
    if (kernel_ms) *kernel_ms = (float)(ctx->kernel_ms_sum - k_sum0);
    return collect_scene_verdict(ctx);
}

// This rank's bands straight into a WHOLE rgba8 frame (own or peer memory), optional completion signal: body of
// rtb_render_bands_frame_signal and of the group renders.  Enqueue only.
int enqueue_bands_frame(rtb_ctx* ctx, int width, int height, const float* cam16, const CameraFrame& cf, int band_rows,
                        int band_rank, int band_world, void* d_frame_rgba8, void* d_signal, cudaStream_t stream) {
    int rc = order_streams(ctx, stream);
    if (rc != RTB_OK) return rc;
    const FrameSet* fs = nullptr;
    rc = prepare_frame(ctx, cf, width, height, cam16, band_rows, band_rank, band_world, stream, &fs);
    if (rc != RTB_OK) return rc;
    const int local_rows = rtb_band_local_rows(height, band_rows, band_rank, band_world);
    if (local_rows <= 0 && d_signal) {  // nothing to render: the signal still has to arrive
        static_assert(sizeof(unsigned int) == 4, "signal words are 32 bit");
        const cudaError_t e = launch_signal_only(static_cast<unsigned int*>(d_signal), stream);
        if (e != cudaSuccess) return fail(ctx, RTB_ERR_CUDA, std::string("signal launch: ") + cudaGetErrorString(e));
    }
    return enqueue_grid(ctx, cf.cam, fs, width, height, band_rows, band_rank, band_world, local_rows, d_frame_rgba8,
                        nullptr, nullptr, stream, 1, 0, nullptr, d_signal);
}

}  // namespace rtb

extern "C" {

int rtb_render(rtb_ctx* ctx, int width, int height, const float origin0[4], const float dU[4],
               const float dV[4], const float rayDir[4], uint8_t* rgba8_out, int32_t* prim_id_out,
               int32_t* rgba_i32_out, float* kernel_ms, float* total_ms) {
    if (ctx && ctx->multi)
        return group_render_local(ctx, width, height, origin0, dU, dV, rayDir, rgba8_out, prim_id_out, rgba_i32_out,
                                  kernel_ms, total_ms);
    int rc = check_frame_args(ctx, width, height, origin0, dU, dV, rayDir);
    if (rc != RTB_OK) return rc;
    RTB_CUDA(ctx, cudaSetDevice(ctx->device));
    CameraFrame cf;
    if (!camera_prepare(origin0, dU, dV, rayDir, &cf))
        return fail(ctx, RTB_ERR_CAMERA, "degenerate camera (dU, dV, rayDir do not span a ray grid)");
    void* const host[3] = {rgba8_out, prim_id_out, rgba_i32_out};
    rc = render_single(ctx, width, height, origin0, dU, dV, rayDir, cf, host, kernel_ms, total_ms);
    if (rc != RTB_OK) drain(ctx, nullptr);  // the call is blocking: nothing may still write into the caller's buffers
    return rc;
}

int rtb_render_bands_host(rtb_ctx* ctx, int width, int height, const float origin0[4], const float dU[4],
                          const float dV[4], const float rayDir[4], int band_rows, int band_rank, int band_world,
                          uint8_t* host_frame_rgba8, float* kernel_ms) {
    RTB_SINGLE_ONLY(ctx);
    int rc = check_frame_args(ctx, width, height, origin0, dU, dV, rayDir);
    if (rc != RTB_OK) return rc;
    if (band_rows <= 0 || band_rows % 4 != 0 || band_world <= 0 || band_rank < 0 || band_rank >= band_world ||
        !host_frame_rgba8)
        return fail(ctx, RTB_ERR_INVALID_ARG, "band_rows must be a positive multiple of 4; 0 <= rank < world");
    RTB_CUDA(ctx, cudaSetDevice(ctx->device));
    CameraFrame cf;
    if (!camera_prepare(origin0, dU, dV, rayDir, &cf))
        return fail(ctx, RTB_ERR_CAMERA, "degenerate camera (dU, dV, rayDir do not span a ray grid)");
    float cam16[16];
    cam16_of(origin0, dU, dV, rayDir, cam16);
    void* const host[3] = {host_frame_rgba8, nullptr, nullptr};
    rc = bands_host_render(ctx, width, height, cam16, cf, band_rows, band_rank, band_world, host, kernel_ms);
    if (rc != RTB_OK) drain(ctx, nullptr);
    return rc;
}

int rtb_host_register(void* p, size_t bytes) {
    if (!p || bytes == 0) return RTB_ERR_INVALID_ARG;
    const cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable);
    if (e == cudaErrorMemoryAllocation) return RTB_ERR_OUT_OF_MEMORY;
    if (e != cudaSuccess) (void)cudaGetLastError();
    return e == cudaSuccess ? RTB_OK : RTB_ERR_CUDA;
}

int rtb_host_unregister(void* p) {
    if (!p) return RTB_ERR_INVALID_ARG;
    const cudaError_t e = cudaHostUnregister(p);
    if (e != cudaSuccess) (void)cudaGetLastError();
    return e == cudaSuccess ? RTB_OK : RTB_ERR_CUDA;
}

int rtb_render_rays(rtb_ctx* ctx, size_t n_rays, const float* rayOrigins, const float rayDir[4],
                    uint8_t* rgba8_out, int32_t* prim_id_out, int32_t* rgba_i32_out, float* kernel_ms,
                    float* total_ms) {
    if (!ctx) return RTB_ERR_INVALID_ARG;
    if (ctx->multi) ctx = group_root_local(ctx);  // explicit origins: rendered by the first device alone
    if (!rayDir || (n_rays > 0 && !rayOrigins)) return fail(ctx, RTB_ERR_INVALID_ARG, "null ray array");
    if (!ctx->have_scene) return fail(ctx, RTB_ERR_NO_SCENE, "rtb_set_scene has not been called");
    int rc = collect_scene_verdict(ctx);
    if (rc != RTB_OK) return rc;
    RTB_CUDA(ctx, cudaSetDevice(ctx->device));
    const double dir3[3] = {rayDir[0], rayDir[1], rayDir[2]};
    if (!(dir3[0] * dir3[0] + dir3[1] * dir3[1] + dir3[2] * dir3[2] > 0.0))
        return fail(ctx, RTB_ERR_CAMERA, "rayDir.xyz is zero");
    const size_t bytes[3] = {4 * n_rays, 4 * n_rays, 16 * n_rays};
    void* host[3] = {rgba8_out, prim_id_out, rgba_i32_out};
    cudaStream_t st = ctx->stream;
    rc = order_streams(ctx, st);
    if (rc != RTB_OK) return rc;
    RTB_CUDA(ctx, cudaStreamSynchronize(st));  // buffers may be re-allocated below
    for (int i = 0; i < 3; ++i)
        if (host[i] && n_rays) {
            rc = ensure_device(ctx, &ctx->d_out[i], &ctx->d_out_cap[i], bytes[i]);
            if (rc != RTB_OK) return rc;
        }
    void* dorig = ctx->d_origins;
    rc = ensure_device(ctx, &dorig, &ctx->d_origins_cap, 16 * (n_rays ? n_rays : 1));
    ctx->d_origins = static_cast<float4*>(dorig);
    if (rc != RTB_OK) return rc;
    rc = ensure_host_scene(ctx);
    if (rc != RTB_OK) return rc;
    const size_t nrec = 4 * (size_t)ctx->nc + 2 * (size_t)ctx->ns;
    rc = ensure_frame_set(ctx, &ctx->cur, nrec + 8, 0);
    if (rc == RTB_OK) rc = ensure_host_records(ctx, nrec + 8);
    if (rc != RTB_OK) return rc;
    RTB_CUDA(ctx, cudaEventSynchronize(ctx->ev_rec));
    RTB_CUDA(ctx, cudaEventRecord(ctx->ev_total0, st));
    // world-space records: 4 float4 per cube, then 2 per sphere (centre, radius)
    ctx->cur.rec_valid = ctx->cur.tab_valid = false;
    float* h = reinterpret_cast<float*>(ctx->h_rec);
    for (int c = 0; c < ctx->nc; ++c) pack_box_world(dir3, ctx->boxes[c], h + 16 * (size_t)c);
    float* hs = h + 16 * (size_t)ctx->nc;
    for (int s = 0; s < ctx->ns; ++s) {
        std::memcpy(hs + 8 * (size_t)s, &ctx->sph_o[4 * (size_t)s], 16);
        hs[8 * (size_t)s + 4] = ctx->sph_r[s];
        hs[8 * (size_t)s + 5] = hs[8 * (size_t)s + 6] = hs[8 * (size_t)s + 7] = 0.f;
    }
    auto body = [&]() -> int {
        if (nrec)
            RTB_CUDA(ctx, cudaMemcpyAsync(ctx->cur.rec, ctx->h_rec, nrec * sizeof(float4), cudaMemcpyHostToDevice, st));
        RTB_CUDA(ctx, cudaEventRecord(ctx->ev_rec, st));
        if (n_rays)
            RTB_CUDA(ctx, cudaMemcpyAsync(ctx->d_origins, rayOrigins, 16 * n_rays, cudaMemcpyHostToDevice, st));
        ctx->h2d_bytes += nrec * sizeof(float4) + 16 * n_rays;
        RaysParams p;
        p.n_rays = n_rays;
        p.origins = ctx->d_origins;
        for (int k = 0; k < 4; ++k) p.dir[k] = rayDir[k];
        p.n_cubes = ctx->nc; p.n_spheres = ctx->ns;
        p.cube_rec = ctx->cur.rec;
        p.sph_rec = ctx->cur.rec + 4 * (size_t)ctx->nc;
        p.cube_col = ctx->d_cube_col;
        p.sph_col = ctx->d_sph_col;
        p.out_rgba8 = host[0] ? static_cast<uint32_t*>(ctx->d_out[0]) : nullptr;
        p.out_id = host[1] ? static_cast<int32_t*>(ctx->d_out[1]) : nullptr;
        p.out_i32 = host[2] ? static_cast<int4*>(ctx->d_out[2]) : nullptr;
        EventPair ev;
        int rc2 = take_events(ctx, &ev);
        if (rc2 != RTB_OK) return rc2;
        ctx->ev_free.push_back(ev);  // used synchronously below; stays owned by the pool
        RTB_CUDA(ctx, cudaEventRecord(ev.a, st));
        RTB_CUDA(ctx, launch_rays(p, st));
        RTB_CUDA(ctx, cudaEventRecord(ev.b, st));
        ctx->launches += (n_rays ? 1 : 0);
        for (int i = 0; i < 3; ++i)
            if (host[i] && n_rays)
                RTB_CUDA(ctx, cudaMemcpyAsync(host[i], ctx->d_out[i], bytes[i], cudaMemcpyDeviceToHost, st));
        RTB_CUDA(ctx, cudaEventRecord(ctx->ev_total1, st));
        RTB_CUDA(ctx, cudaStreamSynchronize(st));
        float kms = 0.f;
        RTB_CUDA(ctx, cudaEventElapsedTime(&kms, ev.a, ev.b));
        if (kernel_ms) *kernel_ms = kms;
        if (total_ms) RTB_CUDA(ctx, cudaEventElapsedTime(total_ms, ctx->ev_total0, ctx->ev_total1));
        return RTB_OK;
    };
    rc = body();
    if (rc != RTB_OK) drain(ctx, nullptr);
    return rc;
}

int rtb_render_sequence_device(rtb_ctx* ctx, int width, int height, int n_frames, const float* cameras,
                               void* d_rgba8, float* kernel_ms_total) {
    if (!ctx) return RTB_ERR_INVALID_ARG;
    if (ctx->multi) return group_render_sequence_local(ctx, width, height, n_frames, cameras, d_rgba8, kernel_ms_total);
    if (n_frames < 0 || (n_frames > 0 && (!cameras || !d_rgba8)))
        return fail(ctx, RTB_ERR_INVALID_ARG, "null cameras / output");
    if (n_frames == 0) { if (kernel_ms_total) *kernel_ms_total = 0.f; return RTB_OK; }
    int rc = check_frame_args(ctx, width, height, cameras, cameras + 4, cameras + 8, cameras + 12);
    if (rc != RTB_OK) return rc;
    RTB_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    rc = order_streams(ctx, st);
    if (rc != RTB_OK) return rc;
    RTB_CUDA(ctx, cudaStreamSynchronize(st));
    rc = resolve_pending(ctx);
    if (rc != RTB_OK) return rc;
    const double k0 = ctx->kernel_ms_sum;
    const size_t frame_bytes = 4 * (size_t)width * height;
    // The frame sets of up to kMaxCamSlots frames are packed ahead on pack_stream (small FP64 launches per
    // frame, ~20 us each) while the frames render: frame f's launch only waits for ITS pack.  Consecutive frames go
    // to alternating lanes (the context's stream / seq_stream, each with its own bin words): frames are independent,
    // so the next frame's first CTAs run in the SMs the current frame's ragged end leaves idle, and a binned frame's
    // bin launch overlaps the previous frame's render (RTB_SEQ_LANES=1: one lane, as before).
    static const bool two_lanes = [] { const char* e = std::getenv("RTB_SEQ_LANES"); return !(e && e[0] == '1'); }();
    if (two_lanes && !ctx->seq_stream) {
        RTB_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->seq_stream, cudaStreamNonBlocking));
        RTB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_seq_fork, cudaEventDisableTiming));
        RTB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_seq_join, cudaEventDisableTiming));
    }
    for (int f0 = 0; f0 < n_frames;) {
        rc = prepack_cameras(ctx, width, height, 16, 0, 1, n_frames - f0, cameras + 16 * (size_t)f0, st);
        if (rc != RTB_OK) { cudaDeviceSynchronize(); drain(ctx, nullptr); ctx->slots_used = 0; return rc; }
        const int batch = ctx->slots_used;
        if (two_lanes) {  // lane 1 starts behind whatever the context's stream holds
            RTB_CUDA(ctx, cudaEventRecord(ctx->ev_seq_fork, st));
            RTB_CUDA(ctx, cudaStreamWaitEvent(ctx->seq_stream, ctx->ev_seq_fork, 0));
        }
        for (int s = 0; s < batch; ++s) {
            const int f = f0 + s;
            const float* c = cameras + 16 * (size_t)f;
            CameraFrame cf;
            camera_prepare(c, c + 4, c + 8, c + 12, &cf);  // checked by prepack_cameras
            const FrameSet* fs = &ctx->slots[(size_t)s];
            const int lane = two_lanes ? (f & 1) : 0;
            cudaStream_t ls = lane ? ctx->seq_stream : st;
            rc = cudaStreamWaitEvent(ls, fs->ready, 0) == cudaSuccess ? RTB_OK : fail(ctx, RTB_ERR_CUDA, "cudaStreamWaitEvent");
            if (rc == RTB_OK)
                rc = enqueue_grid(ctx, cf.cam, fs, width, height, 16, 0, 1, height,
                                  static_cast<uint8_t*>(d_rgba8) + frame_bytes * (size_t)f, nullptr, nullptr, ls, 0, 0, nullptr,
                                  nullptr, lane);
            if (rc != RTB_OK) { cudaDeviceSynchronize(); drain(ctx, nullptr); ctx->slots_used = 0; return rc; }
        }
        if (two_lanes) {  // the next batch's packs (and the caller) see both lanes through the context's stream
            RTB_CUDA(ctx, cudaEventRecord(ctx->ev_seq_join, ctx->seq_stream));
            RTB_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_seq_join, 0));
        }
        f0 += batch;
    }
    ctx->slots_used = 0;  // nothing of this sequence is kept for later calls
    RTB_CUDA(ctx, cudaStreamSynchronize(st));
    rc = resolve_pending(ctx);
    if (rc != RTB_OK) return rc;
    if (kernel_ms_total) *kernel_ms_total = (float)(ctx->kernel_ms_sum - k0);
    return collect_scene_verdict(ctx);
}

int rtb_prepack_cameras_bands(rtb_ctx* ctx, int width, int height, int band_rows, int band_rank, int band_world,
                              int n_cameras, const float* cameras, void* stream) {
    if (!ctx) return RTB_ERR_INVALID_ARG;
    RTB_SINGLE_ONLY(ctx);
    if (n_cameras < 0 || (n_cameras > 0 && !cameras)) return fail(ctx, RTB_ERR_INVALID_ARG, "null cameras");
    if (band_rows <= 0 || band_rows % 4 != 0 || band_world <= 0 || band_rank < 0 || band_rank >= band_world)
        return fail(ctx, RTB_ERR_INVALID_ARG, "band_rows must be a positive multiple of 4; 0 <= rank < world");
    if (n_cameras == 0) { ctx->slots_used = 0; return RTB_OK; }
    int rc = check_frame_args(ctx, width, height, cameras, cameras + 4, cameras + 8, cameras + 12);
    if (rc != RTB_OK) return rc;
    RTB_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : ctx->stream;
    rc = order_streams(ctx, st);
    if (rc == RTB_OK) rc = prepack_cameras(ctx, width, height, band_rows, band_rank, band_world, n_cameras, cameras, st);
    if (rc != RTB_OK) ctx->slots_used = 0;
    return rc;
}

int rtb_prepack_cameras(rtb_ctx* ctx, int width, int height, int n_cameras, const float* cameras, void* stream) {
    return rtb_prepack_cameras_bands(ctx, width, height, 16, 0, 1, n_cameras, cameras, stream);
}

int rtb_host_alloc(void** out, size_t bytes) {
    if (!out) return RTB_ERR_INVALID_ARG;
    *out = nullptr;
    const cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocPortable);  // usable from every device
    if (e == cudaErrorMemoryAllocation) return RTB_ERR_OUT_OF_MEMORY;
    return e == cudaSuccess ? RTB_OK : RTB_ERR_CUDA;
}

void rtb_host_free(void* p) { if (p) cudaFreeHost(p); }

int rtb_device_info(const rtb_ctx* ctx, int* sm_count, int* sm_clock_khz, char* name, size_t name_cap) {
    if (!ctx) return RTB_ERR_INVALID_ARG;
    if (sm_count) *sm_count = ctx->sm_count;
    if (sm_clock_khz) *sm_clock_khz = ctx->clock_khz;
    if (name && name_cap) std::snprintf(name, name_cap, "%s", ctx->name);
    return RTB_OK;
}

int rtb_set_option(rtb_ctx* ctx, int option, int value) {
    if (!ctx) return RTB_ERR_INVALID_ARG;
    switch (option) {
        case RTB_OPT_TILE_CULL:
            if (value < 0 || value > 3) return fail(ctx, RTB_ERR_INVALID_ARG, "RTB_OPT_TILE_CULL takes 0, 1, 2 or 3 (auto)");
            ctx->tile_cull = value;
            break;
        case RTB_OPT_DEFER_SCENE_CHECK:
            if (value < 0 || value > 1) return fail(ctx, RTB_ERR_INVALID_ARG, "RTB_OPT_DEFER_SCENE_CHECK takes 0 or 1");
            ctx->defer_scene_check = value;
            break;
        default:
            return fail(ctx, RTB_ERR_INVALID_ARG, "unknown option");
    }
    if (ctx->multi) group_set_option_local(ctx, option, value);
    return RTB_OK;
}

int rtb_get_option(const rtb_ctx* ctx, int option, int* value) {
    if (!ctx || !value) return RTB_ERR_INVALID_ARG;
    switch (option) {
        case RTB_OPT_TILE_CULL: *value = ctx->tile_cull; return RTB_OK;
        case RTB_OPT_DEFER_SCENE_CHECK: *value = ctx->defer_scene_check; return RTB_OK;
        default: return RTB_ERR_INVALID_ARG;
    }
}

unsigned long long rtb_launch_count(const rtb_ctx* ctx) {
    if (!ctx) return 0ull;
    return ctx->multi ? group_launch_count_local(ctx) : ctx->launches;
}

void rtb_transfer_bytes(const rtb_ctx* ctx, unsigned long long* h2d, unsigned long long* d2h) {
    if (ctx && ctx->multi) { group_transfer_bytes_local(ctx, h2d, d2h); return; }
    if (h2d) *h2d = ctx ? ctx->h2d_bytes : 0ull;
    if (d2h) *d2h = ctx ? ctx->d2h_bytes : 0ull;
}

float rtb_kernel_ms_avg(rtb_ctx* ctx, int reset, int* n_launches) {
    if (!ctx) return 0.f;
    cudaSetDevice(ctx->device);
    resolve_pending(ctx);
    const int n = ctx->kernel_ms_n;
    const float avg = n ? (float)(ctx->kernel_ms_sum / n) : 0.f;
    if (n_launches) *n_launches = n;
    if (reset) { ctx->kernel_ms_sum = 0.0; ctx->kernel_ms_n = 0; }
    return avg;
}

int rtb_debug_pack(int numSpheres, const float* sphereOrigins, const float* sphereRadius, int numCubes,
                   const float* cubeVertices, const float origin0[4], const float dU[4], const float dV[4],
                   const float rayDir[4], float* cube_rec, float* sphere_rec, float* cam, int* bad_cube_index) {
    if (numSpheres < 0 || numCubes < 0 || !origin0 || !dU || !dV || !rayDir) return RTB_ERR_INVALID_ARG;
    if (bad_cube_index) *bad_cube_index = -1;
    CameraFrame cf;
    if (!camera_prepare(origin0, dU, dV, rayDir, &cf)) return RTB_ERR_CAMERA;
    for (int c = 0; c < numCubes; ++c) {
        Box b;
        if (!box_from_vertices(cubeVertices + 144 * (size_t)c, &b)) {
            if (bad_cube_index) *bad_cube_index = c;
            return RTB_ERR_NOT_A_BOX;
        }
        if (cube_rec) pack_box(cf, b, reinterpret_cast<RtbSlab*>(cube_rec) + 3 * (size_t)c);
    }
    if (sphere_rec)
        for (int s = 0; s < numSpheres; ++s)
            pack_sphere(cf, sphereOrigins + 4 * (size_t)s, sphereRadius[s],
                        reinterpret_cast<RtbSphere*>(sphere_rec) + s);
    if (cam) std::memcpy(cam, &cf.cam, sizeof(RtbCam));
    return RTB_OK;
}

int rtb_debug_pack_bounds(int numSpheres, const float* sphereOrigins, const float* sphereRadius, int numCubes,
                          const float* cubeVertices, int width, int height, const float origin0[4],
                          const float dU[4], const float dV[4], const float rayDir[4], float* bound_rec) {
    if (numSpheres < 0 || numCubes < 0 || width <= 0 || height <= 0 || !origin0 || !dU || !dV || !rayDir || !bound_rec)
        return RTB_ERR_INVALID_ARG;
    CameraFrame cf;
    if (!camera_prepare(origin0, dU, dV, rayDir, &cf)) return RTB_ERR_CAMERA;
    const FrameExtent fe = frame_extent(cf, width, height);
    RtbBound* out = reinterpret_cast<RtbBound*>(bound_rec);
    for (int c = 0; c < numCubes; ++c) {
        Box b;
        if (!box_from_vertices(cubeVertices + 144 * (size_t)c, &b)) return RTB_ERR_NOT_A_BOX;
        RtbSlab slabs[3];
        pack_box(cf, b, slabs);
        pack_bound_box(fe, slabs, out + c);
    }
    for (int s = 0; s < numSpheres; ++s) {
        RtbSphere sp;
        pack_sphere(cf, sphereOrigins + 4 * (size_t)s, sphereRadius[s], &sp);
        pack_bound_sphere(fe, sp, out + numCubes + s);
    }
    return RTB_OK;
}

int rtb_debug_pack_row_table(int n_bound, const float* bound_rec, int width, int height, const float origin0[4],
                             const float dU[4], const float dV[4], const float rayDir[4], float* table) {
    if (n_bound < 0 || width <= 0 || height <= 0 || !origin0 || !dU || !dV || !rayDir || !table || (n_bound && !bound_rec))
        return RTB_ERR_INVALID_ARG;
    CameraFrame cf;
    if (!camera_prepare(origin0, dU, dV, rayDir, &cf)) return RTB_ERR_CAMERA;
    const FrameExtent fe = frame_extent(cf, width, height);
    const RtbBound* b = reinterpret_cast<const RtbBound*>(bound_rec);
    for (int y = 0; y < height; ++y) {
        const float v = rtb_cam_v(cf.cam, (float)y);
        for (int i = 0; i < n_bound; ++i) pack_row_entry(b[i], v, fe.uabs, table + 2 * ((size_t)y * n_bound + i));
    }
    return RTB_OK;
}

int rtb_debug_device_records(rtb_ctx* ctx, int width, int height, const float origin0[4], const float dU[4],
                             const float dV[4], const float rayDir[4], float* out, size_t out_cap_floats,
                             size_t* n_floats, size_t* bound_offset_floats) {
    if (ctx && ctx->multi) ctx = group_root_local(ctx);
    int rc = check_frame_args(ctx, width, height, origin0, dU, dV, rayDir);
    if (rc != RTB_OK) return rc;
    RTB_CUDA(ctx, cudaSetDevice(ctx->device));
    CameraFrame cf;
    if (!camera_prepare(origin0, dU, dV, rayDir, &cf)) return fail(ctx, RTB_ERR_CAMERA, "degenerate camera");
    const size_t n = 4 * frame_records(ctx);
    if (n_floats) *n_floats = n;
    if (bound_offset_floats) *bound_offset_floats = 4 * bound_offset(ctx);
    if (!out) return RTB_OK;
    if (out_cap_floats < n) return fail(ctx, RTB_ERR_INVALID_ARG, "record buffer too small");
    rc = order_streams(ctx, ctx->stream);
    if (rc != RTB_OK) return rc;
    RTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float cam16[16];
    cam16_of(origin0, dU, dV, rayDir, cam16);
    rc = enqueue_pack(ctx, cf, width, height, cam16, &ctx->cur, ctx->stream);
    if (rc != RTB_OK) return rc;
    RTB_CUDA(ctx, cudaMemcpyAsync(out, ctx->cur.rec, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    RTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return RTB_OK;
}

int rtb_debug_device_row_table(rtb_ctx* ctx, int width, int height, const float origin0[4], const float dU[4],
                               const float dV[4], const float rayDir[4], int band_rows, int band_rank, int band_world,
                               float* table, size_t table_cap_floats, int* entries_per_row) {
    if (ctx && ctx->multi) ctx = group_root_local(ctx);
    int rc = check_frame_args(ctx, width, height, origin0, dU, dV, rayDir);
    if (rc != RTB_OK) return rc;
    if (band_rows <= 0 || band_rows % 4 != 0 || band_world <= 0 || band_rank < 0 || band_rank >= band_world)
        return fail(ctx, RTB_ERR_INVALID_ARG, "band_rows must be a positive multiple of 4; 0 <= rank < world");
    RTB_CUDA(ctx, cudaSetDevice(ctx->device));
    CameraFrame cf;
    if (!camera_prepare(origin0, dU, dV, rayDir, &cf)) return fail(ctx, RTB_ERR_CAMERA, "degenerate camera");
    if (entries_per_row) *entries_per_row = bound_count(ctx);
    const size_t n = 2 * table_entries(ctx, height);
    if (!table) return RTB_OK;
    if (table_cap_floats < n) return fail(ctx, RTB_ERR_INVALID_ARG, "table buffer too small");
    rc = order_streams(ctx, ctx->stream);
    if (rc != RTB_OK) return rc;
    RTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float cam16[16];
    cam16_of(origin0, dU, dV, rayDir, cam16);
    rc = enqueue_pack(ctx, cf, width, height, cam16, &ctx->cur, ctx->stream);
    if (rc != RTB_OK) return rc;
    rc = ensure_frame_set(ctx, &ctx->cur, frame_records(ctx) + 32, table_entries(ctx, height));
    if (rc != RTB_OK) return rc;
    RTB_CUDA(ctx, cudaMemsetAsync(ctx->cur.tab, 0, n * sizeof(float), ctx->stream));  // rows of other ranks' bands read as 0
    rc = enqueue_table(ctx, cf, &ctx->cur, band_rows, band_rank, band_world, ctx->stream);
    if (rc != RTB_OK) return rc;
    RTB_CUDA(ctx, cudaMemcpyAsync(table, ctx->cur.tab, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    RTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return RTB_OK;
}

int rtb_debug_grid_shape(int width, int rows, int sm_count, int out[4]) {
    if (width <= 0 || rows <= 0 || sm_count <= 0 || !out) return RTB_ERR_INVALID_ARG;
    const GridLaunchShape s = grid_launch_shape(width, rows, sm_count);
    out[0] = s.px;
    out[1] = s.wx * 8 * s.px;
    out[2] = s.wy * 4;
    out[3] = (int)(s.grid.x * s.grid.y);
    return RTB_OK;
}

}  // extern "C"
