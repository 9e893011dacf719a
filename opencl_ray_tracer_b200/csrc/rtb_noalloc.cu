// rtb_noalloc.cu -- rtb_render_device_noalloc (include/rtb.h): one frame, device pointers in and out, and NO
// allocation of any kind inside the call.  For callers that cannot call cudaMalloc: code under CUDA graph capture,
// tight serving loops, harnesses that forbid it.  It has no counterpart in the
// reference.
//
// The binned render (bin_tiles_kernel + raycast_binned_kernel, raycast_kernels.cu) needs scratch memory: the boxes
// recovered from the cubes' vertices, one frame's exact + bound records, and the per-tile primitive bit words.  A
// context (rtb_init) keeps them in grow-only cudaMalloc'ed buffers; here they live in ONE file-scope
// `static __device__` array of a fixed size, which every GPU that loads librtb.so holds until the process exits.
//
// In host code the array's name is a host-side stand-in, not a device address: the address a kernel can use is
// fetched once per device, with that device current, by cudaGetSymbolAddress (scratch_for_device) and only that
// pointer is ever handed to a launch.
//
// Every call on a device shares the array, which keeps the last call's contents.  What must start from a known
// value is set inside the call's own stream order (the bad-cube word: cudaMemsetAsync in launch_boxes_from_vertices;
// every record, the gap between the exact and the bound records, and every bin word the render reads are rewritten
// by the call's own kernels), so two calls in a row on ONE stream are independent of each other.  Calls that overlap
// on the GPU (different streams of one device, no event between them) race on the array.
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <mutex>

#include "../../include/rtb.h"
#include "raycast_kernels.cuh"
#include "scene_pack.h"
#include "scene_pack_device.cuh"

namespace {

using rtb::Box;

constexpr size_t align256(size_t n) { return (n + 255) & ~(size_t)255; }

// Layout of the scratch array (bytes; every section 256-byte aligned).
constexpr size_t kMaxPrims = RTB_NOALLOC_MAX_PRIMITIVES;
constexpr size_t kMaxBinWords = RTB_NOALLOC_MAX_BIN_WORDS;
constexpr size_t kOffBad = 0;                                            // int: lowest index of a cube that is not a box
constexpr size_t kOffBoxes = 256;                                        // Box[kMaxPrims]
constexpr size_t kOffRec = kOffBoxes + align256(sizeof(Box) * kMaxPrims);
// records: 3 per cube + 1 per sphere (<= 3 kMaxPrims), rounded up to 8, + 2 per bound (padded to kBoundGroup)
constexpr size_t kMaxRecFloat4 = ((3 * kMaxPrims + 7) & ~(size_t)7) + 2 * (kMaxPrims + rtb::kBoundGroup) + 32;
constexpr size_t kOffBins = kOffRec + align256(sizeof(float4) * kMaxRecFloat4);
constexpr size_t kScratchBytes = kOffBins + align256(sizeof(uint32_t) * kMaxBinWords);
static_assert(kScratchBytes <= ((size_t)1 << 20), "the static scratch array is meant to stay modest (<= 1 MiB)");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
unsigned char* g_scratch_ptr[kMaxDevices] = {nullptr};
int g_sm_count[kMaxDevices] = {0};

int fail(char* err, size_t cap, int status, const char* fmt, ...) {
    if (err && cap) {
        va_list ap;
        va_start(ap, fmt);
        std::vsnprintf(err, cap, fmt, ap);
        va_end(ap);
    }
    return status;
}

// Device address of the scratch array on `dev`, which must be the current device.  Fetched once per device.
int scratch_for_device(int dev, unsigned char** out, int* sm_count, char* err, size_t cap) {
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_ptr[dev]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || p == nullptr)
            return fail(err, cap, RTB_ERR_CUDA, "cudaGetSymbolAddress(scratch) on device %d: %s", dev, cudaGetErrorString(e));
        int sms = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess)
            return fail(err, cap, RTB_ERR_CUDA, "cudaDeviceGetAttribute on device %d: %s", dev, cudaGetErrorString(e));
        g_sm_count[dev] = sms;
        g_scratch_ptr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_scratch_ptr[dev];
    *sm_count = g_sm_count[dev];
    return RTB_OK;
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace

extern "C" int rtb_render_device_noalloc(int device, int width, int height, const float origin0[4], const float dU[4],
                                         const float dV[4], const float rayDir[4], int numSpheres,
                                         const float* d_sphereOrigins, const float* d_sphereRadius,
                                         const float* d_sphereColours, int numCubes, const float* d_cubeVertices,
                                         const float* d_cubeColours, void* d_rgba8, void* d_prim_id, void* d_rgba_i32,
                                         int* d_bad_cube, void* cuda_stream, char* err, size_t err_cap) {
    using namespace rtb;
    if (err && err_cap) err[0] = 0;
    if (width <= 0 || height <= 0 || !origin0 || !dU || !dV || !rayDir)
        return fail(err, err_cap, RTB_ERR_INVALID_ARG, "bad frame size or null camera vector");
    if ((long long)width * height > (1LL << 31) - 1) return fail(err, err_cap, RTB_ERR_INVALID_ARG, "frame too large");
    if (numSpheres < 0 || numCubes < 0 || (numSpheres > 0 && (!d_sphereOrigins || !d_sphereRadius || !d_sphereColours)) ||
        (numCubes > 0 && (!d_cubeVertices || !d_cubeColours)))
        return fail(err, err_cap, RTB_ERR_INVALID_ARG, "negative primitive count or null scene array");
    if ((size_t)numSpheres + (size_t)numCubes > kMaxPrims)
        return fail(err, err_cap, RTB_ERR_INVALID_ARG,
                    "%d spheres + %d cubes exceed RTB_NOALLOC_MAX_PRIMITIVES = %d (the static scratch array is fixed; "
                    "use a context: rtb_init / rtb_set_scene / rtb_render_bands_device)",
                    numSpheres, numCubes, (int)kMaxPrims);
    const int n_bound = (int)(((size_t)numCubes + (size_t)numSpheres + kBoundGroup - 1) / kBoundGroup * kBoundGroup);
    const size_t words = grid_bin_words(width, height, n_bound);
    if (words > kMaxBinWords)
        return fail(err, err_cap, RTB_ERR_INVALID_ARG,
                    "a %d x %d frame of %d primitives needs %zu tile words, more than RTB_NOALLOC_MAX_BIN_WORDS = %zu "
                    "(64 x 32-pixel tiles x ceil(primitives / 32)); use a context",
                    width, height, numCubes + numSpheres, words, kMaxBinWords);
    if (!aligned16(d_sphereOrigins) || !aligned16(d_sphereColours) || !aligned16(d_cubeVertices) || !aligned16(d_cubeColours) ||
        (reinterpret_cast<uintptr_t>(d_sphereRadius) & 3) || (reinterpret_cast<uintptr_t>(d_rgba8) & 3) ||
        (reinterpret_cast<uintptr_t>(d_prim_id) & 3) || !aligned16(d_rgba_i32) || (reinterpret_cast<uintptr_t>(d_bad_cube) & 3))
        return fail(err, err_cap, RTB_ERR_INVALID_ARG, "a device pointer is not aligned for its element type (float4 arrays: 16 bytes)");

    cudaError_t e;
    int dev = device;
    if (dev < 0) {
        e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return fail(err, err_cap, RTB_ERR_NO_DEVICE, "cudaGetDevice: %s", cudaGetErrorString(e));
    } else {
        e = cudaSetDevice(dev);
        if (e != cudaSuccess) return fail(err, err_cap, RTB_ERR_NO_DEVICE, "cudaSetDevice(%d): %s", dev, cudaGetErrorString(e));
    }
    if (dev >= kMaxDevices) return fail(err, err_cap, RTB_ERR_NO_DEVICE, "device %d out of range", dev);

    CameraFrame cf;
    if (!camera_prepare(origin0, dU, dV, rayDir, &cf))
        return fail(err, err_cap, RTB_ERR_CAMERA, "degenerate camera (dU, dV, rayDir do not span a ray grid)");

    unsigned char* scratch = nullptr;
    int sm_count = 0;
    const int rc = scratch_for_device(dev, &scratch, &sm_count, err, err_cap);
    if (rc != RTB_OK) return rc;
    int* const s_bad = reinterpret_cast<int*>(scratch + kOffBad);
    Box* const s_boxes = reinterpret_cast<Box*>(scratch + kOffBoxes);
    float4* const s_rec = reinterpret_cast<float4*>(scratch + kOffRec);
    uint32_t* const s_bins = reinterpret_cast<uint32_t*>(scratch + kOffBins);

    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);  // NULL: the default stream, and the call waits for it
#define NA_CUDA(call)                                                                                      \
    do {                                                                                                   \
        const cudaError_t e__ = (call);                                                                    \
        if (e__ != cudaSuccess) return fail(err, err_cap, RTB_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e__)); \
    } while (0)

    // 1. boxes from the 36 vertices of every cube (the bad-cube word is reset first, in stream order)
    NA_CUDA(launch_boxes_from_vertices(d_cubeVertices, numCubes, s_boxes, s_bad, stream));
    if (d_bad_cube) NA_CUDA(cudaMemcpyAsync(d_bad_cube, s_bad, sizeof(int), cudaMemcpyDeviceToDevice, stream));
    // 2. this frame's exact + bound records
    const size_t exact = 3 * (size_t)numCubes + (size_t)numSpheres;
    const size_t bound_off = (exact + 7) & ~(size_t)7;
    PackFrameParams pp;
    pp.cf = cf;
    pp.fe = frame_extent(cf, width, height);
    pp.nc = numCubes; pp.ns = numSpheres; pp.n_bound = n_bound;
    pp.boxes = s_boxes;
    pp.sph_o = reinterpret_cast<const float4*>(d_sphereOrigins);
    pp.sph_r = d_sphereRadius;
    pp.rec = s_rec;
    pp.bound_off = bound_off;
    NA_CUDA(launch_pack_frame(pp, stream));
    // 3. bin + render (tile_cull = 2: byte-identical to the dense kernel)
    GridParams p;
    std::memset(&p, 0, sizeof p);
    p.width = width; p.height = height;
    p.band_rows = 16; p.band_rank = 0; p.band_world = 1;
    p.local_rows = height;
    p.out_global_rows = 0;
    p.n_cubes = numCubes; p.n_spheres = numSpheres;
    p.cube_rec = s_rec;
    p.sph_rec = s_rec + 3 * (size_t)numCubes;
    p.bound_rec = s_rec + bound_off;
    p.n_bound = n_bound;
    p.row_tab = nullptr;
    p.n_pad = n_bound;
    p.cube_col = reinterpret_cast<const float4*>(d_cubeColours);
    p.sph_col = reinterpret_cast<const float4*>(d_sphereColours);
    p.out_rgba8 = static_cast<uint32_t*>(d_rgba8);
    p.out_id = static_cast<int32_t*>(d_prim_id);
    p.out_i32 = static_cast<int4*>(d_rgba_i32);
    p.vec_ok = (width % 4 == 0) && aligned16(d_rgba8) && aligned16(d_prim_id) && aligned16(d_rgba_i32);
    p.progress = nullptr;
    p.done_counter = nullptr;
    p.done_signal = nullptr;
    p.tile_cull = 2;
    p.bins = s_bins;
    p.bin_words = (n_bound + 31) / 32;
    p.cam = cf.cam;
    NA_CUDA(launch_grid(p, sm_count, stream));
    if (!cuda_stream) NA_CUDA(cudaStreamSynchronize(stream));
#undef NA_CUDA
    return RTB_OK;
}
