// rtb_group.cu -- multi-GPU frames behind the C ABI (include/rtb.h: rtb_init(n_devices > 1), rtb_group_*).
//
// The reference renders a frame with ONE blocking call on ONE device (MainState::executeRayTracerOpenCL,
// RayTrace/states/MainState.cpp:217-224, :641-934).  Here a frame is sharded over the GPUs of one box as
// interleaved row bands and is still one call per frame and rank:
//
//   local group   rtb_init(n_devices > 1): one process drives every device (a worker thread per device);
//                 rtb_set_scene / rtb_render / rtb_render_sequence_device of the returned context fan out.
//   process group rtb_group_join: one process per GPU (how bench.py runs under torchrun).  The ranks meet in a
//                 POSIX shared-memory control block (no NCCL, no Python in the frame loop):
//                   * rank 0 owns a ring of frame slots in its device memory, mapped into every rank (CUDA IPC);
//                     each rank's dense kernel stores its bands into the slot over NVLink, and the launch's last
//                     CTA bumps THAT RANK's done word in rank 0's memory (one word per rank: a rank that runs
//                     ahead can never stand in for one that lags);
//                   * rank 0's stream waits on the done words (cuStreamWaitValue32) -- no collective;
//                   * a slot is handed back by credit words: one per rank, in that rank's own memory, written by a
//                     tiny kernel of rank 0 through the peer mapping when rank 0 starts its next frame; a rank's
//                     stream waits on its own credit word before it overwrites a slot.
//                   * host frames (rtb_group_render_host): one frame in shared, page-locked host memory; every GPU
//                     copies its bands over its own PCIe link; completion and release are counters in the control
//                     block that the ranks' host threads poll.
//   Waits are stream memory operations or host polls; no kernel ever spins on another rank.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <new>
#include <thread>

#include "rtb_group.h"

using namespace rtb;

namespace {

constexpr uint32_t kShmMagic = 0x52544247u;  // "RTBG"
constexpr int kDoneStrideWords = 8;          // done words sit 32 bytes apart (one sector each)
constexpr double kPeerTimeoutS = 60.0;

struct GroupShm {
    std::atomic<uint32_t> magic;
    uint32_t world, width, height, slots, band_rows;
    std::atomic<uint32_t> joined;                         // ranks that have mapped everything
    std::atomic<uint32_t> left;                           // ranks that have left
    std::atomic<uint32_t> frames_ready;
    std::atomic<uint32_t> mailbox_ready[RTB_MAX_GROUP];
    std::atomic<uint32_t> error;                          // a rank failed: everybody stops waiting
    std::atomic<uint64_t> barrier_count;                  // rtb_group_barrier: monotonic arrivals
    std::atomic<uint64_t> host_done[RTB_MAX_GROUP];       // host frames delivered by rank r
    std::atomic<uint64_t> host_released;                  // host frames rank 0 is done with
    unsigned char frames_handle[64];
    unsigned char mailbox_handle[RTB_MAX_GROUP][64];
};

// Persistent worker threads of a local group: run(fn) calls fn(r) for r = 1..n-1 on the workers and fn(0) inline.
class Workers {
public:
    explicit Workers(int n) : n_(n) {
        for (int r = 1; r < n; ++r) threads_.emplace_back([this, r] { loop(r); });
    }
    ~Workers() {
        {
            std::lock_guard<std::mutex> lk(m_);
            stop_ = true;
            ++gen_;
        }
        cv_.notify_all();
        for (auto& t : threads_) t.join();
    }
    void run(const std::function<void(int)>& fn) {
        {
            std::lock_guard<std::mutex> lk(m_);
            fn_ = &fn;
            pending_ = n_ - 1;
            ++gen_;
        }
        cv_.notify_all();
        fn(0);
        std::unique_lock<std::mutex> lk(m_);
        done_cv_.wait(lk, [this] { return pending_ == 0; });
        fn_ = nullptr;
    }

private:
    void loop(int r) {
        unsigned long long seen = 0;
        for (;;) {
            const std::function<void(int)>* fn;
            {
                std::unique_lock<std::mutex> lk(m_);
                cv_.wait(lk, [&] { return gen_ != seen; });
                seen = gen_;
                if (stop_) return;
                fn = fn_;
            }
            (*fn)(r);
            {
                std::lock_guard<std::mutex> lk(m_);
                --pending_;
            }
            done_cv_.notify_one();
        }
    }
    int n_;
    std::vector<std::thread> threads_;
    std::mutex m_;
    std::condition_variable cv_, done_cv_;
    const std::function<void(int)>* fn_ = nullptr;
    unsigned long long gen_ = 0;
    int pending_ = 0;
    bool stop_ = false;
};

__global__ void release_credits_kernel(unsigned int* const* words, int n, unsigned int value) {
    const int r = threadIdx.x;
    if (r < n && words[r] != nullptr) {
        __threadfence_system();
        *reinterpret_cast<volatile unsigned int*>(words[r]) = value;  // this rank is the only writer; values only grow
    }
}

double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

}  // namespace

struct rtb_group {
    bool local = false;            // every rank is a device of this process
    int world = 1, rank = 0;       // process group: this process's rank
    int width = 0, height = 0, slots = 0, band_rows = 16;
    size_t frame_bytes = 0;        // one slot
    size_t done_off = 0;           // byte offset of the done words behind the slots
    rtb_ctx* ctx[RTB_MAX_GROUP] = {nullptr};  // local: every rank's context; process: only [rank]
    uint8_t* frames[RTB_MAX_GROUP] = {nullptr};        // rank 0's slot ring as rank r addresses it
    unsigned int* credit_own[RTB_MAX_GROUP] = {nullptr};   // rank r's credit word (r's memory, r's address)
    unsigned int* credit_at0[RTB_MAX_GROUP] = {nullptr};   // the same words as rank 0 addresses them
    unsigned int** d_credit_ptrs = nullptr;                // device copy of credit_at0 (rank 0's device)
    unsigned long long frame_no[RTB_MAX_GROUP] = {0};      // frames enqueued so far by rank r
    unsigned long long host_frame_no = 0;
    cudaStream_t release_stream = nullptr;                 // rank 0: credits are written off the frame's stream
    cudaEvent_t ev_release = nullptr;
    // process group
    std::string shm_name, host_name;
    GroupShm* shm = nullptr;
    uint8_t* host_frame = nullptr;
    size_t host_frame_bytes = 0;
    bool host_registered = false;
    // local group
    Workers* workers = nullptr;
    std::string last_error;
};

namespace {

unsigned int* done_word(const rtb_group* g, int viewer_rank, int r) {
    return reinterpret_cast<unsigned int*>(g->frames[viewer_rank] + g->done_off) + (size_t)r * kDoneStrideWords;
}

int gfail(rtb_group* g, rtb_ctx* ctx, int status, const std::string& msg) {
    if (g) g->last_error = msg;
    if (g && g->shm) g->shm->error.store(1u);
    return fail(ctx, status, msg);
}

// Polls until pred() or a peer reported an error / the timeout passed.
template <class Pred>
bool poll_until(const rtb_group* g, Pred pred) {
    const double t0 = now_s();
    for (unsigned spin = 0;; ++spin) {
        if (pred()) return true;
        if (g->shm && g->shm->error.load(std::memory_order_relaxed)) return false;
        if ((spin & 1023u) == 1023u) {
            if (now_s() - t0 > kPeerTimeoutS) return false;
            std::this_thread::yield();
        }
    }
}

// One rank's share of frame f = g->frame_no[r] into slot f % slots, enqueued on `stream` (see the header comment).
int group_enqueue_rank(rtb_group* g, int r, const float* cam16, const CameraFrame& cf, cudaStream_t stream, void** d_frame) {
    rtb_ctx* ctx = g->ctx[r];
    RTB_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!ctx->stream_wait_value32) return fail(ctx, RTB_ERR_UNSUPPORTED, "stream memory operations are unavailable");
    int rc = check_frame_args(ctx, g->width, g->height, cam16, cam16 + 4, cam16 + 8, cam16 + 12);
    if (rc != RTB_OK) return rc;
    const unsigned long long f = g->frame_no[r];
    const int slot = (int)(f % (unsigned long long)g->slots);
    uint8_t* frame = g->frames[r] + (size_t)slot * g->frame_bytes;
    rc = order_streams(ctx, stream);
    if (rc != RTB_OK) return rc;
    if (r == 0) {
        // rank 0 is done with frame f-1 (whatever the caller enqueued on this stream since the last call has run
        // by the time this kernel does): hand its slot back to the ranks
        // -- on a side stream behind an event of this one: the frame's stream does not spend a launch on it
        if (f > 0 && g->world > 1) {
            if (!g->release_stream) {
                int lo = 0, hi = 0;
                RTB_CUDA(ctx, cudaDeviceGetStreamPriorityRange(&lo, &hi));
                RTB_CUDA(ctx, cudaStreamCreateWithPriority(&g->release_stream, cudaStreamNonBlocking, hi));
                RTB_CUDA(ctx, cudaEventCreateWithFlags(&g->ev_release, cudaEventDisableTiming));
            }
            RTB_CUDA(ctx, cudaEventRecord(g->ev_release, stream));
            RTB_CUDA(ctx, cudaStreamWaitEvent(g->release_stream, g->ev_release, 0));
            release_credits_kernel<<<1, 32, 0, g->release_stream>>>(g->d_credit_ptrs, g->world, (unsigned int)f);
            RTB_CUDA(ctx, cudaGetLastError());
            ctx->launches += 1;
        }
    } else if (f >= (unsigned long long)g->slots) {
        // frame f - slots lived in this slot: wait until rank 0 has released it (credit = frames released so far)
        const unsigned int need = (unsigned int)(f - (unsigned long long)g->slots + 1ull);
        if (ctx->stream_wait_value32(stream, reinterpret_cast<CUdeviceptr>(g->credit_own[r]), need, CU_STREAM_WAIT_VALUE_GEQ) !=
            CUDA_SUCCESS)
            return fail(ctx, RTB_ERR_CUDA, "cuStreamWaitValue32 (credit) failed");
    }
    // "this rank's bands of frame f have landed": a stream write of the rank's done word (in rank 0's memory) behind the
    // launch -- the word is written once the kernel and its stores have completed, and the launch carries no fences.
    // RTB_GROUP_STREAM_SIGNAL=0 (or no stream memory operations): the launch's last CTA bumps the word itself, behind a
    // system-scope fence per CTA.  Measured on C3 at 2 GPUs: rank 1's kernel 0.3048 ms against 0.3183 ms with the
    // in-kernel signal (binned frame: 0.0405 against 0.0511 ms); the step, which pays ~4 us for the stream write,
    // 0.3181 against 0.3275 ms (profiles/r02_ab_signal_n2.txt).
    static const bool stream_signal = [] { const char* e = std::getenv("RTB_GROUP_STREAM_SIGNAL"); return !(e && e[0] == '0'); }();
    const bool by_stream = stream_signal && ctx->stream_write_value32 != nullptr && r != 0;
    rc = enqueue_bands_frame(ctx, g->width, g->height, cam16, cf, g->band_rows, r, g->world, frame,
                             (r == 0 || by_stream) ? nullptr : done_word(g, r, r), stream);
    if (rc != RTB_OK) return rc;
    if (by_stream && ctx->stream_write_value32(stream, reinterpret_cast<CUdeviceptr>(done_word(g, r, r)), (unsigned int)(f + 1ull),
                                               CU_STREAM_WRITE_VALUE_DEFAULT) != CUDA_SUCCESS)
        return fail(ctx, RTB_ERR_CUDA, "cuStreamWriteValue32 (done) failed");
    if (r == 0) {
        // every other rank's bands of frame f have landed in rank 0's memory: one wait per rank's done word, submitted
        // as ONE batch (a stream wait costs 2-4 us of the stream's time each when issued on its own: 7 of them were
        // 26 us of an 8-GPU frame's 123 us); RTB_GROUP_BATCH_WAIT=0 or no cuStreamBatchMemOp: one call per word
        static const bool batch = [] { const char* e = std::getenv("RTB_GROUP_BATCH_WAIT"); return !(e && e[0] == '0'); }();
        bool done = false;
        if (batch && ctx->stream_batch_memop && g->world > 2) {
            CUstreamBatchMemOpParams ops[RTB_MAX_GROUP];
            std::memset(ops, 0, sizeof ops);
            for (int q = 1; q < g->world; ++q) {
                ops[q - 1].waitValue.operation = CU_STREAM_MEM_OP_WAIT_VALUE_32;
                ops[q - 1].waitValue.address = reinterpret_cast<CUdeviceptr>(done_word(g, 0, q));
                ops[q - 1].waitValue.value = (cuuint32_t)(f + 1ull);
                ops[q - 1].waitValue.flags = CU_STREAM_WAIT_VALUE_GEQ;
            }
            done = ctx->stream_batch_memop(stream, (unsigned int)(g->world - 1), ops, 0) == CUDA_SUCCESS;
        }
        for (int q = 1; q < g->world && !done; ++q)
            if (ctx->stream_wait_value32(stream, reinterpret_cast<CUdeviceptr>(done_word(g, 0, q)), (unsigned int)(f + 1ull),
                                         CU_STREAM_WAIT_VALUE_GEQ) != CUDA_SUCCESS)
                return fail(ctx, RTB_ERR_CUDA, "cuStreamWaitValue32 (done) failed");
        if (d_frame) *d_frame = frame;
    }
    g->frame_no[r] = f + 1;
    return RTB_OK;
}

}  // namespace

// ---- local group: rtb_init(n_devices > 1) -----------------------------------------------------------
namespace rtb {

int group_init_local(rtb_ctx** out, int n, const int* device_ids) {
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) return RTB_ERR_NO_DEVICE;
    int ids[RTB_MAX_GROUP];
    for (int r = 0; r < n; ++r) {
        ids[r] = device_ids ? device_ids[r] : r;
        if (ids[r] < 0 || ids[r] >= count) return RTB_ERR_NO_DEVICE;
        for (int q = 0; q < r; ++q)
            if (ids[q] == ids[r]) return RTB_ERR_INVALID_ARG;  // one context per device
    }
    rtb_group* g = new (std::nothrow) rtb_group();
    rtb_ctx* root = new (std::nothrow) rtb_ctx();
    if (!g || !root) { delete g; delete root; return RTB_ERR_OUT_OF_MEMORY; }
    g->local = true;
    g->world = n;
    g->rank = 0;
    int rc = RTB_OK;
    for (int r = 0; r < n && rc == RTB_OK; ++r) rc = init_single(&g->ctx[r], ids[r]);
    // every device stores into device 0's frames and done words; device 0 stores the credits into the others
    for (int r = 1; r < n && rc == RTB_OK; ++r) {
        int can01 = 0, can10 = 0;
        cudaDeviceCanAccessPeer(&can10, ids[r], ids[0]);
        cudaDeviceCanAccessPeer(&can01, ids[0], ids[r]);
        if (!can01 || !can10) { rc = RTB_ERR_UNSUPPORTED; break; }
        cudaSetDevice(ids[r]);
        cudaError_t e = cudaDeviceEnablePeerAccess(ids[0], 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) rc = RTB_ERR_CUDA;
        cudaSetDevice(ids[0]);
        e = cudaDeviceEnablePeerAccess(ids[r], 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) rc = RTB_ERR_CUDA;
        (void)cudaGetLastError();
    }
    if (rc != RTB_OK) {
        for (int r = 0; r < n; ++r)
            if (g->ctx[r]) rtb_shutdown(g->ctx[r]);
        delete g;
        delete root;
        return rc;
    }
    g->workers = new Workers(n);
    // the front context: no device state of its own; device info = the first device's
    root->device = g->ctx[0]->device;
    root->sm_count = g->ctx[0]->sm_count;
    root->clock_khz = g->ctx[0]->clock_khz;
    std::snprintf(root->name, sizeof root->name, "%.100s x%d", g->ctx[0]->name, n);
    root->multi = g;
    *out = root;
    return RTB_OK;
}

void group_shutdown_local(rtb_ctx* root) {
    rtb_group* g = root->multi;
    if (!g) return;
    delete g->workers;
    g->workers = nullptr;
    for (int r = 0; r < g->world; ++r)
        if (g->ctx[r]) rtb_shutdown(g->ctx[r]);
    delete g;
    root->multi = nullptr;
}

rtb_ctx* group_root_local(rtb_ctx* root) { return root->multi->ctx[0]; }

int group_set_scene_local(rtb_ctx* root, int numSpheres, const float* sphereOrigins, const float* sphereRadius,
                          const float* sphereColours, int numCubes, const float* cubeVertices, const float* cubeColours) {
    rtb_group* g = root->multi;
    int rcs[RTB_MAX_GROUP] = {0};
    g->workers->run([&](int r) {
        rcs[r] = set_scene_single(g->ctx[r], numSpheres, sphereOrigins, sphereRadius, sphereColours, numCubes, cubeVertices,
                                  cubeColours);
    });
    root->have_scene = false;
    for (int r = 0; r < g->world; ++r)
        if (rcs[r] != RTB_OK) return fail(root, rcs[r], g->ctx[r]->last_error);
    root->have_scene = true;
    root->ns = numSpheres;
    root->nc = numCubes;
    return RTB_OK;
}

// One blocking call per frame, like the reference's: every device renders its interleaved 16-row bands and copies
// them over its own PCIe link straight into the caller's host buffers (finished chunks first, while its launch runs).
int group_render_local(rtb_ctx* root, int width, int height, const float origin0[4], const float dU[4], const float dV[4],
                       const float rayDir[4], uint8_t* rgba8_out, int32_t* prim_id_out, int32_t* rgba_i32_out,
                       float* kernel_ms, float* total_ms) {
    rtb_group* g = root->multi;
    if (width <= 0 || height <= 0 || !origin0 || !dU || !dV || !rayDir)
        return fail(root, RTB_ERR_INVALID_ARG, "bad frame size or null camera vector");
    if (!root->have_scene) return fail(root, RTB_ERR_NO_SCENE, "rtb_set_scene has not been called");
    CameraFrame cf;
    if (!camera_prepare(origin0, dU, dV, rayDir, &cf))
        return fail(root, RTB_ERR_CAMERA, "degenerate camera (dU, dV, rayDir do not span a ray grid)");
    float cam16[16];
    std::memcpy(cam16, origin0, 16); std::memcpy(cam16 + 4, dU, 16);
    std::memcpy(cam16 + 8, dV, 16); std::memcpy(cam16 + 12, rayDir, 16);
    void* const host[3] = {rgba8_out, prim_id_out, rgba_i32_out};
    int rcs[RTB_MAX_GROUP] = {0};
    float kms[RTB_MAX_GROUP] = {0.f};
    const double t0 = now_s();
    g->workers->run([&](int r) {
        rtb_ctx* c = g->ctx[r];
        cudaSetDevice(c->device);
        rcs[r] = check_frame_args(c, width, height, origin0, dU, dV, rayDir);
        if (rcs[r] == RTB_OK) rcs[r] = bands_host_render(c, width, height, cam16, cf, 16, r, g->world, host, &kms[r]);
        if (rcs[r] != RTB_OK) drain(c, nullptr);
    });
    const double t1 = now_s();
    float kmax = 0.f;
    for (int r = 0; r < g->world; ++r) {
        if (rcs[r] != RTB_OK) return fail(root, rcs[r], g->ctx[r]->last_error);
        kmax = kms[r] > kmax ? kms[r] : kmax;
    }
    if (kernel_ms) *kernel_ms = kmax;                      // the slowest device's launch
    if (total_ms) *total_ms = (float)((t1 - t0) * 1e3);    // host clock around the fan-out
    return RTB_OK;
}

// n_frames frames kept on the first device (d_rgba8 = its memory): every device stores its bands of every frame
// straight into that buffer over NVLink (peer access); the frames are distinct, so no slot hand-over is needed.
int group_render_sequence_local(rtb_ctx* root, int width, int height, int n_frames, const float* cameras, void* d_rgba8,
                                float* kernel_ms_total) {
    rtb_group* g = root->multi;
    if (n_frames < 0 || (n_frames > 0 && (!cameras || !d_rgba8))) return fail(root, RTB_ERR_INVALID_ARG, "null cameras / output");
    if (kernel_ms_total) *kernel_ms_total = 0.f;
    if (n_frames == 0) return RTB_OK;
    if (!root->have_scene) return fail(root, RTB_ERR_NO_SCENE, "rtb_set_scene has not been called");
    // Culled modes do not shard for frames that stay on a device: the first device renders a binned C3 frame in 0.033 ms
    // (33 000 frames/s on two lanes), while moving (N-1)/N of a frame into its memory takes 0.038 ms at NVLink speed --
    // sharded over 2 / 4 / 8 devices the sequence ran at 18 700 / 18 600 / 17 900 frames/s
    // (profiles/r02_local_group_time.jsonl).  The first device renders such sequences alone.
    if (effective_cull(g->ctx[0]) != 0) {
        const int rc = rtb_render_sequence_device(g->ctx[0], width, height, n_frames, cameras, d_rgba8, kernel_ms_total);
        return rc == RTB_OK ? RTB_OK : fail(root, rc, g->ctx[0]->last_error);
    }
    const size_t frame_bytes = 4 * (size_t)width * height;
    int rcs[RTB_MAX_GROUP] = {0};
    double ksum[RTB_MAX_GROUP] = {0.0};
    g->workers->run([&](int r) {
        rtb_ctx* c = g->ctx[r];
        cudaSetDevice(c->device);
        int rc = check_frame_args(c, width, height, cameras, cameras + 4, cameras + 8, cameras + 12);
        if (rc == RTB_OK) rc = resolve_pending(c);
        const double k0 = c->kernel_ms_sum;
        cudaStream_t st = c->stream;
        for (int f0 = 0; f0 < n_frames && rc == RTB_OK;) {
            rc = order_streams(c, st);
            if (rc == RTB_OK) rc = prepack_cameras(c, width, height, 16, r, g->world, n_frames - f0, cameras + 16 * (size_t)f0, st);
            const int batch = c->slots_used;
            for (int s = 0; s < batch && rc == RTB_OK; ++s) {
                const float* cam = cameras + 16 * (size_t)(f0 + s);
                CameraFrame cf;
                camera_prepare(cam, cam + 4, cam + 8, cam + 12, &cf);  // checked by prepack_cameras
                rc = enqueue_bands_frame(c, width, height, cam, cf, 16, r, g->world,
                                         static_cast<uint8_t*>(d_rgba8) + frame_bytes * (size_t)(f0 + s), nullptr, st);
            }
            if (batch <= 0 && rc == RTB_OK) rc = fail(c, RTB_ERR_CUDA, "no frame slot could be packed");
            f0 += batch;
        }
        c->slots_used = 0;
        if (rc == RTB_OK && cudaStreamSynchronize(st) != cudaSuccess) rc = fail(c, RTB_ERR_CUDA, "cudaStreamSynchronize");
        if (rc == RTB_OK) rc = resolve_pending(c);
        if (rc == RTB_OK) rc = collect_scene_verdict(c);
        if (rc != RTB_OK) { cudaDeviceSynchronize(); drain(c, nullptr); }
        ksum[r] = c->kernel_ms_sum - k0;
        rcs[r] = rc;
    });
    double kmax = 0.0;
    for (int r = 0; r < g->world; ++r) {
        if (rcs[r] != RTB_OK) return fail(root, rcs[r], g->ctx[r]->last_error);
        kmax = ksum[r] > kmax ? ksum[r] : kmax;
    }
    if (kernel_ms_total) *kernel_ms_total = (float)kmax;
    return RTB_OK;
}

void group_set_option_local(rtb_ctx* root, int option, int value) {
    rtb_group* g = root->multi;
    for (int r = 0; r < g->world; ++r) rtb_set_option(g->ctx[r], option, value);
}

unsigned long long group_launch_count_local(const rtb_ctx* root) {
    unsigned long long n = 0;
    for (int r = 0; r < root->multi->world; ++r) n += root->multi->ctx[r]->launches;
    return n;
}

void group_transfer_bytes_local(const rtb_ctx* root, unsigned long long* h2d, unsigned long long* d2h) {
    unsigned long long a = 0, b = 0;
    for (int r = 0; r < root->multi->world; ++r) { a += root->multi->ctx[r]->h2d_bytes; b += root->multi->ctx[r]->d2h_bytes; }
    if (h2d) *h2d = a;
    if (d2h) *d2h = b;
}

}  // namespace rtb

// ---- process group: one process per GPU ------------------------------------------------------------
namespace {

void* map_shm(const std::string& name, size_t bytes, bool create) {
    int fd = -1;
    if (create) {
        shm_unlink(name.c_str());  // a stale segment of a crashed job
        fd = shm_open(name.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
        if (fd >= 0 && ftruncate(fd, (off_t)bytes) != 0) { close(fd); shm_unlink(name.c_str()); return nullptr; }
    } else {
        const double t0 = now_s();
        while (fd < 0) {
            fd = shm_open(name.c_str(), O_RDWR, 0600);
            if (fd >= 0) {
                struct stat st;
                if (fstat(fd, &st) == 0 && (size_t)st.st_size >= bytes) break;
                close(fd);  // created but not yet sized
                fd = -1;
            }
            if (now_s() - t0 > kPeerTimeoutS) return nullptr;
            usleep(200);
        }
    }
    if (fd < 0) return nullptr;
    void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    return p == MAP_FAILED ? nullptr : p;
}

}  // namespace

extern "C" {

int rtb_group_join(rtb_ctx* ctx, const char* name, int rank, int world, int width, int height, int band_rows, int slots,
                   rtb_group** out) {
    if (!ctx || !out) return RTB_ERR_INVALID_ARG;
    *out = nullptr;
    if (ctx->multi) return fail(ctx, RTB_ERR_UNSUPPORTED, "rtb_group_join takes a single-device context (one process per GPU)");
    if (!name || !name[0] || world < 1 || world > RTB_MAX_GROUP || rank < 0 || rank >= world || width <= 0 || height <= 0 ||
        band_rows <= 0 || band_rows % 4 != 0 || slots < 1 || slots > 64)
        return fail(ctx, RTB_ERR_INVALID_ARG, "bad group arguments (1 <= world <= 16, band_rows a multiple of 4, 1 <= slots <= 64)");
    if (!ctx->stream_wait_value32) return fail(ctx, RTB_ERR_UNSUPPORTED, "stream memory operations are unavailable");
    RTB_CUDA(ctx, cudaSetDevice(ctx->device));
    rtb_group* g = new (std::nothrow) rtb_group();
    if (!g) return RTB_ERR_OUT_OF_MEMORY;
    g->world = world; g->rank = rank;
    g->width = width; g->height = height; g->slots = slots; g->band_rows = band_rows;
    g->ctx[rank] = ctx;
    g->frame_bytes = (4 * (size_t)width * height + 255) & ~(size_t)255;
    g->done_off = g->frame_bytes * (size_t)slots;
    g->shm_name = std::string(name[0] == '/' ? "" : "/") + name;
    g->host_name = g->shm_name + "_frame";
    auto bail = [&](int status, const std::string& msg) {
        if (g->shm) g->shm->error.store(1u);
        rtb_group_leave(g);
        return fail(ctx, status, msg);
    };
    g->shm = static_cast<GroupShm*>(map_shm(g->shm_name, sizeof(GroupShm), rank == 0));
    if (!g->shm) { delete g; return fail(ctx, RTB_ERR_CUDA, "could not map the group's control block " + g->shm_name); }
    GroupShm* s = g->shm;
    if (rank == 0) {
        std::memset(static_cast<void*>(s), 0, sizeof(GroupShm));
        s->world = (uint32_t)world; s->width = (uint32_t)width; s->height = (uint32_t)height;
        s->slots = (uint32_t)slots; s->band_rows = (uint32_t)band_rows;
        s->magic.store(kShmMagic, std::memory_order_release);
    } else {
        if (!poll_until(g, [&] { return s->magic.load(std::memory_order_acquire) == kShmMagic; }))
            return bail(RTB_ERR_CUDA, "group rank 0 did not show up");
        if (s->world != (uint32_t)world || s->width != (uint32_t)width || s->height != (uint32_t)height ||
            s->slots != (uint32_t)slots || s->band_rows != (uint32_t)band_rows)
            return bail(RTB_ERR_INVALID_ARG, "group arguments differ between the ranks");
    }
    // device memory: rank 0 the slot ring + done words, every rank its credit word; CUDA IPC handles through the block
    void* mail = nullptr;
    if (cudaMalloc(&mail, 256) != cudaSuccess || cudaMemset(mail, 0, 256) != cudaSuccess) return bail(RTB_ERR_CUDA, "credit word allocation failed");
    g->credit_own[rank] = static_cast<unsigned int*>(mail);
    {
        cudaIpcMemHandle_t h;
        if (cudaIpcGetMemHandle(&h, mail) != cudaSuccess) return bail(RTB_ERR_CUDA, "cudaIpcGetMemHandle (credit word) failed");
        std::memcpy(s->mailbox_handle[rank], &h, 64);
        s->mailbox_ready[rank].store(1u, std::memory_order_release);
    }
    if (rank == 0) {
        void* p = nullptr;
        cudaError_t e = cudaMalloc(&p, g->done_off + 4096);
        if (e != cudaSuccess) return bail(e == cudaErrorMemoryAllocation ? RTB_ERR_OUT_OF_MEMORY : RTB_ERR_CUDA, "frame slot allocation failed");
        g->frames[0] = static_cast<uint8_t*>(p);
        cudaIpcMemHandle_t h;
        if (cudaMemset(g->frames[0] + g->done_off, 0, 4096) != cudaSuccess || cudaIpcGetMemHandle(&h, p) != cudaSuccess)
            return bail(RTB_ERR_CUDA, "cudaIpcGetMemHandle (frame slots) failed");
        std::memcpy(s->frames_handle, &h, 64);
        s->frames_ready.store(1u, std::memory_order_release);
        unsigned int* ptrs[RTB_MAX_GROUP] = {nullptr};
        for (int r = 1; r < world; ++r) {
            if (!poll_until(g, [&] { return s->mailbox_ready[r].load(std::memory_order_acquire) != 0u; }))
                return bail(RTB_ERR_CUDA, "a group rank did not show up");
            cudaIpcMemHandle_t mh;
            std::memcpy(&mh, s->mailbox_handle[r], 64);
            void* q = nullptr;
            if (cudaIpcOpenMemHandle(&q, mh, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess)
                return bail(RTB_ERR_CUDA, std::string("cudaIpcOpenMemHandle (credit word): ") + cudaGetErrorString(cudaGetLastError()));
            g->credit_at0[r] = ptrs[r] = static_cast<unsigned int*>(q);
        }
        if (cudaMalloc(&g->d_credit_ptrs, sizeof ptrs) != cudaSuccess ||
            cudaMemcpy(g->d_credit_ptrs, ptrs, sizeof ptrs, cudaMemcpyHostToDevice) != cudaSuccess)
            return bail(RTB_ERR_CUDA, "credit pointer table allocation failed");
    } else {
        if (!poll_until(g, [&] { return s->frames_ready.load(std::memory_order_acquire) != 0u; }))
            return bail(RTB_ERR_CUDA, "group rank 0 did not publish its frame slots");
        cudaIpcMemHandle_t h;
        std::memcpy(&h, s->frames_handle, 64);
        void* p = nullptr;
        if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess)
            return bail(RTB_ERR_CUDA, std::string("cudaIpcOpenMemHandle (frame slots): ") + cudaGetErrorString(cudaGetLastError()));
        g->frames[rank] = static_cast<uint8_t*>(p);
    }
    // the shared host frame of rtb_group_render_host, page-locked in every rank
    g->host_frame_bytes = 4 * (size_t)width * height;
    g->host_frame = static_cast<uint8_t*>(map_shm(g->host_name, g->host_frame_bytes, rank == 0));
    if (!g->host_frame) return bail(RTB_ERR_CUDA, "could not map the group's host frame " + g->host_name);
    if (cudaHostRegister(g->host_frame, g->host_frame_bytes, cudaHostRegisterPortable) == cudaSuccess) g->host_registered = true;
    else (void)cudaGetLastError();  // not page-locked: the copies still work (staged by the driver), only slower
    // everybody has mapped everything before anybody renders (or unlinks the names)
    s->joined.fetch_add(1u, std::memory_order_acq_rel);
    if (!poll_until(g, [&] { return s->joined.load(std::memory_order_acquire) >= (uint32_t)world; }))
        return bail(RTB_ERR_CUDA, "group join timed out");
    if (rank == 0) {  // the names are no longer needed: the segments live until the last rank unmaps them
        shm_unlink(g->shm_name.c_str());
        shm_unlink(g->host_name.c_str());
    }
    *out = g;
    return RTB_OK;
}

int rtb_group_leave(rtb_group* g) {
    if (!g) return RTB_ERR_INVALID_ARG;
    if (g->local) return RTB_ERR_INVALID_ARG;  // owned by its context: rtb_shutdown
    rtb_ctx* ctx = g->ctx[g->rank];
    if (ctx) {
        cudaSetDevice(ctx->device);
        cudaDeviceSynchronize();
    }
    GroupShm* s = g->shm;
    if (s) {
        // nobody frees memory a peer may still be storing into: every rank drains its device, then all meet
        s->left.fetch_add(1u, std::memory_order_acq_rel);
        poll_until(g, [&] { return s->left.load(std::memory_order_acquire) >= (uint32_t)g->world; });
    }
    if (g->host_frame) {
        if (g->host_registered) cudaHostUnregister(g->host_frame);
        munmap(g->host_frame, g->host_frame_bytes);
    }
    if (g->release_stream) {
        cudaStreamSynchronize(g->release_stream);
        cudaStreamDestroy(g->release_stream);
        cudaEventDestroy(g->ev_release);
    }
    if (g->rank == 0) {
        for (int r = 1; r < g->world; ++r)
            if (g->credit_at0[r]) cudaIpcCloseMemHandle(g->credit_at0[r]);
        cudaFree(g->d_credit_ptrs);
        cudaFree(g->frames[0]);
    } else if (g->frames[g->rank]) {
        cudaIpcCloseMemHandle(g->frames[g->rank]);
    }
    cudaFree(g->credit_own[g->rank]);
    (void)cudaGetLastError();
    if (s) {
        if (g->rank == 0) {
            shm_unlink(g->shm_name.c_str());
            shm_unlink(g->host_name.c_str());
        }
        munmap(s, sizeof(GroupShm));
    }
    delete g;
    return RTB_OK;
}

int rtb_group_info(const rtb_group* g, int* rank, int* world, int* slots, int* band_rows, size_t* slot_bytes) {
    if (!g) return RTB_ERR_INVALID_ARG;
    if (rank) *rank = g->rank;
    if (world) *world = g->world;
    if (slots) *slots = g->slots;
    if (band_rows) *band_rows = g->band_rows;
    if (slot_bytes) *slot_bytes = g->frame_bytes;
    return RTB_OK;
}

const char* rtb_group_last_error(const rtb_group* g) { return g ? g->last_error.c_str() : ""; }

int rtb_group_render(rtb_group* g, const float origin0[4], const float dU[4], const float dV[4], const float rayDir[4],
                     void* cuda_stream, void** d_frame_out) {
    if (!g || g->local) return RTB_ERR_INVALID_ARG;
    rtb_ctx* ctx = g->ctx[g->rank];
    if (!origin0 || !dU || !dV || !rayDir) return gfail(g, ctx, RTB_ERR_INVALID_ARG, "null camera vector");
    if (d_frame_out) *d_frame_out = nullptr;
    CameraFrame cf;
    if (!camera_prepare(origin0, dU, dV, rayDir, &cf))
        return gfail(g, ctx, RTB_ERR_CAMERA, "degenerate camera (dU, dV, rayDir do not span a ray grid)");
    float cam16[16];
    std::memcpy(cam16, origin0, 16); std::memcpy(cam16 + 4, dU, 16);
    std::memcpy(cam16 + 8, dV, 16); std::memcpy(cam16 + 12, rayDir, 16);
    cudaStream_t stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->stream;
    const int rc = group_enqueue_rank(g, g->rank, cam16, cf, stream, d_frame_out);
    if (rc != RTB_OK) {
        drain(ctx, stream);
        return gfail(g, ctx, rc, ctx->last_error);
    }
    if (!cuda_stream) RTB_CUDA(ctx, cudaStreamSynchronize(stream));
    return RTB_OK;
}

int rtb_group_prepack(rtb_group* g, int n_cameras, const float* cameras, void* cuda_stream) {
    if (!g || g->local) return RTB_ERR_INVALID_ARG;
    return rtb_prepack_cameras_bands(g->ctx[g->rank], g->width, g->height, g->band_rows, g->rank, g->world, n_cameras, cameras,
                                     cuda_stream);
}

int rtb_group_render_host(rtb_group* g, const float origin0[4], const float dU[4], const float dV[4], const float rayDir[4],
                          uint8_t** host_frame_out, float* kernel_ms) {
    if (!g || g->local) return RTB_ERR_INVALID_ARG;
    rtb_ctx* ctx = g->ctx[g->rank];
    if (host_frame_out) *host_frame_out = nullptr;
    int rc = check_frame_args(ctx, g->width, g->height, origin0, dU, dV, rayDir);
    if (rc != RTB_OK) return gfail(g, ctx, rc, ctx->last_error);
    RTB_CUDA(ctx, cudaSetDevice(ctx->device));
    CameraFrame cf;
    if (!camera_prepare(origin0, dU, dV, rayDir, &cf))
        return gfail(g, ctx, RTB_ERR_CAMERA, "degenerate camera (dU, dV, rayDir do not span a ray grid)");
    float cam16[16];
    std::memcpy(cam16, origin0, 16); std::memcpy(cam16 + 4, dU, 16);
    std::memcpy(cam16 + 8, dV, 16); std::memcpy(cam16 + 12, rayDir, 16);
    GroupShm* s = g->shm;
    const unsigned long long f = g->host_frame_no;
    if (g->rank == 0) {
        s->host_released.store(f, std::memory_order_release);  // the caller is done with frame f-1: the buffer is free again
    } else if (!poll_until(g, [&] { return s->host_released.load(std::memory_order_acquire) >= f; })) {
        return gfail(g, ctx, RTB_ERR_CUDA, "group rank 0 did not release the host frame (timeout or a failed rank)");
    }
    void* const host[3] = {g->host_frame, nullptr, nullptr};
    rc = bands_host_render(ctx, g->width, g->height, cam16, cf, g->band_rows, g->rank, g->world, host, kernel_ms);
    if (rc != RTB_OK) {
        drain(ctx, nullptr);
        return gfail(g, ctx, rc, ctx->last_error);
    }
    s->host_done[g->rank].store(f + 1ull, std::memory_order_release);
    g->host_frame_no = f + 1ull;
    if (g->rank == 0) {
        for (int r = 1; r < g->world; ++r)
            if (!poll_until(g, [&] { return s->host_done[r].load(std::memory_order_acquire) >= f + 1ull; }))
                return gfail(g, ctx, RTB_ERR_CUDA, "a group rank did not deliver its bands (timeout or a failed rank)");
        if (host_frame_out) *host_frame_out = g->host_frame;
    }
    return RTB_OK;
}

int rtb_group_barrier(rtb_group* g) {
    if (!g || g->local) return RTB_ERR_INVALID_ARG;
    GroupShm* s = g->shm;
    const uint64_t mine = s->barrier_count.fetch_add(1ull, std::memory_order_acq_rel) + 1ull;
    const uint64_t target = (mine + (uint64_t)g->world - 1ull) / (uint64_t)g->world * (uint64_t)g->world;  // end of this round
    if (!poll_until(g, [&] { return s->barrier_count.load(std::memory_order_acquire) >= target; }))
        return gfail(g, g->ctx[g->rank], RTB_ERR_CUDA, "group barrier timed out (or a rank failed)");
    return RTB_OK;
}

}  // extern "C"
