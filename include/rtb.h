/* rtb.h -- C ABI of the B200 ray-casting backend ("rtb").
 *
 * This is the drop-in boundary for ONE path of RichardHancock/OpenCL-Ray-Tracer: the per-frame
 * primary-ray cast that the reference runs either on the CPU (MainState::executeRayTracerCPU,
 * RayTrace/states/MainState.cpp:936-972) or through OpenCL (MainState::openCLInit :1181-1326 +
 * MainState::executeRayTracerOpenCL :641-934 + the kernel RayTrace/resources/shaders/
 * rayTracer.cl:111-202).  A third `Mode::CUDA` behind the same toggle (MainState.cpp:217-224)
 * calls these entry points; INTEGRATION.md shows the binding.
 *
 * Conventions: plain C types only; every function returns RTB_OK (0) or an rtb_status code and
 * never throws; the context is opaque and owned by the library; output buffers are owned by the
 * caller; calls block until the result is where the caller asked for it (like the reference's
 * blocking clEnqueueWriteBuffer / clEnqueueMapBuffer sequence); one context is used from one
 * thread at a time.  There is no CPU fallback: without a CUDA device rtb_init fails with
 * RTB_ERR_NO_DEVICE.
 *
 * Scene arrays have EXACTLY the layout of the OpenCL kernel's arguments (rayTracer.cl:111-114),
 * i.e. the flattening done at MainState.cpp:646-655:
 *   sphereOrigins  4*ns floats (x,y,z,w)      sphereRadius  ns floats     sphereColours 4*ns floats
 *   cubeVertices   4*36*nc floats (12 triangles x 3 vertices per cube, Cube::getTriangles() order)
 *   cubeColours    4*nc floats
 * Colours are (r,g,b,255) with r,g,b in [0,1] as the reference stores them.
 *
 * Camera: the reference has one shared direction `rayDir` and an explicit origin per pixel,
 * rayOrigins[y*W+x] = (x, y, 0, 1) (MainState.cpp:37-50).  The grid entry points take the origin
 * grid in affine form origin(x,y) = origin0 + x*dU + y*dV (reference values: origin0=(0,0,0,1),
 * dU=(1,0,0,0), dV=(0,1,0,0), rayDir=(0,0,-1,-1)); rtb_render_rays keeps the explicit form.
 *
 * Outputs (each nullable):
 *   rgba8     4*N bytes   the (uint8_t)-narrowed frame of MainState.cpp:1023-1037, row-major, y down
 *   prim_id   N int32     winning primitive: cubes 0..nc-1, spheres nc..nc+ns-1, -1 = miss (extension)
 *   rgba_i32  4*N int32   the reference's `std::vector<int> pixels` stream (MainState.cpp:952-955,
 *                         rayTracer.cl:198-201), for assignment straight into `pixels`
 */
#ifndef RTB_H_
#define RTB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rtb_ctx rtb_ctx;

typedef enum rtb_status {
    RTB_OK = 0,
    RTB_ERR_INVALID_ARG = 1,  /* null pointer, non-positive size, ...                          */
    RTB_ERR_NO_DEVICE = 2,    /* no usable CUDA device / device id out of range                */
    RTB_ERR_CUDA = 3,         /* a CUDA runtime call failed; see rtb_last_error_message        */
    RTB_ERR_NO_SCENE = 4,     /* render called before rtb_set_scene                            */
    RTB_ERR_NOT_A_BOX = 5,    /* a cube's 36 vertices are not an affine image of Cube.cpp's    */
    RTB_ERR_CAMERA = 6,       /* degenerate camera: dU, dV, rayDir do not span a ray grid      */
    RTB_ERR_OUT_OF_MEMORY = 7,
    RTB_ERR_UNSUPPORTED = 8   /* e.g. no peer access between the devices of a multi-device context */
} rtb_status;

#define RTB_MAX_GROUP 16      /* devices of one context / ranks of one group */

/* Replaces MainState::openCLInit (MainState.cpp:1181-1326: platform/device pick, context, queue,
 * program build, kernel).  device_ids may be NULL (= devices 0 .. n_devices-1).  Unlike the reference,
 * failures are reported, not printed and ignored.
 * n_devices = 1: one device.  n_devices = 2 .. RTB_MAX_GROUP: the context drives that many GPUs of this box from
 * this one process (a worker thread per device, peer access over NVLink): rtb_set_scene uploads to every device,
 * rtb_render shards each frame as interleaved 16-row bands and every GPU delivers its bands over its own PCIe
 * link into the caller's buffers, rtb_render_sequence_device keeps the frames on the first device (the others
 * store their bands into it over NVLink).  Still ONE blocking call per frame, like the reference's
 * (MainState.cpp:217-224).  The device-pointer entry points (rtb_render_bands_*, rtb_peer_frame_*, ...) are the
 * building blocks of the one-process-per-GPU form (rtb_group_join below) and need a single-device context. */
int rtb_init(rtb_ctx** out, int n_devices, const int* device_ids);

/* Replaces the four clRelease* calls of MainState::~MainState (MainState.cpp:75-78). */
void rtb_shutdown(rtb_ctx* ctx);

/* Replaces MainState::getErrorString (MainState.cpp:1101-1179). */
const char* rtb_error_string(int status);

/* Detail text of the last failure on this context (CUDA error string, offending cube index...). */
const char* rtb_last_error_message(const rtb_ctx* ctx);

/* Per-context switches; every render entry point below obeys them.  No counterpart in the
 * reference (its only switch is enum Mode, MainState.h:48-52).
 *   RTB_OPT_TILE_CULL (0 | 1 | 2 | 3, default 3): which kernels render.  The pixels are byte-identical in all
 *     modes (the culls are statements about the values the per-ray bound would compute).
 *     3 = auto (the default): 2 for scenes of 64 primitives or more, else 0 -- the fastest byte-identical mode is
 *         what a caller gets without asking.
 *     0 = the dense O(rays x primitives) kernel that BASELINE.json's roofline point is quoted on.
 *     1 = raycast_cull_kernel: every CTA first tests each primitive's conservative screen-space bound
 *         against its whole 256 x 8 pixel tile and runs the per-ray tests only for the survivors:
 *         O(tiles x primitives + rays x survivors) in one launch.
 *     2 = binned: bin_tiles_kernel makes that decision once per frame (two-level test, one 32-bit word
 *         per 64 x 32 pixel tile and 32 primitives, grow-only device buffer owned by the context),
 *         raycast_binned_kernel then renders every tile from its words.  Two launches per frame, the
 *         second a programmatic dependent of the first; the fastest mode on SURVEY.md s8d's scenes.
 *         With mode 1 or 2 rtb_render copies the frame to the host in one piece behind the launch
 *         (the launch is much shorter than the PCIe time, there is nothing to overlap).
 *   RTB_OPT_DEFER_SCENE_CHECK (0 | 1, default 0): 0 = rtb_set_scene waits for the device's box check and returns
 *     RTB_ERR_NOT_A_BOX itself.  1 = rtb_set_scene returns as soon as the arrays are staged (no wait for the
 *     device); the verdict is returned by the next render call instead (by a blocking one after its own wait for
 *     the frame, so nothing ever waits for the scene upload alone).
 */
typedef enum rtb_option { RTB_OPT_TILE_CULL = 1, RTB_OPT_DEFER_SCENE_CHECK = 2 } rtb_option;
int rtb_set_option(rtb_ctx* ctx, int option, int value);
int rtb_get_option(const rtb_ctx* ctx, int option, int* value);

/* Replaces the 6 clCreateBuffer + clEnqueueWriteBuffer pairs for the scene
 * (MainState.cpp:677-839) and the clSetKernelArg calls 1-7 (:747-753).  The arrays are copied to
 * the device as they are (the caller may free them on return) and stay resident until the next
 * rtb_set_scene; a device kernel recovers each cube's box from its 36 vertices and the call fails
 * with RTB_ERR_NOT_A_BOX (lowest offending index in rtb_last_error_message) if one is not a box. */
int rtb_set_scene(rtb_ctx* ctx, int numSpheres, const float* sphereOrigins,
                  const float* sphereRadius, const float* sphereColours, int numCubes,
                  const float* cubeVertices, const float* cubeColours);

/* Replaces the rest of MainState::executeRayTracerOpenCL (rays upload :841-851, launch :858-869,
 * blocking read-back :876-886, pixels.assign :907): renders one width x height frame into HOST
 * buffers.  kernel_ms (nullable) = device time of the ray-cast kernel alone; total_ms (nullable)
 * = upload + kernel + download, the analogue of the reference's OpenCL timer (:662-894). */
int rtb_render(rtb_ctx* ctx, int width, int height, const float origin0[4], const float dU[4],
               const float dV[4], const float rayDir[4], uint8_t* rgba8_out, int32_t* prim_id_out,
               int32_t* rgba_i32_out, float* kernel_ms, float* total_ms);

/* Explicit-origin form, argument-for-argument the OpenCL kernel's `rayOrigins` + `rayDir`
 * (rayTracer.cl:111-114; MainState.cpp:841-851): n_rays origins of 4 floats each, host memory. */
int rtb_render_rays(rtb_ctx* ctx, size_t n_rays, const float* rayOrigins, const float rayDir[4],
                    uint8_t* rgba8_out, int32_t* prim_id_out, int32_t* rgba_i32_out,
                    float* kernel_ms, float* total_ms);

/* Device-resident variant for multi-GPU sharding and for kernel-only timing.  Renders the
 * interleaved row bands owned by `band_rank` out of `band_world` (band b = rows
 * [b*band_rows, (b+1)*band_rows) belongs to rank b % band_world) into PACKED device buffers:
 * local row = (b / band_world)*band_rows + (y - b*band_rows).  band_world = 1 renders the whole
 * frame.  band_rows must be a positive multiple of 4.  d_* are device pointers on this context's
 * device (nullable); cuda_stream is a cudaStream_t (NULL = the context's own stream).  The call
 * only enqueues work when cuda_stream != NULL; with the context's stream it also synchronises. */
int rtb_render_bands_device(rtb_ctx* ctx, int width, int height, const float origin0[4],
                            const float dU[4], const float dV[4], const float rayDir[4],
                            int band_rows, int band_rank, int band_world, void* d_rgba8,
                            void* d_prim_id, void* d_rgba_i32, void* cuda_stream);

/* Number of local rows rank `band_rank` owns for a frame of `height` rows. */
int rtb_band_local_rows(int height, int band_rows, int band_rank, int band_world);

/* Row-band render of this rank's bands straight into a WHOLE rgba8 frame (row = frame row).
 * d_frame_rgba8 may be this GPU's memory or another GPU's frame mapped with rtb_peer_frame_open:
 * the kernel's 128-bit framebuffer stores then travel over NVLink / NVSwitch and no gather or
 * de-interleave pass is needed (the caller orders "all ranks done" with one tiny collective).
 * New (the reference is single-device); replaces the per-render clEnqueueMapBuffer read-back of
 * MainState.cpp:876-886 for the multi-GPU frame. */
int rtb_render_bands_frame(rtb_ctx* ctx, int width, int height, const float origin0[4],
                           const float dU[4], const float dV[4], const float rayDir[4], int band_rows,
                           int band_rank, int band_world, void* d_frame_rgba8, void* cuda_stream);

/* Row-band render of this rank's bands straight into a WHOLE rgba8 frame in HOST memory (row = frame
 * row): the multi-GPU counterpart of rtb_render.  With the frame in memory shared by the ranks'
 * processes (e.g. POSIX shared memory, pinned in each with rtb_host_register) every GPU delivers its
 * bands over its own PCIe link, in parallel, and finished chunks of bands leave while the launch is
 * still running.  Blocking; the caller orders "all ranks done" (a barrier).  band_world = 1 is a
 * whole-frame render.  kernel_ms nullable.  Replaces the clEnqueueMapBuffer read-back of
 * MainState.cpp:876-886 for the multi-GPU frame. */
int rtb_render_bands_host(rtb_ctx* ctx, int width, int height, const float origin0[4], const float dU[4],
                          const float dV[4], const float rayDir[4], int band_rows, int band_rank,
                          int band_world, uint8_t* host_frame_rgba8, float* kernel_ms);

/* Page-locks caller-owned host memory (e.g. a shared-memory frame) for direct DMA; undo with
 * rtb_host_unregister before freeing it.  Optional: pageable buffers work, only slower. */
int rtb_host_register(void* p, size_t bytes);
int rtb_host_unregister(void* p);

/* As rtb_render_bands_frame, and the launch's last CTA adds 1 (system-scope atomic) to the 32-bit
 * word d_signal -- typically the signal word of rank 0's peer frame, at
 * frame + rtb_peer_frame_signal_offset(bytes) -- once all of this rank's stores are ordered before
 * it.  Rank 0 then orders "every rank's bands have landed" on its stream with
 * rtb_stream_wait_signal(ctx, signal, world * frames_so_far, stream): no collective on the frame's
 * critical path.  rtb_stream_wait_signal returns RTB_ERR_UNSUPPORTED without stream memory
 * operations (use a collective as the fence then). */
int rtb_render_bands_frame_signal(rtb_ctx* ctx, int width, int height, const float origin0[4],
                                  const float dU[4], const float dV[4], const float rayDir[4],
                                  int band_rows, int band_rank, int band_world, void* d_frame_rgba8,
                                  void* d_signal, void* cuda_stream);
int rtb_stream_wait_signal(rtb_ctx* ctx, void* d_signal, unsigned int value, void* cuda_stream);
size_t rtb_peer_frame_signal_offset(size_t bytes);

/* A device frame buffer other processes (one per GPU) can map: create on the owning rank (returns
 * the device pointer and a 64-byte CUDA IPC handle to hand to the peers, e.g. with a broadcast),
 * open on the peers, close on both (owner != 0 frees the allocation). */
/* The allocation is bytes rounded up to 256, followed by 256 zeroed bytes of signal words. */
int rtb_peer_frame_create(rtb_ctx* ctx, size_t bytes, void** d_ptr, unsigned char handle_out[64]);
int rtb_peer_frame_open(rtb_ctx* ctx, const unsigned char handle[64], void** d_ptr);
int rtb_peer_frame_close(rtb_ctx* ctx, void* d_ptr, int owner);

/* Multi-GPU gather epilogue.  After the ranks' packed band buffers have been gathered back to back
 * (rank 0's rows, then rank 1's, ...; rank r contributes rtb_band_local_rows(height, band_rows, r,
 * band_world) rows) into d_packed, this copies every row to its place in the final frame d_frame.
 * elem_bytes = 4 (rgba8 / prim_id) or 16 (rgba_i32).  Stream semantics as above.  (The reference has
 * no counterpart: it is single-device.) */
int rtb_assemble_bands_device(rtb_ctx* ctx, int width, int height, int band_rows, int band_world,
                              int elem_bytes, const void* d_packed, void* d_frame, void* cuda_stream);

/* n_frames frames of the same scene with per-frame cameras (cameras = n_frames x 16 floats:
 * origin0, dU, dV, rayDir), kept on the device: d_rgba8 receives n_frames packed frames
 * (4*width*height bytes each).  Used for the orbit sequence (BASELINE.json config 5). */
int rtb_render_sequence_device(rtb_ctx* ctx, int width, int height, int n_frames,
                               const float* cameras, void* d_rgba8, float* kernel_ms_total);

/* Packs the per-frame records of up to 256 cameras AHEAD of their frames (16 floats each: origin0, dU, dV,
 * rayDir), on a side stream of the context and behind whatever is already enqueued on `stream` (NULL = the
 * context's own stream).  A later rtb_render* call whose camera and frame size equal one of them skips its
 * own pack and only makes its stream wait for that camera's: the FP64 pack launch (one thread per primitive,
 * ~20 us) leaves the frame's critical path -- what rtb_render_sequence_device does internally, offered for
 * callers that drive the frames themselves (one process per GPU, rtb_render_bands_frame per frame).  Every call
 * packs anew (nothing is reused from an earlier call); the slots are dropped by the next rtb_prepack_cameras,
 * rtb_render_sequence_device or rtb_set_scene.  Cameras beyond the capacity are simply packed by their own
 * render call.  No counterpart in the reference (it has one fixed camera, MainState.cpp:37-50). */
int rtb_prepack_cameras(rtb_ctx* ctx, int width, int height, int n_cameras, const float* cameras, void* stream);
/* The same for a rank that renders only its row bands (the dense kernel's row table is then built for those rows
 * only; a render call with other bands simply builds its own). */
int rtb_prepack_cameras_bands(rtb_ctx* ctx, int width, int height, int band_rows, int band_rank, int band_world,
                              int n_cameras, const float* cameras, void* stream);

/* ---- multi-GPU frames, one process per GPU: ONE call per frame and rank, no collective, no Python -------------
 * The ranks of a job (all on one box) meet in a POSIX shared-memory control block called `name` (any string unique
 * to the job; rank 0 creates it).  Rank 0 owns a ring of `slots` frame slots in its device memory, mapped into every
 * rank (CUDA IPC).  rtb_group_render enqueues, on the caller's stream (NULL = the context's, then it also waits):
 *   every rank   wait for its credit word (slot free again) -> [record pack] -> dense / binned launch whose row stores
 *                go straight into rank 0's slot over NVLink -> a stream write of this rank's done word in rank 0's memory
 *                (ordered behind the launch and its stores; without cuStreamWriteValue32 the launch's last CTA bumps it);
 *   rank 0       additionally: after its own launch ONE batched stream wait on every other rank's done word (one word per
 *                rank: a rank that runs ahead cannot stand in for one that lags), and -- on a side stream, behind an event
 *                of this one -- the credits for the previous frame's slot (a 1-warp kernel writing every rank's credit
 *                word through the peer mapping).  *d_frame_out (rank 0) = the slot: complete in stream order behind this
 *                call, valid until rank 0's NEXT rtb_group_render call is reached in stream order (whatever rank 0
 *                enqueues on the stream in between -- a copy, a consumer kernel -- runs before the slot is handed back).
 * With slots >= 2 the other ranks run up to slots-1 frames ahead of rank 0's consumer.  Every rank must call it
 * once per frame with the same camera.  Replaces, for the sharded frame, the per-render sequence of
 * MainState.cpp:641-934 (and round 1's Python loop with an NCCL all-reduce per frame).
 * rtb_group_render_host is the blocking host-buffer form (the multi-GPU rtb_render): every GPU copies its bands
 * over its own PCIe link into ONE frame in shared, page-locked host memory; on rank 0 the call returns when every
 * rank has delivered and *host_frame_out points at the frame (valid until rank 0's next call); completion and
 * release travel through counters in the control block, not through a collective.
 * rtb_group_prepack = rtb_prepack_cameras_bands for this rank's bands.  rtb_group_barrier: host barrier of the ranks.
 * rtb_group_leave: collective; drains this rank's device first. */
typedef struct rtb_group rtb_group;
int rtb_group_join(rtb_ctx* ctx, const char* name, int rank, int world, int width, int height, int band_rows, int slots,
                   rtb_group** out);
int rtb_group_leave(rtb_group* g);
int rtb_group_info(const rtb_group* g, int* rank, int* world, int* slots, int* band_rows, size_t* slot_bytes);
const char* rtb_group_last_error(const rtb_group* g);
int rtb_group_render(rtb_group* g, const float origin0[4], const float dU[4], const float dV[4], const float rayDir[4],
                     void* cuda_stream, void** d_frame_out);
int rtb_group_render_host(rtb_group* g, const float origin0[4], const float dU[4], const float dV[4], const float rayDir[4],
                          uint8_t** host_frame_out, float* kernel_ms);
int rtb_group_prepack(rtb_group* g, int n_cameras, const float* cameras, void* cuda_stream);
int rtb_group_barrier(rtb_group* g);

/* One frame with NO allocation inside the call (no cudaMalloc / cudaFree / allocating library routine, no context).
 * This entry point has no counterpart in the reference; it was added for
 * callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that forbid it.
 * Inputs and outputs are the CALLER's device buffers on `device` (< 0: the current device): the six scene arrays in
 * the layout of rtb_set_scene (float4 arrays 16-byte aligned), and any of d_rgba8 / d_prim_id / d_rgba_i32 (nullable,
 * whole width x height frames).  d_bad_cube (nullable, device int) receives the lowest index of a cube that is not a
 * box, or 0x7F7F7F7F if every cube is one (the verdict rtb_set_scene returns as RTB_ERR_NOT_A_BOX; reading it is the
 * caller's business, nothing here waits for the device).  cuda_stream != NULL: the call only enqueues (4 kernels, one
 * memset, at most one 4-byte copy); NULL: default stream, and the call waits for it.
 * The binned kernels run (RTB_OPT_TILE_CULL 2: byte-identical to every other mode); their scratch memory -- recovered
 * boxes, the frame's records, the per-tile primitive words -- is a file-scope static __device__ array of about 0.7 MB
 * that every GPU which loads the library holds until the process exits.  LIMITS: numSpheres + numCubes <=
 * RTB_NOALLOC_MAX_PRIMITIVES and ceil(width / 64) * ceil(height / 32) * ceil(primitives / 32) <=
 * RTB_NOALLOC_MAX_BIN_WORDS (3840 x 2160 at 1024 primitives fits); beyond them the call returns RTB_ERR_INVALID_ARG
 * with the reason in err.  All calls on a device share the array: calls in one stream are independent (everything
 * read is rewritten by the call itself), calls that OVERLAP on the GPU (two streams, no event between them) race.
 * err (nullable, err_cap bytes) receives the detail text of a failure.  (The first call on a device loads the
 * kernels' code, which the driver may allocate for: make one call before starting a graph capture.) */
#define RTB_NOALLOC_MAX_PRIMITIVES 1024
#define RTB_NOALLOC_MAX_BIN_WORDS 131072
int rtb_render_device_noalloc(int device, int width, int height, const float origin0[4], const float dU[4],
                              const float dV[4], const float rayDir[4], int numSpheres, const float* d_sphereOrigins,
                              const float* d_sphereRadius, const float* d_sphereColours, int numCubes,
                              const float* d_cubeVertices, const float* d_cubeColours, void* d_rgba8, void* d_prim_id,
                              void* d_rgba_i32, int* d_bad_cube, void* cuda_stream, char* err, size_t err_cap);

/* Pinned host allocations for the e2e path (pageable buffers work too, only slower). */
int rtb_host_alloc(void** out, size_t bytes);
void rtb_host_free(void* p);

/* Introspection. */
int rtb_device_info(const rtb_ctx* ctx, int* sm_count, int* sm_clock_khz, char* name, size_t name_cap);
/* Kernel launches issued by this context so far (ray-cast + helper kernels). */
unsigned long long rtb_launch_count(const rtb_ctx* ctx);
/* Bytes this context has copied host->device / device->host so far (scene, records, frames). */
void rtb_transfer_bytes(const rtb_ctx* ctx, unsigned long long* h2d, unsigned long long* d2h);
/* Average device time (ms) of the ray-cast kernel launches since the last call with reset != 0;
 * measured with CUDA events on the launching stream.  n_launches (nullable) receives the count. */
float rtb_kernel_ms_avg(rtb_ctx* ctx, int reset, int* n_launches);

/* Debug/test entry: runs only the host-side packer (double-precision camera-space precompute)
 * and copies the packed records out.  cube_rec: 12 floats per cube, sphere_rec: 4 floats per
 * sphere, cam: 8 floats.  No device work.  ctx may be NULL-device-free: see rtb_pack_only_new. */
int rtb_debug_pack(int numSpheres, const float* sphereOrigins, const float* sphereRadius,
                   int numCubes, const float* cubeVertices, const float origin0[4],
                   const float dU[4], const float dV[4], const float rayDir[4], float* cube_rec,
                   float* sphere_rec, float* cam, int* bad_cube_index);

/* Debug/test entry: the conservative conic bound records (8 floats per primitive, cubes then
 * spheres; raycast_math.h RtbBound) the grid kernel's dense loop evaluates for a width x height
 * frame of this camera.  Host only.  Replaces nothing in the reference: the bound is an internal
 * accelerator of the rayTracer.cl:140-184 loops and never changes a pixel. */
int rtb_debug_pack_bounds(int numSpheres, const float* sphereOrigins, const float* sphereRadius,
                          int numCubes, const float* cubeVertices, int width, int height,
                          const float origin0[4], const float dU[4], const float dV[4],
                          const float rayDir[4], float* bound_rec);

/* Debug/test entry: the dense kernel's row table (raycast_math.h RtbRowEntry: 2 floats per (frame row, bound
 * record), table[(y * n_bound + i) * 2]) as the HOST packer builds it from n_bound bound records of 8 floats.
 * Host only.  rtb_debug_device_row_table returns the table the DEVICE builds (row_table_kernel) for the rows of
 * the given bands (other rows 0); the two must agree bit for bit.  *entries_per_row = bound records incl. padding. */
int rtb_debug_pack_row_table(int n_bound, const float* bound_rec, int width, int height, const float origin0[4],
                             const float dU[4], const float dV[4], const float rayDir[4], float* table);
int rtb_debug_device_row_table(rtb_ctx* ctx, int width, int height, const float origin0[4], const float dU[4],
                               const float dV[4], const float rayDir[4], int band_rows, int band_rank, int band_world,
                               float* table, size_t table_cap_floats, int* entries_per_row);

/* Debug/test entry: packs the records of one frame ON THE DEVICE (the per-frame path of rtb_render*)
 * and copies them out: 4 floats per record in the layout of rtb_debug_pack (3 per cube, 1 per sphere)
 * followed, from float index *bound_offset_floats, by 8 floats per bound.  With out == NULL only the
 * sizes are returned.  Must equal the host packer's records bit for bit. */
int rtb_debug_device_records(rtb_ctx* ctx, int width, int height, const float origin0[4],
                             const float dU[4], const float dV[4], const float rayDir[4], float* out,
                             size_t out_cap_floats, size_t* n_floats, size_t* bound_offset_floats);

/* Debug/test entry: the tiling the grid kernel uses for a width x rows launch on a GPU with
 * sm_count SMs: out = {rays per thread, tile width, tile height, CTAs}.  Host only. */
int rtb_debug_grid_shape(int width, int rows, int sm_count, int out[4]);

const char* rtb_version(void);

#ifdef __cplusplus
}
#endif
#endif /* RTB_H_ */
