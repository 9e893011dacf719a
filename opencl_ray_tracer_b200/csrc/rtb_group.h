// rtb_group.h -- multi-GPU frames behind the C ABI (rtb_group.cu): the parts rtb_api.cu forwards to when a context
// was created with rtb_init(n_devices > 1) ("local" group: every device in this process).
#ifndef RTB_GROUP_H_
#define RTB_GROUP_H_

#include "rtb_internal.h"

namespace rtb {

int group_init_local(rtb_ctx** out, int n_devices, const int* device_ids);
void group_shutdown_local(rtb_ctx* root);
rtb_ctx* group_root_local(rtb_ctx* root);  // the first device's own single-device context
int group_set_scene_local(rtb_ctx* root, int numSpheres, const float* sphereOrigins, const float* sphereRadius,
                          const float* sphereColours, int numCubes, const float* cubeVertices, const float* cubeColours);
int group_render_local(rtb_ctx* root, int width, int height, const float origin0[4], const float dU[4], const float dV[4],
                       const float rayDir[4], uint8_t* rgba8_out, int32_t* prim_id_out, int32_t* rgba_i32_out,
                       float* kernel_ms, float* total_ms);
int group_render_sequence_local(rtb_ctx* root, int width, int height, int n_frames, const float* cameras, void* d_rgba8,
                                float* kernel_ms_total);
void group_set_option_local(rtb_ctx* root, int option, int value);
unsigned long long group_launch_count_local(const rtb_ctx* root);
void group_transfer_bytes_local(const rtb_ctx* root, unsigned long long* h2d, unsigned long long* d2h);

}  // namespace rtb

#endif
