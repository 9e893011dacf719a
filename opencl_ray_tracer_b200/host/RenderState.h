// RenderState.h -- headless host mirror of the reference's render state `MainState`
// (RayTrace/states/MainState.h:48-142, MainState.cpp:26-71, :135-239): same members on the ray-cast
// path (rayDir, the pixel grid of ray origins, scene vectors, `std::vector<int> pixels`,
// `timeTaken` in microseconds, currentScene / sceneChange, the Mode toggle) with a third mode,
// Mode::CUDA, whose work is done by the sm_100a kernels behind include/rtb.h.
//
// The reference's own Mode::CPU and Mode::OpenCL paths are NOT provided by this library (no CPU
// fallback, no OpenCL): selecting them makes update() return RTB_ERR_UNSUPPORTED.  SDL windowing,
// fonts and input are out of scope; F1/F2/Space become toggleMode()/nextScene()/requestRender().
#ifndef RTB_HOST_RENDERSTATE_H_
#define RTB_HOST_RENDERSTATE_H_

#include <cstdint>
#include <string>
#include <vector>

#include "../../include/rtb.h"
#include "Ray.h"
#include "Scene.h"

namespace rtbh {

class RenderState {
public:
    enum Mode { CPU, OpenCL, CUDA };

    // gpus > 1: the frame is sharded over devices device .. device+gpus-1 of this box (rtb_init with n_devices > 1)
    RenderState(int width, int height, int device = 0, int gpus = 1);
    ~RenderState();
    RenderState(const RenderState&) = delete;
    RenderState& operator=(const RenderState&) = delete;

    int initStatus() const { return initStatus_; }  // result of rtb_init (cf. openCLInit)

    // F1 / F2 / Space of MainState::update (MainState.cpp:137-177)
    void toggleMode();
    void setMode(Mode m) { currentMode = m; start = true; }
    void nextScene();
    void setScene(int scene, uint64_t seed = 1);
    void setRandomScene(int ns, int nc, uint64_t seed);
    void requestRender() { start = true; }
    // rtb_set_option(RTB_OPT_TILE_CULL): 0 dense kernel, 1 per-tile cull, 2 binned (same pixels, include/rtb.h).
    int setTileCull(int mode);

    // One pass of MainState::update's render branch (MainState.cpp:180-237): rebuild the scene if it
    // changed, clear `pixels`, run the current mode, convert to the RGBA8 image.  Returns an
    // rtb_status.
    int update();

    int executeRayTracerCUDA();        // third path behind the toggle (cf. :641-934, :936-972)
    void generateImageFromPixels();    // (uint8_t) narrowing of :1023-1037 into `image`
    unsigned encodePNG(const char* filename) const;  // :410-417

    // --- state, named as in MainState.h --------------------------------------------------------
    int width, height;
    int pixelCount;
    Vec4 rayDir;                       // perspective(45, 4/3, 0, 100) * (0,0,1,1) = (0,0,-1,-1)
    Vec4 rayOrigin0, rayOriginDU, rayOriginDV;  // rayOrigins[y*W+x] = origin0 + x*dU + y*dV
    Scene scene;
    std::vector<int> pixels;           // 4 ints per pixel, RGBA
    std::vector<unsigned char> image;  // RGBA8 framebuffer (stands in for the SDL Texture)
    std::vector<int32_t> primIds;      // extension: winning primitive per pixel
    uint64_t timeTaken;                // microseconds, like PerformanceCounter
    float kernelMs, totalMs;
    int currentScene;
    uint64_t sceneSeed;
    bool sceneChange;
    bool start;
    Mode currentMode;
    std::string lastError;

private:
    rtb_ctx* ctx_;
    int initStatus_;
    int randomNs_, randomNc_;
};

}  // namespace rtbh

#endif
