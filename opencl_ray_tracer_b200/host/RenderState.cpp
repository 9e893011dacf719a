#include "RenderState.h"

#include <chrono>

#include "PngWriter.h"

namespace rtbh {

RenderState::RenderState(int w, int h, int device, int gpus)
    : width(w), height(h), pixelCount(w * h), timeTaken(0), kernelMs(0.f), totalMs(0.f), currentScene(1),
      sceneSeed(1), sceneChange(true), start(true), currentMode(CUDA), ctx_(nullptr), randomNs_(0),
      randomNc_(0) {
    // MainState.cpp:37-50.  glm::perspective(45, 4/3, zNear = 0, 100) * (0,0,1,1): only column 2
    // (0, 0, -(f+n)/(f-n), -1) and column 3 (0, 0, -2fn/(f-n), 0) contribute, giving (0,0,-1,-1).
    const float zNear = 0.0f, zFar = 100.0f;
    const float m22 = -(zFar + zNear) / (zFar - zNear);
    const float m32 = -(2.0f * zFar * zNear) / (zFar - zNear);
    rayDir = {0.0f, 0.0f, m22 * 1.0f + m32 * 1.0f, -1.0f * 1.0f};
    rayOrigin0 = {0.f, 0.f, 0.f, 1.f};
    rayOriginDU = {1.f, 0.f, 0.f, 0.f};
    rayOriginDV = {0.f, 1.f, 0.f, 0.f};
    int dev[RTB_MAX_GROUP];
    if (gpus < 1) gpus = 1;
    if (gpus > RTB_MAX_GROUP) gpus = RTB_MAX_GROUP;
    for (int k = 0; k < gpus; ++k) dev[k] = device + k;
    initStatus_ = rtb_init(&ctx_, gpus, dev);  // replaces openCLInit(), MainState.cpp:52
    if (initStatus_ != RTB_OK) lastError = rtb_error_string(initStatus_);
}

RenderState::~RenderState() { rtb_shutdown(ctx_); }

void RenderState::toggleMode() {  // F1 cycles CPU -> OpenCL -> CUDA -> CPU
    currentMode = (currentMode == CPU) ? OpenCL : (currentMode == OpenCL ? CUDA : CPU);
    start = true;
}

void RenderState::nextScene() {  // F2: 1 -> 2 -> 3 -> 1 (MainState.cpp:156-172)
    currentScene = (currentScene >= 3) ? 1 : currentScene + 1;
    sceneChange = true;
    start = true;
}

void RenderState::setScene(int s, uint64_t seed) {
    currentScene = s;
    sceneSeed = seed;
    sceneChange = true;
    start = true;
}

void RenderState::setRandomScene(int ns, int nc, uint64_t seed) {
    currentScene = 0;
    randomNs_ = ns; randomNc_ = nc;
    sceneSeed = seed;
    sceneChange = true;
    start = true;
}

int RenderState::update() {
    if (!start) return RTB_OK;
    start = false;
    if (initStatus_ != RTB_OK) return initStatus_;
    if (sceneChange) {  // MainState.cpp:185-210
        scene.clear();
        if (currentScene == 1) scene.createScene1();
        else if (currentScene == 2) scene.createScene2(sceneSeed);
        else if (currentScene == 3) scene.createScene3(sceneSeed);
        else scene.createRandom(width, height, randomNs_, randomNc_, sceneSeed);
        std::vector<float> so, sr, sc, cv, cc;
        scene.flatten(&so, &sr, &sc, &cv, &cc);
        const int rc = rtb_set_scene(ctx_, (int)sr.size(), so.data(), sr.data(), sc.data(),
                                     (int)scene.cubes.size(), cv.data(), cc.data());
        if (rc != RTB_OK) { lastError = rtb_last_error_message(ctx_); return rc; }
        sceneChange = false;
    }
    pixels.clear();  // MainState.cpp:213-215
    int rc;
    if (currentMode == CUDA) {
        rc = executeRayTracerCUDA();
    } else {
        lastError = "Mode::CPU / Mode::OpenCL are the reference's own paths; this library ships only Mode::CUDA";
        return RTB_ERR_UNSUPPORTED;
    }
    if (rc != RTB_OK) return rc;
    generateImageFromPixels();
    return RTB_OK;
}

int RenderState::setTileCull(int mode) {
    if (!ctx_) return initStatus_;
    const int rc = rtb_set_option(ctx_, RTB_OPT_TILE_CULL, mode);
    if (rc != RTB_OK) lastError = rtb_last_error_message(ctx_);
    start = true;
    return rc;
}

int RenderState::executeRayTracerCUDA() {
    pixels.resize((size_t)pixelCount * 4);
    primIds.resize((size_t)pixelCount);
    const auto t0 = std::chrono::steady_clock::now();
    const int rc = rtb_render(ctx_, width, height, &rayOrigin0.x, &rayOriginDU.x, &rayOriginDV.x, &rayDir.x,
                              nullptr, primIds.data(), pixels.data(), &kernelMs, &totalMs);
    const auto t1 = std::chrono::steady_clock::now();
    timeTaken = (uint64_t)std::chrono::duration_cast<std::chrono::microseconds>(t1 - t0).count();
    if (rc != RTB_OK) lastError = rtb_last_error_message(ctx_);
    return rc;
}

void RenderState::generateImageFromPixels() {
    image.resize(pixels.size());
    for (size_t i = 0; i < pixels.size(); ++i) image[i] = (unsigned char)pixels[i];
}

unsigned RenderState::encodePNG(const char* filename) const {
    return rtbh::encodePNG(filename, image, (unsigned)width, (unsigned)height);
}

}  // namespace rtbh
