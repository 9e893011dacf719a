// raycast_cli.cpp -- headless stand-in for the reference's main loop (RayTrace/main.cpp:24-88):
// builds a RenderState, selects scene / mode from flags (the reference's F1 / F2 / Space keys),
// renders once and optionally writes a PNG.  Prints the same "Time Taken" line the reference
// prints (MainState.cpp:903, :969) plus rays/s.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "RenderState.h"

using namespace rtbh;

static void usage() {
    std::printf(
        "raycast_cli [--width W] [--height H] [--scene 1|2|3|random] [--spheres N] [--cubes N] [--seed S]\n"
        "            [--mode cuda|cpu|opencl] [--device D] [--gpus N] [--frames K] [--tile-cull 0|1|2|3] [--out file.png]\n");
}

int main(int argc, char** argv) {
    int W = 640, H = 480, ns = 100, nc = 100, device = 0, gpus = 1, frames = 1, tileCull = -1;  // 640x480: defaultSettings.xml
    unsigned long long seed = 1;
    std::string scene = "1", mode = "cuda", out;
    for (int i = 1; i < argc; ++i) {
        auto need = [&](const char* f) { if (i + 1 >= argc) { std::fprintf(stderr, "%s needs a value\n", f); std::exit(2); } return argv[++i]; };
        if (!std::strcmp(argv[i], "--width")) W = std::atoi(need("--width"));
        else if (!std::strcmp(argv[i], "--height")) H = std::atoi(need("--height"));
        else if (!std::strcmp(argv[i], "--scene")) scene = need("--scene");
        else if (!std::strcmp(argv[i], "--spheres")) ns = std::atoi(need("--spheres"));
        else if (!std::strcmp(argv[i], "--cubes")) nc = std::atoi(need("--cubes"));
        else if (!std::strcmp(argv[i], "--seed")) seed = std::strtoull(need("--seed"), nullptr, 10);
        else if (!std::strcmp(argv[i], "--mode")) mode = need("--mode");
        else if (!std::strcmp(argv[i], "--device")) device = std::atoi(need("--device"));
        else if (!std::strcmp(argv[i], "--gpus")) gpus = std::atoi(need("--gpus"));
        else if (!std::strcmp(argv[i], "--frames")) frames = std::atoi(need("--frames"));
        else if (!std::strcmp(argv[i], "--tile-cull")) tileCull = std::atoi(need("--tile-cull"));
        else if (!std::strcmp(argv[i], "--out")) out = need("--out");
        else { usage(); return 2; }
    }
    if (W < 16 || H < 16) { std::fprintf(stderr, "resolution too small\n"); return 2; }
    RenderState st(W, H, device, gpus);
    if (st.initStatus() != RTB_OK) {
        std::fprintf(stderr, "rtb_init failed: %s (no CPU fallback)\n", rtb_error_string(st.initStatus()));
        return 1;
    }
    if (tileCull != -1 && st.setTileCull(tileCull) != RTB_OK) {  // unset: the library default (auto)
        std::fprintf(stderr, "--tile-cull %d: %s\n", tileCull, st.lastError.c_str());
        return 2;
    }
    if (scene == "random") st.setRandomScene(ns, nc, seed);
    else st.setScene(std::atoi(scene.c_str()), seed);
    st.setMode(mode == "cuda" ? RenderState::CUDA : (mode == "cpu" ? RenderState::CPU : RenderState::OpenCL));
    for (int f = 0; f < frames; ++f) {
        st.requestRender();
        const int rc = st.update();
        if (rc != RTB_OK) {
            std::fprintf(stderr, "render failed: %s: %s\n", rtb_error_string(rc), st.lastError.c_str());
            return 1;
        }
        const double prims = (double)st.scene.cubes.size() + (double)st.scene.sphereRadius.size();
        std::printf("Time Taken: %llu microseconds (kernel %.3f ms, upload+kernel+download %.3f ms)  "
                    "%.1f Mrays/s  %.3e ray-primitive tests/s (kernel)\n",
                    (unsigned long long)st.timeTaken, st.kernelMs, st.totalMs,
                    st.pixelCount / (st.kernelMs * 1e3), st.pixelCount * prims / (st.kernelMs * 1e-3));
    }
    if (!out.empty()) {
        const unsigned e = st.encodePNG(out.c_str());
        if (e) { std::fprintf(stderr, "encoder error %u\n", e); return 1; }
        std::printf("wrote %s\n", out.c_str());
    }
    return 0;
}
