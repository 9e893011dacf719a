// host_capi.cpp -- small C surface over the C++ host mirror so that Python (bench.py, tests) uses
// the SAME scene builders / camera / PNG writer as the C++ driver.  Pure host code, no CUDA.
#include <cstring>
#include <vector>

#include "Cube.h"
#include "PngWriter.h"
#include "Scene.h"

using namespace rtbh;

extern "C" {

struct rtbh_scene {
    std::vector<float> so, sr, sc, cv, cc, recipes;
};

// kind 1,2,3 = the reference's createScene1/2/3 (seed feeds the PCG stream of 2 and 3);
// kind 0 = createRandom(width, height, ns, nc, seed).
rtbh_scene* rtbh_scene_create(int kind, unsigned long long seed, int width, int height, int ns, int nc) {
    Scene s;
    if (kind == 1) s.createScene1();
    else if (kind == 2) s.createScene2(seed);
    else if (kind == 3) s.createScene3(seed);
    else if (kind == 0 && width > 10 && height > 10 && ns >= 0 && nc >= 0) s.createRandom(width, height, ns, nc, seed);
    else return nullptr;
    rtbh_scene* h = new rtbh_scene();
    s.flatten(&h->so, &h->sr, &h->sc, &h->cv, &h->cc);
    h->recipes = s.cubeRecipes;
    return h;
}

void rtbh_scene_counts(const rtbh_scene* h, int* ns, int* nc) {
    *ns = (int)h->sr.size();
    *nc = (int)(h->cc.size() / 4);
}

void rtbh_scene_export(const rtbh_scene* h, float* so, float* sr, float* sc, float* cv, float* cc) {
    if (!h->so.empty()) std::memcpy(so, h->so.data(), h->so.size() * 4);
    if (!h->sr.empty()) std::memcpy(sr, h->sr.data(), h->sr.size() * 4);
    if (!h->sc.empty()) std::memcpy(sc, h->sc.data(), h->sc.size() * 4);
    if (!h->cv.empty()) std::memcpy(cv, h->cv.data(), h->cv.size() * 4);
    if (!h->cc.empty()) std::memcpy(cc, h->cc.data(), h->cc.size() * 4);
}

// 7 floats per cube (scale, rz, ry, rx [radians], tx, ty, tz) for randomly generated cubes; returns
// the number of cubes that have a recipe (0 for scenes 1 and 2).  out may be NULL to query.
int rtbh_scene_export_recipes(const rtbh_scene* h, float* out) {
    if (out && !h->recipes.empty()) std::memcpy(out, h->recipes.data(), h->recipes.size() * 4);
    return (int)(h->recipes.size() / 7);
}

void rtbh_scene_free(rtbh_scene* h) { delete h; }

// One cube from a list of operations (0 = scale, 1 = rotate [radians], 2 = translate).
void rtbh_cube_build(const float* colour4, int n_ops, const int* kinds, const float* vec3s, float* verts_out) {
    Cube c({colour4[0], colour4[1], colour4[2], colour4[3]});
    for (int i = 0; i < n_ops; ++i) {
        const Vec3 v = {vec3s[3 * i], vec3s[3 * i + 1], vec3s[3 * i + 2]};
        if (kinds[i] == 0) c.scale(v);
        else if (kinds[i] == 1) c.rotate(v);
        else c.translate(v);
    }
    std::memcpy(verts_out, c.triangles().data(), 36 * sizeof(Vec4));
}

float rtbh_convert_angle_to_radian(float a) { return convertAngleToRadian(a); }

void rtbh_orbit_camera(int width, int height, int k, float* out16) { orbitCamera(width, height, k, out16); }

unsigned rtbh_png_write(const char* filename, const unsigned char* rgba, unsigned w, unsigned h) {
    return encodePNG(filename, rgba, w, h);
}

}  // extern "C"
