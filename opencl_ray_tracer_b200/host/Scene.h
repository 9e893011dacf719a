// Scene.h -- the scene members of the reference's render state (RayTrace/states/MainState.h:104-112:
// sphereOrigins / sphereRadius / sphereColours / cubes) and its three builders
// (MainState.cpp:419-462, :464-594, :596-639) plus the seeded synthetic generator used by the
// benchmark configs.
#ifndef RTB_HOST_SCENE_H_
#define RTB_HOST_SCENE_H_

#include <cstdint>
#include <vector>

#include "Cube.h"
#include "Random.h"

namespace rtbh {

const float PI = 3.1415926535f;                                          // misc/Utility.h:19
inline float convertAngleToRadian(float angle) { return (angle * PI / 180.0f); }  // Utility.cpp:343-347

struct Scene {
    std::vector<Vec4> sphereOrigins;
    std::vector<float> sphereRadius;
    std::vector<Vec4> sphereColours;
    std::vector<Cube> cubes;
    // For cubes made by randomCubes(): the 8 draws behind each cube (scale, rz, ry, rx in radians,
    // tx, ty, tz -- the arguments of its scale / rotate x3 / translate calls), so that a checker
    // can rebuild the same cube with the reference's own Cube class.  Empty for scenes 1 and 2.
    std::vector<float> cubeRecipes;

    void clear();
    void createScene1();                 // deterministic
    void createScene2(uint64_t seed);    // fixed geometry, random colours
    void createScene3(uint64_t seed);    // 100 + 100 fully random
    // createScene3's recipe with x in U(0, width-10), y in U(0, height-10), ns spheres, nc cubes.
    void createRandom(int width, int height, int ns, int nc, uint64_t seed);

    // Flattening of MainState.cpp:646-655 (the OpenCL kernel's argument layout).
    void flatten(std::vector<float>* sph_o, std::vector<float>* sph_r, std::vector<float>* sph_c,
                 std::vector<float>* cube_v, std::vector<float>* cube_c) const;

private:
    void randomSpheres(Random& rng, int n, float xmax, float ymax);
    void randomCubes(Random& rng, int n, float xmax, float ymax);
};

// Camera k of the orbit sequence (BASELINE.json config 5): the reference grid and rayDir rotated
// by 3*k degrees about the y axis through (width/2, height/2, -60).  out16 = origin0,dU,dV,rayDir.
void orbitCamera(int width, int height, int k, float out16[16]);

}  // namespace rtbh

#endif
