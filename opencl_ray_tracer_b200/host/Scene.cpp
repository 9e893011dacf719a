#include "Scene.h"

#include <cmath>

namespace rtbh {

namespace {

struct Rot { float x, y, z; };  // radians, one Cube::rotate call

Rot degX(float d) { return {convertAngleToRadian(d), 0.0f, 0.0f}; }
Rot degY(float d) { return {0.0f, convertAngleToRadian(d), 0.0f}; }
Rot degZ(float d) { return {0.0f, 0.0f, convertAngleToRadian(d)}; }

Cube makeCube(const Vec4& colour, float s, std::initializer_list<Rot> rots, Vec3 t) {
    Cube c(colour);
    c.scale({s, s, s});
    for (const Rot& r : rots) c.rotate({r.x, r.y, r.z});
    c.translate(t);
    return c;
}

Vec4 randomColour(Random& rng) {
    const float r = rng.getFloat(0.05f, 1.0f);
    const float g = rng.getFloat(0.05f, 1.0f);
    const float b = rng.getFloat(0.05f, 1.0f);
    return {r, g, b, 255.0f};
}

// the four cubes scenes 1 and 2 share (MainState.cpp:435-461, :504-530)
void addBaseCubes(std::vector<Cube>& cubes) {
    cubes.push_back(makeCube({1.f, 1.f, 0.f, 255.f}, 40.f, {degZ(30.f), degY(30.f)}, {70.f, 60.f, -60.f}));
    cubes.push_back(makeCube({0.f, 1.f, 1.f, 255.f}, 30.f, {degZ(80.f), degY(250.f)}, {150.f, 60.f, -70.f}));
    cubes.push_back(makeCube({0.f, 0.f, 1.f, 255.f}, 10.f, {degZ(160.f), degX(210.f)}, {150.f, 400.f, -40.f}));
    cubes.push_back(makeCube({1.f, 0.f, 0.f, 255.f}, 50.f, {degZ(80.f), degY(250.f)}, {450.f, 200.f, -80.f}));
}

}  // namespace

void Scene::clear() {
    sphereOrigins.clear();
    sphereRadius.clear();
    sphereColours.clear();
    cubes.clear();
    cubeRecipes.clear();
}

void Scene::createScene1() {  // MainState.cpp:419-462
    sphereOrigins = {{300.f, 250.f, -85.f, 1.f}, {500.f, 250.f, -85.f, 1.f}};
    sphereRadius = {50.f, 30.f};
    sphereColours = {{0.f, 1.f, 1.f, 255.f}, {1.f, 0.f, 1.f, 255.f}};
    addBaseCubes(cubes);
}

void Scene::createScene2(uint64_t seed) {  // MainState.cpp:464-594
    Random rng(seed);
    const float pos[8][3] = {{100, 150, -85}, {300, 400, -65}, {350, 150, -85}, {200, 250, -85},
                             {200, 350, -45}, {600, 450, -125}, {20, 450, -64}, {620, 250, -115}};
    const float rad[8] = {50, 30, 15, 25, 20, 42, 42, 32};
    for (int i = 0; i < 8; ++i) {
        sphereOrigins.push_back({pos[i][0], pos[i][1], pos[i][2], 1.f});
        sphereRadius.push_back(rad[i]);
    }
    for (int i = 0; i < 8; ++i) sphereColours.push_back(randomColour(rng));
    addBaseCubes(cubes);
    Vec4 c;
    c = randomColour(rng);
    cubes.push_back(makeCube(c, 30.f, {degX(170.f), degY(150.f)}, {450.f, 400.f, -60.f}));
    c = randomColour(rng);
    cubes.push_back(makeCube(c, 50.f, {degZ(80.f), degX(350.f)}, {50.f, 300.f, -100.f}));
    c = randomColour(rng);
    cubes.push_back(makeCube(c, 70.f, {degX(160.f), degY(250.f)}, {530.f, 300.f, -100.f}));
    c = randomColour(rng);
    cubes.push_back(makeCube(c, 25.f, {degZ(190.f), degY(140.f)}, {230.f, 150.f, -40.f}));
    c = randomColour(rng);
    // the second rotation mixes 150 degrees about x with a literal 9.9 rad about z (:574)
    cubes.push_back(makeCube(c, 50.f, {degY(130.f), Rot{convertAngleToRadian(150.f), 0.f, 9.9f}, degZ(50.f)},
                             {510.f, 50.f, -90.f}));
    c = randomColour(rng);
    cubes.push_back(makeCube(c, 24.f, {degZ(280.f), degY(20.f)}, {350.f, 340.f, -40.f}));
}

void Scene::randomSpheres(Random& rng, int n, float xmax, float ymax) {  // MainState.cpp:599-616
    for (int i = 0; i < n; ++i) {
        const float x = rng.getFloat(0.0f, xmax);
        const float y = rng.getFloat(0.0f, ymax);
        const float z = -rng.getFloat(20.0f, 100.0f);
        sphereOrigins.push_back({x, y, z, 1.0f});
        sphereRadius.push_back(rng.getFloat(5.0f, 30.0f));
        sphereColours.push_back(randomColour(rng));
    }
}

void Scene::randomCubes(Random& rng, int n, float xmax, float ymax) {  // MainState.cpp:618-638
    for (int i = 0; i < n; ++i) {
        Cube cube(randomColour(rng));
        const float s = rng.getFloat(5.0f, 30.0f);
        cube.scale({s, s, s});
        const float rz = convertAngleToRadian(rng.getFloat(0.0f, 359.0f));
        cube.rotate({0.0f, 0.0f, rz});
        const float ry = convertAngleToRadian(rng.getFloat(0.0f, 359.0f));
        cube.rotate({0.0f, ry, 0.0f});
        const float rx = convertAngleToRadian(rng.getFloat(0.0f, 359.0f));
        cube.rotate({rx, 0.0f, 0.0f});
        const float x = rng.getFloat(0.0f, xmax);
        const float y = rng.getFloat(0.0f, ymax);
        const float z = -rng.getFloat(30.0f, 100.0f);
        cube.translate({x, y, z});
        cubes.push_back(cube);
        cubeRecipes.insert(cubeRecipes.end(), {s, rz, ry, rx, x, y, z});
    }
}

void Scene::createScene3(uint64_t seed) {
    Random rng(seed);
    randomSpheres(rng, 100, 630.0f, 470.0f);
    randomCubes(rng, 100, 630.0f, 470.0f);
}

void Scene::createRandom(int width, int height, int ns, int nc, uint64_t seed) {
    Random rng(seed);
    randomSpheres(rng, ns, (float)(width - 10), (float)(height - 10));
    randomCubes(rng, nc, (float)(width - 10), (float)(height - 10));
}

void Scene::flatten(std::vector<float>* so, std::vector<float>* sr, std::vector<float>* sc,
                    std::vector<float>* cv, std::vector<float>* cc) const {
    so->clear(); sr->clear(); sc->clear(); cv->clear(); cc->clear();
    for (size_t i = 0; i < sphereOrigins.size(); ++i) {
        const Vec4 &o = sphereOrigins[i], &c = sphereColours[i];
        so->insert(so->end(), {o.x, o.y, o.z, o.w});
        sr->push_back(sphereRadius[i]);
        sc->insert(sc->end(), {c.x, c.y, c.z, c.w});
    }
    for (const Cube& cube : cubes) {
        for (const Vec4& p : cube.triangles()) cv->insert(cv->end(), {p.x, p.y, p.z, p.w});
        const Vec4 c = cube.getColour();
        cc->insert(cc->end(), {c.x, c.y, c.z, c.w});
    }
}

void orbitCamera(int width, int height, int k, float out[16]) {
    const double th = (3.0 * k) * 3.14159265358979323846 / 180.0;
    const double c = std::cos(th), s = std::sin(th);
    const double px = width * 0.5, py = height * 0.5, pz = -60.0;
    auto rotv = [&](double x, double y, double z, double* o) {  // rotation about +y
        o[0] = c * x + s * z; o[1] = y; o[2] = -s * x + c * z;
    };
    double o0[3], du[3], dv[3], d[3];
    rotv(0.0 - px, 0.0 - py, 0.0 - pz, o0);
    o0[0] += px; o0[1] += py; o0[2] += pz;
    rotv(1, 0, 0, du);
    rotv(0, 1, 0, dv);
    rotv(0, 0, -1, d);
    const double v[16] = {o0[0], o0[1], o0[2], 1.0, du[0], du[1], du[2], 0.0,
                          dv[0], dv[1], dv[2], 0.0, d[0],  d[1],  d[2],  -1.0};
    for (int i = 0; i < 16; ++i) out[i] = (float)v[i];
}

}  // namespace rtbh
