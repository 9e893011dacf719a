#include "Cube.h"

#include "../csrc/scene_pack.h"  // rtb::kCubeVertexSigns: the reference's 36-corner order

namespace rtbh {

Cube::Cube(const Vec4& colour) : colour_(colour) {
    triangles_.reserve(36);
    for (int j = 0; j < 36; ++j) {
        const signed char* s = rtb::kCubeVertexSigns[j];
        triangles_.push_back({(float)s[0], (float)s[1], (float)s[2], 1.0f});
    }
}

void Cube::apply(const Mat4& m) {
    for (Vec4& p : triangles_) p = mul(m, p);
}

void Cube::rotate(Vec3 r) {
    Mat4 m = rtbh::rotate(Mat4::identity(), r.z, {0.f, 0.f, 1.f});
    m = rtbh::rotate(m, r.y, {0.f, 1.f, 0.f});
    m = rtbh::rotate(m, r.x, {1.f, 0.f, 0.f});
    apply(m);
}

void Cube::scale(Vec3 s) { apply(rtbh::scale(Mat4::identity(), s)); }

void Cube::translate(Vec3 t) { apply(rtbh::translate(Mat4::identity(), t)); }

}  // namespace rtbh
