// vecmath.h -- the handful of 4-vector / 4x4-matrix operations the reference's scene code takes
// from glm 0.9.6.1, restated with the SAME floating-point operation order so that scenes built
// here are bit-identical to scenes built by the reference's Cube class (checked against
// oracle/_ref in tests/test_host_scene.py).  Column-major like glm: M.col[c][r].
//   mat*vec    : RayTrace/glm/detail/type_mat4x4.inl:571-617  ((m0*x + m1*y) + (m2*z + m3*w))
//   translate  : RayTrace/glm/gtc/matrix_transform.inl:40-49
//   rotate     : RayTrace/glm/gtc/matrix_transform.inl:51-84
//   scale      : RayTrace/glm/gtc/matrix_transform.inl:117-129
#ifndef RTB_HOST_VECMATH_H_
#define RTB_HOST_VECMATH_H_

#include <cmath>

namespace rtbh {

struct Vec3 { float x, y, z; };

struct Vec4 {
    float x, y, z, w;
    float& operator[](int i) { return (&x)[i]; }
    const float& operator[](int i) const { return (&x)[i]; }
};

inline Vec4 operator*(const Vec4& a, float s) { return {a.x * s, a.y * s, a.z * s, a.w * s}; }
inline Vec4 operator+(const Vec4& a, const Vec4& b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }

struct Mat4 {
    Vec4 col[4];
    static Mat4 identity() {
        return {{{1.f, 0.f, 0.f, 0.f}, {0.f, 1.f, 0.f, 0.f}, {0.f, 0.f, 1.f, 0.f}, {0.f, 0.f, 0.f, 1.f}}};
    }
};

inline Vec4 mul(const Mat4& m, const Vec4& v) {
    const Vec4 add0 = m.col[0] * v.x + m.col[1] * v.y;
    const Vec4 add1 = m.col[2] * v.z + m.col[3] * v.w;
    return add0 + add1;
}

inline Mat4 translate(const Mat4& m, const Vec3& v) {
    Mat4 r = m;
    r.col[3] = m.col[0] * v.x + m.col[1] * v.y + m.col[2] * v.z + m.col[3];
    return r;
}

inline Mat4 scale(const Mat4& m, const Vec3& v) {
    Mat4 r;
    r.col[0] = m.col[0] * v.x;
    r.col[1] = m.col[1] * v.y;
    r.col[2] = m.col[2] * v.z;
    r.col[3] = m.col[3];
    return r;
}

inline Mat4 rotate(const Mat4& m, float angle, const Vec3& v) {
    const float c = std::cos(angle);
    const float s = std::sin(angle);
    // normalize(v) = v * (1 / sqrt(dot(v, v))), dot3 = x*x + y*y + z*z left to right
    const float inv = 1.0f / std::sqrt(v.x * v.x + v.y * v.y + v.z * v.z);
    const float a[3] = {v.x * inv, v.y * inv, v.z * inv};
    const float omc = 1.0f - c;
    const float t[3] = {omc * a[0], omc * a[1], omc * a[2]};
    float R[3][3];
    R[0][0] = c + t[0] * a[0];
    R[0][1] = 0 + t[0] * a[1] + s * a[2];
    R[0][2] = 0 + t[0] * a[2] - s * a[1];
    R[1][0] = 0 + t[1] * a[0] - s * a[2];
    R[1][1] = c + t[1] * a[1];
    R[1][2] = 0 + t[1] * a[2] + s * a[0];
    R[2][0] = 0 + t[2] * a[0] + s * a[1];
    R[2][1] = 0 + t[2] * a[1] - s * a[0];
    R[2][2] = c + t[2] * a[2];
    Mat4 r;
    r.col[0] = m.col[0] * R[0][0] + m.col[1] * R[0][1] + m.col[2] * R[0][2];
    r.col[1] = m.col[0] * R[1][0] + m.col[1] * R[1][1] + m.col[2] * R[1][2];
    r.col[2] = m.col[0] * R[2][0] + m.col[1] * R[2][1] + m.col[2] * R[2][2];
    r.col[3] = m.col[3];
    return r;
}

}  // namespace rtbh

#endif
