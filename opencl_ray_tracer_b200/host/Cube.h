// Cube.h -- host mirror of the reference's Cube (RayTrace/Cube.h:7-59, RayTrace/Cube.cpp):
// 36 world-space vertices (12 triangles) + a colour; scale / rotate / translate bake a matrix
// into the vertices.  Same public interface, same vertex order, same float results.
#ifndef RTB_HOST_CUBE_H_
#define RTB_HOST_CUBE_H_

#include <vector>

#include "vecmath.h"

namespace rtbh {

class Cube {
public:
    explicit Cube(const Vec4& colour);
    std::vector<Vec4> getTriangles() const { return triangles_; }
    const std::vector<Vec4>& triangles() const { return triangles_; }  // no-copy accessor
    void rotate(Vec3 newRotation);   // radians; applied as Rz * Ry * Rx like Cube.cpp:53-62
    void scale(Vec3 newScale);
    void translate(Vec3 newTranslation);
    Vec4 getColour() const { return colour_; }

private:
    void apply(const Mat4& m);
    std::vector<Vec4> triangles_;
    Vec4 colour_;
};

}  // namespace rtbh

#endif
