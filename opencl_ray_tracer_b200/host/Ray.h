// Ray.h -- host mirror of RayTrace/Ray.h:6-13 (origin + direction, both 4-vectors).
#ifndef RTB_HOST_RAY_H_
#define RTB_HOST_RAY_H_
#include "vecmath.h"
namespace rtbh {
struct Ray {
    Vec4 origin;
    Vec4 direction;
};
}  // namespace rtbh
#endif
