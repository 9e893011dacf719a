// Random.h -- host mirror of RayTrace/misc/Random.{h,cpp}: getFloat(min,max) = min + u*(max-min)
// with u in [0,1] (Random.cpp:29-42).  The reference seeds libc rand() from the clock
// (main.cpp:41), so its scenes 2/3 are not reproducible; here the uniform comes from a
// self-contained, explicitly seeded PCG32 stream and every scene names its seed.
#ifndef RTB_HOST_RANDOM_H_
#define RTB_HOST_RANDOM_H_

#include <cstdint>

namespace rtbh {

class Random {
public:
    explicit Random(uint64_t seed) { init(seed); }
    void init(uint64_t seed) {
        state_ = 0u;
        inc_ = (seed << 1u) | 1u;
        next();
        state_ += 0x853c49e6748fea9bULL;
        next();
    }
    uint32_t next() {  // PCG-XSH-RR 64/32
        const uint64_t old = state_;
        state_ = old * 6364136223846793005ULL + inc_;
        const uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
        const uint32_t rot = (uint32_t)(old >> 59u);
        return (xorshifted >> rot) | (xorshifted << ((32u - rot) & 31u));
    }
    float getFloat(float min, float max) {
        const float random = (float)(next() >> 8) / 16777215.0f;  // 24 bits -> [0, 1]
        const float diff = max - min;
        const float r = random * diff;
        return min + r;
    }

private:
    uint64_t state_, inc_;
};

}  // namespace rtbh

#endif
