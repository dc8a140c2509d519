// The float expressions the mock data model evaluates in place of the reference's Eigen code, in the operation
// order SURVEY 8a P1 / P3 / P4 derives for Eigen (a0 + (a1 + a2) for 3-term products, K * P34 first, true division;
// normalized(): z > 0 ? v / sqrt(z) : v).  TEST INFRASTRUCTURE: written independently of libslamgpu so that
// MapMatcherGpu::self_check / audit have something to disagree with.
#pragma once
#include <cmath>
namespace mockmath {
inline float sum3(float a0, float a1, float a2)
{
    volatile float t = a1 + a2;
    volatile float s = a0 + t;
    return s;
}
inline void normalized3(const float v[3], float out[3])
{
    volatile float a = v[0] * v[0], b = v[1] * v[1], c = v[2] * v[2];
    const float z = sum3(a, b, c);
    if (z > 0.0F) {
        const float n = std::sqrt(z);
        out[0] = v[0] / n; out[1] = v[1] / n; out[2] = v[2] / n;
    } else {
        out[0] = v[0]; out[1] = v[1]; out[2] = v[2];
    }
}
inline float norm3(const float v[3])
{
    volatile float a = v[0] * v[0], b = v[1] * v[1], c = v[2] * v[2];
    return std::sqrt(sum3(a, b, c));
}
} // namespace mockmath
