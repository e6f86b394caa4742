// Leaf primitive distance functions, evaluated at the frame-local point q.
// Formulas restate scene/primitives.py of the reference (line numbers cited per function).
#pragma once
#include "rm_math.cuh"
#include "rm_tape.h"

namespace rm {

// |q| - r                                                    primitives.py:11-17
template <bool FAST>
__device__ __forceinline__ float prim_sphere(const float4 p0, float x, float y, float z) {
    return Math<FAST>::sqrt(fmaf(x, x, fmaf(y, y, z * z))) - p0.x;
}

// a = |q| - h ; |max(a,0)| + min(max3(a), 0)                 primitives.py:25-33
template <bool FAST>
__device__ __forceinline__ float prim_box(const float4 p0, float x, float y, float z) {
    const float ax = fabsf(x) - p0.x, ay = fabsf(y) - p0.y, az = fabsf(z) - p0.z;
    const float mx = fmaxf(ax, fmaxf(ay, az));
    const float ox = fmaxf(ax, 0.f), oy = fmaxf(ay, 0.f), oz = fmaxf(az, 0.f);
    return Math<FAST>::sqrt(fmaf(ox, ox, fmaf(oy, oy, oz * oz))) + fminf(mx, 0.f);
}

// capsule: p0 = (start, radius), p1 = AB, p2 = AB/|AB|^2     primitives.py:51-61
template <bool FAST>
__device__ __forceinline__ float prim_line(const float4 p0, const float4 p1, const float4 p2,
                                           float x, float y, float z) {
    const float apx = x - p0.x, apy = y - p0.y, apz = z - p0.z;
    float h = fmaf(apx, p2.x, fmaf(apy, p2.y, apz * p2.z));
    // march steps (FAST): FFMA.SAT, no extra instruction; differs from torch.clamp only for a NaN h
    // (saturates to 0), which needs a non-finite position -- the step result is non-finite either way
    h = FAST ? __saturatef(h) : clamp_nan(h, 0.f, 1.f);
    const float wx = fmaf(h, p1.x, -apx), wy = fmaf(h, p1.y, -apy), wz = fmaf(h, p1.z, -apz);
    return Math<FAST>::sqrt(fmaf(wx, wx, fmaf(wy, wy, wz * wz))) - p0.w;
}

// disk in the yz-plane: |(q.x, max(|q.yz| - r, 0))|          primitives.py:69-82
template <bool FAST>
__device__ __forceinline__ float prim_disk(const float4 p0, float x, float y, float z) {
    const float rim = Math<FAST>::sqrt(fmaf(y, y, z * z)) - p0.x;
    const float m = rim > 0.f ? rim : 0.f;
    return Math<FAST>::sqrt(fmaf(x, x, m * m));
}

// torus around y: |(|q.xz| - R1, q.y)| - R2                  primitives.py:91-102
template <bool FAST>
__device__ __forceinline__ float prim_torus(const float4 p0, float x, float y, float z) {
    const float ring = Math<FAST>::sqrt(fmaf(x, x, z * z)) - p0.x;
    return Math<FAST>::sqrt(fmaf(ring, ring, y * y)) - p0.y;
}

}  // namespace rm
