// Camera ray generation, the tetrahedral stencil and the per-pixel shaders (device functions).
#pragma once
#include "rm_math.cuh"

namespace rm {

constexpr int POSE_STRIDE = 20;   // R[9], t[3], fwd[3], q[4], pad
constexpr int CMAP_SIZE = 4096;
constexpr float TAU_F = 6.283185307179586f;

struct CamParams {
    int n_views, width, height;
    float focal, half_w, neg_half_h;       // sensor_width/2, -sensor_height/2
    float inv_w, inv_h;
};

struct Pose {
    float R[9];
    float t[3];
    float fwd[3];
    float q[4];
};

__device__ __forceinline__ Pose load_pose(const float* poses, int view) {
    Pose p;
    const float* s = poses + (size_t)view * POSE_STRIDE;
#pragma unroll
    for (int i = 0; i < 9; ++i) p.R[i] = __ldg(s + i);
#pragma unroll
    for (int i = 0; i < 3; ++i) { p.t[i] = __ldg(s + 9 + i); p.fwd[i] = __ldg(s + 12 + i); }
#pragma unroll
    for (int i = 0; i < 4; ++i) p.q[i] = __ldg(s + 15 + i);
    return p;
}

// Camera-frame ray of pixel (row, col): closed form of F.affine_grid(theta=[[sw/2,0,0],[0,-sh/2,0]],
// align_corners=False) padded with z=0 and normalize(pos - (0,0,-f)) (ray_marching.py:24-50).
__device__ __forceinline__ void base_ray(const CamParams& c, int row, int col, float (&o)[3], float (&v)[3]) {
    const float gx = (float)(2 * col + 1) * c.inv_w - 1.0f;
    const float gy = (float)(2 * row + 1) * c.inv_h - 1.0f;
    o[0] = c.half_w * gx;
    o[1] = c.neg_half_h * gy;
    o[2] = 0.f;
    const float n = __fsqrt_rn(o[0] * o[0] + o[1] * o[1] + c.focal * c.focal);
    v[0] = __fdiv_rn(o[0], n);
    v[1] = __fdiv_rn(o[1], n);
    v[2] = __fdiv_rn(c.focal, n);
}

__device__ __forceinline__ void mat_apply(const float (&R)[9], const float (&x)[3], float (&y)[3]) {
    y[0] = fmaf(R[0], x[0], fmaf(R[1], x[1], R[2] * x[2]));
    y[1] = fmaf(R[3], x[0], fmaf(R[4], x[1], R[5] * x[2]));
    y[2] = fmaf(R[6], x[0], fmaf(R[7], x[1], R[8] * x[2]));
}
__device__ __forceinline__ void mat_apply_t(const float (&R)[9], const float (&x)[3], float (&y)[3]) {
    y[0] = fmaf(R[0], x[0], fmaf(R[3], x[1], R[6] * x[2]));
    y[1] = fmaf(R[1], x[0], fmaf(R[4], x[1], R[7] * x[2]));
    y[2] = fmaf(R[2], x[0], fmaf(R[5], x[1], R[8] * x[2]));
}

// Tetrahedral stencil constants (ray_marching.py:94-113): offsets eps*normalize((+-1,0,-s),(0,+-1,s)),
// s = sqrt(1/2); Ainv = inverse of (off[1..3] - off[0]).  With a = eps*sqrt(2/3), b = eps*sqrt(1/3):
// off = (a,0,-b), (-a,0,-b), (0,a,b), (0,-a,b) and
// Ainv = [[-1/(2a),0,0],[-1/(4a),1/(2a),-1/(2a)],[-1/(4b),1/(4b),1/(4b)]]... computed on the host in
// float64 and passed in, so the kernel uses exactly the reference's float32 constants.
struct Stencil {
    float off[4][3];
    float inv[9];
    float lap_scale;      // 6 / eps^2
};

// n = normalize(Ainv (o[1..3] - o[0]), eps=0); lap = (f - mean(o)) * 6/eps^2   (ray_marching.py:115-125)
__device__ __forceinline__ void stencil_finish(const Stencil& st, const float (&o)[4], float f,
                                               float (&n)[3], float& lap) {
    const float d0 = o[1] - o[0], d1 = o[2] - o[0], d2 = o[3] - o[0];
    const float nx = (st.inv[0] * d0 + st.inv[1] * d1) + st.inv[2] * d2;
    const float ny = (st.inv[3] * d0 + st.inv[4] * d1) + st.inv[5] * d2;
    const float nz = (st.inv[6] * d0 + st.inv[7] * d1) + st.inv[8] * d2;
    const float len = __fsqrt_rn(nx * nx + ny * ny + nz * nz);
    n[0] = __fdiv_rn(nx, len);      // 0/0 = NaN for missed rays, as F.normalize(eps=0) gives
    n[1] = __fdiv_rn(ny, len);
    n[2] = __fdiv_rn(nz, len);
    const float mean = (((o[0] + o[1]) + o[2]) + o[3]) * 0.25f;
    lap = (f - mean) * st.lap_scale;
}

// ---- shaders (rendering/shader.py) ----------------------------------------------------------
__device__ __forceinline__ float shade_lambertian(const float (&v)[3], const float (&n)[3]) {  // :16-20
    return clamp_nan(-(v[0] * n[0] + v[1] * n[1] + v[2] * n[2]), 0.f, 1.f);
}
__device__ __forceinline__ float shade_log_clamped(float x) {        // :29-33, :46-50 (pre-normalisation)
    return logf(x < 1e-2f ? 1e-2f : x);
}
__device__ __forceinline__ float shade_vignette(const float (&v)[3], const float (&fwd)[3]) {  // :62-66
    const float c = v[0] * fwd[0] + v[1] * fwd[1] + v[2] * fwd[2];
    return c * c * c;
}
__device__ __forceinline__ float gamma233(float x) {                 // .pow(1/2.33)
    return powf(x, 1.0f / 2.33f);
}
// ((l - min)/(max - min))^(1/2.33)                                    :34-38, :51-55
__device__ __forceinline__ float finish_lognorm(float l, float mn, float mx) {
    return gamma233(__fdiv_rn(l - mn, mx - mn));
}
// clamp((1 - L/max|L|)/2, 0, 1)^(1/2.33)                              :81-89
__device__ __forceinline__ float finish_laplacian(float L, float mxabs) {
    return gamma233(clamp_nan((__fdiv_rn(L, mxabs) * -1.f + 1.f) * 0.5f, 0.f, 1.f));
}
// domain_colouring(real, imag): |(re,im)| * cmap[floor((atan2(im,re)/tau + 1/2) * degree * 4096) mod 4096]  :92-118
__device__ __forceinline__ void domain_colour(float re, float im, const float* cmap, int degree, float (&rgb)[3]) {
    const float ang = (atan2f(im, re) / TAU_F + 0.5f) * (float)degree * (float)CMAP_SIZE;
    const float fl = floorf(ang);
    long long idx = (fl == fl) ? (long long)fl : 0;
    idx %= CMAP_SIZE;
    if (idx < 0) idx += CMAP_SIZE;
    const float bright = __fsqrt_rn(re * re + im * im);
#pragma unroll
    for (int c = 0; c < 3; ++c) rgb[c] = bright * __ldg(cmap + idx * 3 + c);
}
// TangentShader (:125-150): T = rot(n - (n.v) v, conj(q_cam)) = R^T (...); colour (T.x, T.y)
__device__ __forceinline__ void shade_tangent(const Pose& cam, const float (&v)[3], const float (&n)[3],
                                              const float* cmap, int degree, float (&rgb)[3]) {
    const float nv = n[0] * v[0] + n[1] * v[1] + n[2] * v[2];
    const float a[3] = {n[0] - nv * v[0], n[1] - nv * v[1], n[2] - nv * v[2]};
    float s[3];
    mat_apply_t(cam.R, a, s);
    domain_colour(s[0], s[1], cmap, degree, rgb);
}
// SpinShader (:157-171): (a, bcd) = (0,n) (x) conj(q_cam); re = a^2-|bcd|^2, im = 2a|bcd|;
// the reference passes (im, re) as (real, imag).
__device__ __forceinline__ void shade_spin(const Pose& cam, const float (&n)[3], const float* cmap,
                                           int degree, float (&rgb)[3]) {
    const float w = cam.q[0], x = -cam.q[1], y = -cam.q[2], z = -cam.q[3];   // conj(q_cam)
    // Hamilton product (0, n) (x) (w, x, y, z)
    const float a = -(n[0] * x + n[1] * y + n[2] * z);
    const float b = n[0] * w + n[1] * z - n[2] * y;
    const float c = n[1] * w + n[2] * x - n[0] * z;
    const float d = n[2] * w + n[0] * y - n[1] * x;
    const float m2 = b * b + c * c + d * d;
    const float re = a * a - m2;
    const float im = __fsqrt_rn(m2) * a * 2.f;
    domain_colour(im, re, cmap, degree, rgb);
}

}  // namespace rm
