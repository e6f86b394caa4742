// Camera ray generation, the tetrahedral stencil and the per-pixel shaders (device functions).
#pragma once
#include <math.h>

#include "rm_host.h"
#include "rm_math.cuh"

namespace rm {

constexpr int POSE_STRIDE = 8;    // q[4] (w,x,y,z), t[3], pad
constexpr int CMAP_SIZE = 4096;
constexpr float TAU_F = 6.283185307179586f;

struct CamParams {
    int n_views, width, height;
    float focal, half_w, neg_half_h;       // sensor_width/2, -sensor_height/2
    float inv_w, inv_h;
};

// Per-view pose, expanded in registers from the 8-float record.
//   R   = I + 2w[v]x + 2(v v^T - |v|^2 I): the polynomial Q.rotation evaluates (quaternion.py:55-72),
//         valid for non-unit q, so d/dq matches the reference's autograd;
//   fwd = third column of QuaternionToSO3(q) (quaternion.py:114-124; vignette shader.py:62-66).
struct Pose {
    float R[9];
    float t[3];
    float fwd[3];
    float q[4];
};

__device__ __forceinline__ Pose load_pose(const float* poses, int view) {
    Pose p;
    const float4 a = __ldg(reinterpret_cast<const float4*>(poses + (size_t)view * POSE_STRIDE));
    const float4 b = __ldg(reinterpret_cast<const float4*>(poses + (size_t)view * POSE_STRIDE) + 1);
    const float w = a.x, x = a.y, y = a.z, z = a.w;
    p.q[0] = w; p.q[1] = x; p.q[2] = y; p.q[3] = z;
    p.t[0] = b.x; p.t[1] = b.y; p.t[2] = b.z;
    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, xz = x * z, yz = y * z;
    const float wx = w * x, wy = w * y, wz = w * z;
    p.R[0] = 1.f - 2.f * (yy + zz); p.R[1] = 2.f * (xy - wz);       p.R[2] = 2.f * (xz + wy);
    p.R[3] = 2.f * (xy + wz);       p.R[4] = 1.f - 2.f * (xx + zz); p.R[5] = 2.f * (yz - wx);
    p.R[6] = 2.f * (xz - wy);       p.R[7] = 2.f * (yz + wx);       p.R[8] = 1.f - 2.f * (xx + yy);
    p.fwd[0] = 2.f * (wy + xz); p.fwd[1] = 2.f * (yz - wx); p.fwd[2] = w * w - xx - yy + zz;
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
    // correctly rounded, branch-free (rm_math.cuh): n >= focal > 0 is far inside the in-range window
    const float n = sqrt_rn_inline(o[0] * o[0] + o[1] * o[1] + c.focal * c.focal);
    v[0] = div_rn_inline(o[0], n);
    v[1] = div_rn_inline(o[1], n);
    v[2] = div_rn_inline(c.focal, n);
}

// (view, row, col) of a global ray index, clamped to the batch (tail lanes recompute the last ray).
// One divide per THREAD: consecutive rays advance the cursor with compares.  Batches below 2^31
// rays take 32-bit divides (inline); the 64-bit divide is a subroutine call.
struct PixelCursor { int view, row, col; };
__device__ __forceinline__ PixelCursor pixel_cursor(const CamParams& cam, int64_t r) {
    const int64_t per_view = (int64_t)cam.width * cam.height;
    const int64_t total = per_view * cam.n_views;
    if (r >= total) r = total - 1;
    PixelCursor c;
    if (total <= 0x7fffffffLL) {
        const unsigned ru = (unsigned)r, pv = (unsigned)per_view;
        c.view = (int)(ru / pv);
        const unsigned pix = ru - (unsigned)c.view * pv;
        c.row = (int)(pix / (unsigned)cam.width);
        c.col = (int)(pix - (unsigned)c.row * (unsigned)cam.width);
    } else {
        c.view = (int)(r / per_view);
        const int pix = (int)(r - (int64_t)c.view * per_view);
        c.row = pix / cam.width;
        c.col = pix - c.row * cam.width;
    }
    return c;
}
// Global ray index of local ray `local` of a shard (RmRenderArgs: ray_begin, strip_rays, strip_stride).
// One divide per thread (32-bit when it fits); the K rays of a thread lie in one strip (strip % 4 == 0).
__device__ __forceinline__ int64_t shard_ray(int64_t begin, int64_t strip, int64_t stride, int64_t local) {
    if (strip <= 0) return begin + local;
    if ((local | strip) <= 0x7fffffffLL) {
        const unsigned s = (unsigned)local / (unsigned)strip;
        return begin + (int64_t)s * stride + (int64_t)((unsigned)local - s * (unsigned)strip);
    }
    const int64_t s = local / strip;
    return begin + s * stride + (local - s * strip);
}
// Host: validates a shard description against the batch size.
inline const char* check_shard(int64_t begin, int64_t n, int64_t strip, int64_t stride, int64_t total) {
    if (begin < 0 || n < 0) return "negative ray range";
    if (strip < 0 || stride < 0) return "negative strip";
    if (strip == 0) return begin + n > total ? "ray range outside the camera batch" : nullptr;
    if (strip % 4 != 0) return "strip_rays must be a multiple of 4";
    if (stride < strip) return "strip_stride smaller than strip_rays";
    if (n == 0) return nullptr;
    const int64_t last = begin + ((n - 1) / strip) * stride + (n - 1) % strip;
    return last >= total ? "interleaved ray range outside the camera batch" : nullptr;
}

__device__ __forceinline__ void pixel_next(const CamParams& cam, PixelCursor& c) {
    if (++c.col == cam.width) {
        c.col = 0;
        if (++c.row == cam.height) { c.row = 0; ++c.view; }
    }
    if (c.view >= cam.n_views) { c.view = cam.n_views - 1; c.row = cam.height - 1; c.col = cam.width - 1; }
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

// World-space ray of global ray index r (clamped to the batch): origin = R o_b + t, dir = R v_b
// (PinholeCamera.forward, ray_marching.py:57-64).  `pose`/`cur_view` cache the expanded pose across
// consecutive rays of one thread.  ob/vb return the camera-frame ray (backward needs them).
__device__ __forceinline__ void world_ray(const CamParams& cam, const float* poses, int64_t r,
                                          Pose& pose, int& cur_view, float (&ob)[3], float (&vb)[3],
                                          float (&ow)[3], float (&vw)[3]) {
    const PixelCursor c = pixel_cursor(cam, r);        // tail lanes recompute the last ray
    if (c.view != cur_view) { pose = load_pose(poses, c.view); cur_view = c.view; }
    base_ray(cam, c.row, c.col, ob, vb);
    mat_apply(pose.R, ob, ow);
    mat_apply(pose.R, vb, vw);
    ow[0] += pose.t[0]; ow[1] += pose.t[1]; ow[2] += pose.t[2];
}

// Tetrahedral stencil constants (ray_marching.py:94-113): offsets eps*normalize((+-1,0,-s),(0,+-1,s)),
// s = sqrt(1/2); Ainv = inverse of (off[1..3] - off[0]).  With a = eps*sqrt(2/3), b = eps*sqrt(1/3):
// off = (a,0,-b), (-a,0,-b), (0,a,b), (0,-a,b) and
// Ainv = [[-1/(2a),0,0],[-1/(4a),1/(2a),-1/(2a)],[-1/(4b),1/(4b),1/(4b)]]... computed on the host in
// float64 and passed in, so the kernel uses exactly the reference's float32 constants.
struct Stencil {
    float off[5][3];      // sample offsets; [4] = centre (0,0,0)
    float w[5][3];        // column of Ainv applied to (o_j - o_0): w[0] = w[4] = 0, w[j] = Ainv[:, j-1]
    float first[5];       // 1 for j == 0 (defines o_0)
    float in_sum[5];      // 1 for the four offset samples (Laplacian mean)
    float inv[9];         // Ainv, row-major
    float lap_scale;      // 6 / eps^2
    // backward: adjoint of sample j = wb[j] . g_u + lap_w[j] * g_lap*lap_scale + sdf_w[j] * g_sdf, where
    // g_u is the adjoint of u = Ainv (o[1..3] - o_0): wb[0] = -(w[1]+w[2]+w[3]), wb[1..3] = w[1..3]
    float wb[5][3];
    float lap_w[5];       // -1/4 for the four offset samples, +1 for the centre
    float sdf_w[5];       // 1 for the centre (control.py:244)
};

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

// ---- host-side construction of the kernel constants ------------------------------------------
inline int make_cam(const RmCamera* c, CamParams* out) {
    if (c == nullptr || c->n_views <= 0 || c->width <= 0 || c->height <= 0)
        return fail(RM_ERR_INVALID, "camera is null or has a non-positive size");
    out->n_views = c->n_views; out->width = c->width; out->height = c->height;
    out->focal = c->focal_length;
    out->half_w = c->sensor_width / 2;
    out->neg_half_h = -c->sensor_height / 2;
    out->inv_w = 1.0f / (float)c->width;
    out->inv_h = 1.0f / (float)c->height;
    return RM_OK;
}

// SDFNormals.__init__ (ray_marching.py:94-113), float64 on the host, rounded once to float32.
inline void make_stencil(float eps, Stencil* st) {
    const double s = sqrt(0.5);
    const double raw[4][3] = {{1, 0, -s}, {-1, 0, -s}, {0, 1, s}, {0, -1, s}};
    double off[4][3];
    for (int j = 0; j < 4; ++j) {
        const double len = sqrt(raw[j][0] * raw[j][0] + raw[j][1] * raw[j][1] + raw[j][2] * raw[j][2]);
        for (int c = 0; c < 3; ++c) {
            off[j][c] = raw[j][c] / len * (double)eps;
            st->off[j][c] = (float)off[j][c];
        }
    }
    double A[3][3], inv[3][3];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) A[r][c] = off[r + 1][c] - off[0][c];
    const double det = A[0][0] * (A[1][1] * A[2][2] - A[1][2] * A[2][1]) -
                       A[0][1] * (A[1][0] * A[2][2] - A[1][2] * A[2][0]) +
                       A[0][2] * (A[1][0] * A[2][1] - A[1][1] * A[2][0]);
    inv[0][0] = (A[1][1] * A[2][2] - A[1][2] * A[2][1]) / det;
    inv[0][1] = (A[0][2] * A[2][1] - A[0][1] * A[2][2]) / det;
    inv[0][2] = (A[0][1] * A[1][2] - A[0][2] * A[1][1]) / det;
    inv[1][0] = (A[1][2] * A[2][0] - A[1][0] * A[2][2]) / det;
    inv[1][1] = (A[0][0] * A[2][2] - A[0][2] * A[2][0]) / det;
    inv[1][2] = (A[0][2] * A[1][0] - A[0][0] * A[1][2]) / det;
    inv[2][0] = (A[1][0] * A[2][1] - A[1][1] * A[2][0]) / det;
    inv[2][1] = (A[0][1] * A[2][0] - A[0][0] * A[2][1]) / det;
    inv[2][2] = (A[0][0] * A[1][1] - A[0][1] * A[1][0]) / det;
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) st->inv[r * 3 + c] = (float)inv[r][c];
    for (int j = 0; j < 5; ++j) {
        st->lap_w[j] = j < 4 ? -0.25f : 1.f;
        st->sdf_w[j] = j == 4 ? 1.f : 0.f;
        for (int c = 0; c < 3; ++c)
            st->wb[j][c] = (j >= 1 && j <= 3) ? (float)inv[c][j - 1]
                         : (j == 0 ? -(float)(inv[c][0] + inv[c][1] + inv[c][2]) : 0.f);
        st->first[j] = j == 0 ? 1.f : 0.f;
        st->in_sum[j] = j < 4 ? 1.f : 0.f;
        for (int c = 0; c < 3; ++c) {
            st->w[j][c] = (j >= 1 && j <= 3) ? (float)inv[c][j - 1] : 0.f;
            if (j == 4) st->off[j][c] = 0.f;
        }
    }
    st->lap_scale = (float)(6.0 / ((double)eps * (double)eps));
}

}  // namespace rm
