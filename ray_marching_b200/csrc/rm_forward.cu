// Forward kernels of the sphere-tracing path + their C ABI (include/rm_b200.h).
//
// Thread mapping: every thread owns K = 4 consecutive rays (pixels), so a warp covers 128
// consecutive pixels of a row: coherent rays, 128-bit stores for 1-channel planes and three
// 128-bit stores for 3-channel planes.  The scene tape is staged once per block into shared
// memory and interpreted in registers (rm_eval.cuh).
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <string.h>

#include "../../include/rm_b200.h"
#include "rm_eval.cuh"
#include "rm_host.h"
#include "rm_shade.cuh"

namespace rm {

constexpr int K = 4;            // rays per thread
constexpr int BLOCK = 128;      // threads per block -> 512 rays per block

// ---- small vector I/O helpers: thread t owns elements [4t, 4t+4) ----------------------------
__device__ __forceinline__ void load3(const float* base, int64_t first, int64_t n, float (&x)[K],
                                      float (&y)[K], float (&z)[K]) {
    if (first + K <= n) {
        const float4* p = reinterpret_cast<const float4*>(base + first * 3);
        const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
        x[0] = a.x; y[0] = a.y; z[0] = a.z; x[1] = a.w; y[1] = b.x; z[1] = b.y;
        x[2] = b.z; y[2] = b.w; z[2] = c.x; x[3] = c.y; y[3] = c.z; z[3] = c.w;
    } else {
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const bool ok = first + k < n;
            x[k] = ok ? __ldg(base + (first + k) * 3 + 0) : 0.f;
            y[k] = ok ? __ldg(base + (first + k) * 3 + 1) : 0.f;
            z[k] = ok ? __ldg(base + (first + k) * 3 + 2) : 0.f;
        }
    }
}
__device__ __forceinline__ void store3(float* base, int64_t first, int64_t n, const float (&x)[K],
                                       const float (&y)[K], const float (&z)[K]) {
    if (base == nullptr) return;
    if (first + K <= n) {
        float4* p = reinterpret_cast<float4*>(base + first * 3);
        p[0] = make_float4(x[0], y[0], z[0], x[1]);
        p[1] = make_float4(y[1], z[1], x[2], y[2]);
        p[2] = make_float4(z[2], x[3], y[3], z[3]);
    } else {
#pragma unroll
        for (int k = 0; k < K; ++k)
            if (first + k < n) {
                base[(first + k) * 3 + 0] = x[k];
                base[(first + k) * 3 + 1] = y[k];
                base[(first + k) * 3 + 2] = z[k];
            }
    }
}
__device__ __forceinline__ void store1(float* base, int64_t first, int64_t n, const float (&x)[K]) {
    if (base == nullptr) return;
    if (first + K <= n) {
        *reinterpret_cast<float4*>(base + first) = make_float4(x[0], x[1], x[2], x[3]);
    } else {
#pragma unroll
        for (int k = 0; k < K; ++k)
            if (first + k < n) base[first + k] = x[k];
    }
}

// ---- rm_sdf_eval ----------------------------------------------------------------------------
__global__ void __launch_bounds__(BLOCK)
sdf_eval_kernel(const int32_t* code, const float* par, int n_instr, int n_par,
                const float* __restrict__ pts, int64_t n, float* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem[];
    const TapeView T = stage_tape(code, par, n_instr, n_par, smem);
    const int64_t first = ((int64_t)blockIdx.x * BLOCK + threadIdx.x) * K;
    if (first >= n) return;
    float x[K], y[K], z[K], d[K];
    load3(pts, first, n, x, y, z);
    eval_scene<K, false>(T, x, y, z, d);
    store1(out, first, n, d);
}

// ---- rm_camera_rays -------------------------------------------------------------------------
__device__ __forceinline__ void make_rays(const CamParams& cam, const float* poses, int64_t ray0,
                                          float (&ox)[K], float (&oy)[K], float (&oz)[K],
                                          float (&vx)[K], float (&vy)[K], float (&vz)[K]) {
    const int64_t per_view = (int64_t)cam.width * cam.height;
    const int64_t total = per_view * cam.n_views;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        int64_t r = ray0 + k;
        if (r >= total) r = total - 1;                 // tail lanes recompute the last ray
        const int view = (int)(r / per_view);
        const int pix = (int)(r - (int64_t)view * per_view);
        const int row = pix / cam.width, col = pix - row * cam.width;
        float o[3], v[3], ow[3], vw[3];
        base_ray(cam, row, col, o, v);
        const float* s = poses + (size_t)view * POSE_STRIDE;
        float R[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) R[i] = __ldg(s + i);
        mat_apply(R, o, ow);                           // Q.rotation(base, q) + t  (ray_marching.py:58-62)
        mat_apply(R, v, vw);
        ox[k] = ow[0] + __ldg(s + 9); oy[k] = ow[1] + __ldg(s + 10); oz[k] = ow[2] + __ldg(s + 11);
        vx[k] = vw[0]; vy[k] = vw[1]; vz[k] = vw[2];
    }
}

__global__ void __launch_bounds__(BLOCK)
camera_kernel(CamParams cam, const float* __restrict__ poses, int64_t ray_begin, int64_t n,
              float* __restrict__ origin, float* __restrict__ dirs) {
    const int64_t first = ((int64_t)blockIdx.x * BLOCK + threadIdx.x) * K;
    if (first >= n) return;
    float ox[K], oy[K], oz[K], vx[K], vy[K], vz[K];
    make_rays(cam, poses, ray_begin + first, ox, oy, oz, vx, vy, vz);
    store3(origin, first, n, ox, oy, oz);
    store3(dirs, first, n, vx, vy, vz);
}

// ---- the march loop (ray_marching.py:78-84): p <- f(p) v + p --------------------------------
template <bool FAST>
__device__ __forceinline__ void march_rays(const TapeView& T, int steps, bool early_exit,
                                           float (&px)[K], float (&py)[K], float (&pz)[K],
                                           const float (&vx)[K], const float (&vy)[K],
                                           const float (&vz)[K]) {
    float d[K];
    for (int s = 0; s < steps; ++s) {
        eval_scene<K, FAST>(T, px, py, pz, d);
        bool moved = false;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const float nx = fmaf(d[k], vx[k], px[k]);
            const float ny = fmaf(d[k], vy[k], py[k]);
            const float nz = fmaf(d[k], vz[k], pz[k]);
            moved |= (nx != px[k]) | (ny != py[k]) | (nz != pz[k]);
            px[k] = nx; py[k] = ny; pz[k] = nz;
        }
        // A ray whose position did not change is at a fixed point of the update: all further
        // steps reproduce it bit for bit, so leaving the loop does not change the result.
        if (early_exit && !moved) break;
    }
}

template <bool FAST>
__global__ void __launch_bounds__(BLOCK)
march_kernel(const int32_t* code, const float* par, int n_instr, int n_par,
             const float* __restrict__ origin, const float* __restrict__ dirs, int64_t n, int steps,
             float* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem[];
    const TapeView T = stage_tape(code, par, n_instr, n_par, smem);
    const int64_t first = ((int64_t)blockIdx.x * BLOCK + threadIdx.x) * K;
    if (first >= n) return;
    float px[K], py[K], pz[K], vx[K], vy[K], vz[K];
    load3(origin, first, n, px, py, pz);
    load3(dirs, first, n, vx, vy, vz);
    march_rays<FAST>(T, steps, false, px, py, pz, vx, vy, vz);
    store3(out, first, n, px, py, pz);
}

// ---- stencil: f at p and at the four tetrahedral offsets -------------------------------------
__device__ __forceinline__ void stencil_eval(const TapeView& T, const Stencil& st, bool need_offsets,
                                             const float (&px)[K], const float (&py)[K],
                                             const float (&pz)[K], float (&f)[K],
                                             float (&nx)[K], float (&ny)[K], float (&nz)[K],
                                             float (&lap)[K]) {
    float o[4][K];
    const int first_j = need_offsets ? 0 : 4;
    for (int j = first_j; j < 5; ++j) {               // one call site: j == 4 is the centre
        float x[K], y[K], z[K], d[K];
        const float dx = j < 4 ? st.off[j & 3][0] : 0.f;
        const float dy = j < 4 ? st.off[j & 3][1] : 0.f;
        const float dz = j < 4 ? st.off[j & 3][2] : 0.f;
#pragma unroll
        for (int k = 0; k < K; ++k) { x[k] = px[k] + dx; y[k] = py[k] + dy; z[k] = pz[k] + dz; }
        eval_scene<K, false>(T, x, y, z, d);
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (j == 0) o[0][k] = d[k];
            else if (j == 1) o[1][k] = d[k];
            else if (j == 2) o[2][k] = d[k];
            else if (j == 3) o[3][k] = d[k];
            else f[k] = d[k];
        }
    }
#pragma unroll
    for (int k = 0; k < K; ++k) {
        if (need_offsets) {
            const float ok[4] = {o[0][k], o[1][k], o[2][k], o[3][k]};
            float n3[3];
            stencil_finish(st, ok, f[k], n3, lap[k]);
            nx[k] = n3[0]; ny[k] = n3[1]; nz[k] = n3[2];
        } else {
            nx[k] = ny[k] = nz[k] = lap[k] = 0.f;
        }
    }
}

__global__ void __launch_bounds__(BLOCK)
normals_kernel(const int32_t* code, const float* par, int n_instr, int n_par, Stencil st,
               const float* __restrict__ pts, int64_t n, float* __restrict__ normals,
               float* __restrict__ lap_out) {
    extern __shared__ __align__(16) unsigned char smem[];
    const TapeView T = stage_tape(code, par, n_instr, n_par, smem);
    const int64_t first = ((int64_t)blockIdx.x * BLOCK + threadIdx.x) * K;
    if (first >= n) return;
    float px[K], py[K], pz[K], f[K], nx[K], ny[K], nz[K], lap[K];
    load3(pts, first, n, px, py, pz);
    stencil_eval(T, st, true, px, py, pz, f, nx, ny, nz, lap);
    store3(normals, first, n, nx, ny, nz);
    store1(lap_out, first, n, lap);
}

}  // namespace rm

#include "rm_render.cuh"
#include "rm_abi_forward.cuh"
