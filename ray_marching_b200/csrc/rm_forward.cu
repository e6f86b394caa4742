// Forward kernels of the sphere-tracing path + their C ABI (include/rm_b200.h).
//
// Thread mapping: every thread owns K consecutive rays (pixels), so a warp covers 32*K
// consecutive pixels of a row: coherent rays, 128-bit stores for 1-channel planes and three
// 128-bit stores per 4 rays for 3-channel planes.  The tape's instruction words arrive in the
// kernel-parameter block (uniform-datapath decode), its parameters are staged once per block
// into shared memory, and the interpreter (rm_eval.cuh) keeps all ray state in registers.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <string.h>

#include "../../include/rm_b200.h"
#include "rm_eval.cuh"
#include "rm_host.h"
#include "rm_shade.cuh"

namespace rm {

constexpr int K = 4;            // rays per thread (multiple of 4: float4 plane I/O)
constexpr int BLOCK = 128;      // threads per block -> 512 rays per block

// ---- small vector I/O helpers: thread t owns elements [K t, K t + K) ------------------------
__device__ __forceinline__ void load3(const float* base, int64_t first, int64_t n, float (&x)[K],
                                      float (&y)[K], float (&z)[K]) {
    if (first + K <= n) {
#pragma unroll
        for (int g = 0; g < K; g += 4) {
            const float4* p = reinterpret_cast<const float4*>(base + (first + g) * 3);
            const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
            x[g] = a.x; y[g] = a.y; z[g] = a.z; x[g + 1] = a.w; y[g + 1] = b.x; z[g + 1] = b.y;
            x[g + 2] = b.z; y[g + 2] = b.w; z[g + 2] = c.x; x[g + 3] = c.y; y[g + 3] = c.z; z[g + 3] = c.w;
        }
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
#pragma unroll
        for (int g = 0; g < K; g += 4) {
            float4* p = reinterpret_cast<float4*>(base + (first + g) * 3);
            p[0] = make_float4(x[g], y[g], z[g], x[g + 1]);
            p[1] = make_float4(y[g + 1], z[g + 1], x[g + 2], y[g + 2]);
            p[2] = make_float4(z[g + 2], x[g + 3], y[g + 3], z[g + 3]);
        }
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
#pragma unroll
        for (int g = 0; g < K; g += 4)
            *reinterpret_cast<float4*>(base + first + g) = make_float4(x[g], x[g + 1], x[g + 2], x[g + 3]);
    } else {
#pragma unroll
        for (int k = 0; k < K; ++k)
            if (first + k < n) base[first + k] = x[k];
    }
}

// ---- rm_sdf_eval ----------------------------------------------------------------------------
template <int CAP>
__global__ void __launch_bounds__(BLOCK)
sdf_eval_kernel(const __grid_constant__ TapeCode<CAP> code, const float* par, int n_instr, int n_par,
                const float* __restrict__ pts, int64_t n, float* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem[];
    const float4* spar = stage_params(par, n_par, smem);
    const int64_t first = ((int64_t)blockIdx.x * BLOCK + threadIdx.x) * K;
    if (first >= n) return;
    float x[K], y[K], z[K], d[K];
    load3(pts, first, n, x, y, z);
    eval_scene<K, false, CAP>(code, n_instr, spar, x, y, z, d);
    store1(out, first, n, d);
}

// ---- rm_camera_rays -------------------------------------------------------------------------
// World-space rays of K consecutive pixels starting at global ray index ray0; `base` optionally
// returns the camera-frame ray (needed by the backward kernel).
__device__ __forceinline__ void make_rays(const CamParams& cam, const float* poses, int64_t ray0,
                                          float (&ox)[K], float (&oy)[K], float (&oz)[K],
                                          float (&vx)[K], float (&vy)[K], float (&vz)[K]) {
    int cur_view = -1;
    Pose pose;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        float ob[3], vb[3], ow[3], vw[3];
        world_ray(cam, poses, ray0 + k, pose, cur_view, ob, vb, ow, vw);
        ox[k] = ow[0]; oy[k] = ow[1]; oz[k] = ow[2];
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
template <bool FAST, int CAP>
__device__ __forceinline__ void march_rays(const TapeCode<CAP>& code, int n_instr, const float4* spar,
                                           int steps, bool early_exit,
                                           float (&px)[K], float (&py)[K], float (&pz)[K],
                                           const float (&vx)[K], const float (&vy)[K],
                                           const float (&vz)[K]) {
    float d[K];
    if (!early_exit) {
        for (int s = 0; s < steps; ++s) {
            eval_scene<K, FAST, CAP>(code, n_instr, spar, px, py, pz, d);
#pragma unroll
            for (int k = 0; k < K; ++k) {
                px[k] = fmaf(d[k], vx[k], px[k]);
                py[k] = fmaf(d[k], vy[k], py[k]);
                pz[k] = fmaf(d[k], vz[k], pz[k]);
            }
        }
        return;
    }
    for (int s = 0; s < steps; ++s) {
        eval_scene<K, FAST, CAP>(code, n_instr, spar, px, py, pz, d);
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
        if (!__any_sync(0xffffffffu, moved)) break;
    }
}

template <bool FAST, int CAP>
__global__ void __launch_bounds__(BLOCK)
march_kernel(const __grid_constant__ TapeCode<CAP> code, const float* par, int n_instr, int n_par,
             const float* __restrict__ origin, const float* __restrict__ dirs, int64_t n, int steps,
             float* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem[];
    const float4* spar = stage_params(par, n_par, smem);
    const int64_t first = ((int64_t)blockIdx.x * BLOCK + threadIdx.x) * K;
    if (first >= n) return;
    float px[K], py[K], pz[K], vx[K], vy[K], vz[K];
    load3(origin, first, n, px, py, pz);
    load3(dirs, first, n, vx, vy, vz);
    march_rays<FAST, CAP>(code, n_instr, spar, steps, false, px, py, pz, vx, vy, vz);
    store3(out, first, n, px, py, pz);
}

// ---- stencil: f at p and at the four tetrahedral offsets -------------------------------------
template <int CAP>
__device__ __forceinline__ void stencil_eval(const TapeCode<CAP>& code, int n_instr, const float4* spar,
                                             const Stencil& st, bool need_offsets,
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
        eval_scene<K, false, CAP>(code, n_instr, spar, x, y, z, d);
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

template <int CAP>
__global__ void __launch_bounds__(BLOCK)
normals_kernel(const __grid_constant__ TapeCode<CAP> code, const float* par, int n_instr, int n_par,
               const __grid_constant__ Stencil st, const float* __restrict__ pts, int64_t n,
               float* __restrict__ normals, float* __restrict__ lap_out) {
    extern __shared__ __align__(16) unsigned char smem[];
    const float4* spar = stage_params(par, n_par, smem);
    const int64_t first = ((int64_t)blockIdx.x * BLOCK + threadIdx.x) * K;
    if (first >= n) return;
    float px[K], py[K], pz[K], f[K], nx[K], ny[K], nz[K], lap[K];
    load3(pts, first, n, px, py, pz);
    stencil_eval<CAP>(code, n_instr, spar, st, true, px, py, pz, f, nx, ny, nz, lap);
    store3(normals, first, n, nx, ny, nz);
    store1(lap_out, first, n, lap);
}

}  // namespace rm

#include "rm_render.cuh"
#include "rm_abi_forward.cuh"
