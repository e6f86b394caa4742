// Forward kernels of the sphere-tracing path + their C ABI (include/rm_b200.h).
//
// Thread mapping: every thread owns K consecutive rays (pixels), so a warp covers 32*K
// consecutive pixels of a row: coherent rays, 128-bit stores for 1-channel planes and three
// 128-bit stores per 4 rays for 3-channel planes.  K = 4 rays x 256 threads with the register
// budget capped at 64 (4 resident blocks = 32 warps per SM) measured best on B200: the path is
// bound by the latency of the decode -> parameter load -> MUFU chains, which resident warps hide
// (profiles/r01: K=4/154 regs 0.65 ms, K=8/229 regs 0.60 ms, K=2/64 regs 0.59 ms, K=4/64 regs
// 0.45 ms on config 2).  The tape's instruction words arrive in the
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

#ifndef RM_K
#define RM_K 4
#endif
#ifndef RM_BLOCK
#define RM_BLOCK 256
#endif
#ifndef RM_MINB
#define RM_MINB 4
#endif

#ifndef RM_SPEC_UNROLL
#define RM_SPEC_UNROLL 2        // march-loop unrolling of the tape-specialised kernel
#endif
#ifndef RM_SFAST
#define RM_SFAST 1              // 1: frames without Laplacian / reported f / G-buffer take march_stencil_fast
#endif
#ifndef RM_STEP_SYNC
#define RM_STEP_SYNC 0          // 1: block barrier before every march step (see march_rays)
#endif
#ifndef RM_PERSIST
#define RM_PERSIST 0            // 1: persistent balanced tiling in render_kernel (measured slower: profiles/r01/variants.md)
#endif

namespace rm {

constexpr int SPEC_UNROLL = RM_SPEC_UNROLL;
constexpr int K = RM_K;         // rays per thread (multiple of 4: float4 plane I/O)
constexpr int BLOCK = RM_BLOCK; // threads per block

// ---- small vector I/O helpers: thread t owns elements [K t, K t + K) ------------------------
// K = 4g: 128-bit accesses; K = 2: 64-bit accesses (still fully coalesced across the warp).
__device__ __forceinline__ void load3(const float* base, int64_t first, int64_t n, float (&x)[K],
                                      float (&y)[K], float (&z)[K]) {
    if (first + K <= n) {
        if constexpr (K % 4 == 0) {
#pragma unroll
            for (int g = 0; g < K; g += 4) {
                const float4* p = reinterpret_cast<const float4*>(base + (first + g) * 3);
                const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
                x[g] = a.x; y[g] = a.y; z[g] = a.z; x[g + 1] = a.w; y[g + 1] = b.x; z[g + 1] = b.y;
                x[g + 2] = b.z; y[g + 2] = b.w; z[g + 2] = c.x; x[g + 3] = c.y; y[g + 3] = c.z; z[g + 3] = c.w;
            }
        } else {
            static_assert(K % 4 == 0 || K == 2, "K must be 2 or a multiple of 4");
            const float2* p = reinterpret_cast<const float2*>(base + first * 3);
            const float2 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
            x[0] = a.x; y[0] = a.y; z[0] = b.x; x[K - 1] = b.y; y[K - 1] = c.x; z[K - 1] = c.y;
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
        if constexpr (K % 4 == 0) {
#pragma unroll
            for (int g = 0; g < K; g += 4) {
                float4* p = reinterpret_cast<float4*>(base + (first + g) * 3);
                p[0] = make_float4(x[g], y[g], z[g], x[g + 1]);
                p[1] = make_float4(y[g + 1], z[g + 1], x[g + 2], y[g + 2]);
                p[2] = make_float4(z[g + 2], x[g + 3], y[g + 3], z[g + 3]);
            }
        } else {
            float2* p = reinterpret_cast<float2*>(base + first * 3);
            p[0] = make_float2(x[0], y[0]); p[1] = make_float2(z[0], x[K - 1]); p[2] = make_float2(y[K - 1], z[K - 1]);
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
        if constexpr (K % 4 == 0) {
#pragma unroll
            for (int g = 0; g < K; g += 4)
                *reinterpret_cast<float4*>(base + first + g) = make_float4(x[g], x[g + 1], x[g + 2], x[g + 3]);
        } else {
            *reinterpret_cast<float2*>(base + first) = make_float2(x[0], x[K - 1]);
        }
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
    PixelCursor c = pixel_cursor(cam, ray0);
    int cur_view = c.view;
    Pose pose = load_pose(poses, c.view);
#pragma unroll
    for (int k = 0; k < K; ++k) {
        float ob[3], vb[3], ow[3], vw[3];
        if (c.view != cur_view) { pose = load_pose(poses, c.view); cur_view = c.view; }
        base_ray(cam, c.row, c.col, ob, vb);
        mat_apply(pose.R, ob, ow);
        mat_apply(pose.R, vb, vw);
        ox[k] = ow[0] + pose.t[0]; oy[k] = ow[1] + pose.t[1]; oz[k] = ow[2] + pose.t[2];
        vx[k] = vw[0]; vy[k] = vw[1]; vz[k] = vw[2];
        pixel_next(cam, c);
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

// ---- rm_camera_update: the camera-state integration of EventAggregator.get_state ------------
// control.py:148-165, for every view:
//   t <- t + Q.rotation(translation_input * t_sens, q)                    (uses the OLD q)
//   q <- normalize(Q.multiply(q, Q.to_versor(orientation_input * r_sens)), eps = 0)
// `inputs` holds n_frames x n_views x 6 floats (translation input xyz, orientation input
// yz/zx/xy); frame `*counter % n_frames` is applied and the device-side counter advanced, so a
// scripted camera path replays inside a CUDA graph without any host work per frame.
__global__ void camera_update_kernel(float* __restrict__ poses, const float* __restrict__ inputs,
                                     int n_views, int n_frames, float t_sens, float r_sens,
                                     int* __restrict__ counter) {
    const int frame = counter != nullptr ? (int)((unsigned)*counter % (unsigned)n_frames) : 0;
    for (int v = threadIdx.x; v < n_views; v += blockDim.x) {
        const float* in = inputs + ((size_t)frame * n_views + v) * 6;
        float* ps = poses + (size_t)v * POSE_STRIDE;
        const float w = ps[0], x = ps[1], y = ps[2], z = ps[3];
        // Q.rotation (quaternion.py:55-72): t = 2 (q_v x V), V' = V + w t + q_v x t
        const float ux = in[0] * t_sens, uy = in[1] * t_sens, uz = in[2] * t_sens;
        const float tx = 2.f * (y * uz - z * uy), ty = 2.f * (z * ux - x * uz), tz = 2.f * (x * uy - y * ux);
        ps[4] += ux + w * tx + (y * tz - z * ty);
        ps[5] += uy + w * ty + (z * tx - x * tz);
        ps[6] += uz + w * tz + (x * ty - y * tx);
        // Q.to_versor (quaternion.py:75-91): (sqrt(1 - |V|^2), V); NaN when |V| > 1, as in the reference
        const float bx = in[3] * r_sens, by = in[4] * r_sens, bz = in[5] * r_sens;
        const float bw = sqrt_rn_inline(1.f - (bx * bx + by * by + bz * bz));
        // Q.multiply (quaternion.py:24-46), scalar first
        const float nw = w * bw - x * bx - y * by - z * bz;
        const float nx = w * bx + x * bw + y * bz - z * by;
        const float ny = w * by - x * bz + y * bw + z * bx;
        const float nz = w * bz + x * by - y * bx + z * bw;
        const float len = sqrt_rn_inline(nw * nw + nx * nx + ny * ny + nz * nz);
        ps[0] = div_rn_inline(nw, len); ps[1] = div_rn_inline(nx, len);
        ps[2] = div_rn_inline(ny, len); ps[3] = div_rn_inline(nz, len);
    }
    __syncthreads();
    if (threadIdx.x == 0 && counter != nullptr) *counter = *counter + 1;
}

// ---- the march loop (ray_marching.py:78-84): p <- f(p) v + p --------------------------------
// CULL: see eval_scene.  `hint` (largest term of the culled union at the last evaluated point; -inf = none)
// and `move` (distance from that point to the current one) are the caller's: a following stencil can go on
// culling from where the march stopped.  The ray moves |f(p)| |v| per step.
template <int N>
struct CullState {
    float hint[N], move[N];
    __device__ __forceinline__ void reset() {
#pragma unroll
        for (int k = 0; k < N; ++k) { hint[k] = -CUDART_INF_F; move[k] = 0.f; }
    }
};

template <bool FAST, int CAP, bool CULL>
__device__ __forceinline__ void march_rays(const TapeCode<CAP>& code, int n_instr, const float4* spar,
                                           int steps, bool early_exit,
                                           float (&px)[K], float (&py)[K], float (&pz)[K],
                                           const float (&vx)[K], const float (&vy)[K],
                                           const float (&vz)[K], CullState<CULL ? K : 1>& cs) {
    float d[K];
    float vlen[CULL ? K : 1];
#pragma unroll
    for (int k = 0; k < (CULL ? K : 1); ++k)
        vlen[k] = CULL ? sqrt_rn_inline(fmaf(vx[k], vx[k], fmaf(vy[k], vy[k], vz[k] * vz[k]))) * 1.0001f : 0.f;
#ifdef RM_SPEC_INC
    // Tape-specialised build: the evaluation is a few dozen straight-line instructions, so the
    // fixed-point test below (12 compares, the old/new register rotation, a vote) would be ~15 % of
    // the loop even when predicated off.  Launches without early exit take this loop instead.
    if (!CULL && !early_exit) {
#pragma unroll SPEC_UNROLL
        for (int s = 0; s < steps; ++s) {
            eval_scene<K, FAST, CAP, 0>(code, n_instr, spar, px, py, pz, d);
#pragma unroll
            for (int k = 0; k < K; ++k) {
                px[k] = fmaf(d[k], vx[k], px[k]);
                py[k] = fmaf(d[k], vy[k], py[k]);
                pz[k] = fmaf(d[k], vz[k], pz[k]);
            }
        }
        return;
    }
#endif
    for (int s = 0; s < steps; ++s) {                 // one interpreter instance (instruction cache)
#if RM_STEP_SYNC
        // keep the block's warps in phase: they then walk the interpreter's code together and share its
        // instruction-cache lines (block-uniform trip count: legal whenever early_exit is off)
        if (!early_exit) __syncthreads();
#endif
        eval_scene<K, FAST, CAP, CULL ? 2 : 0>(code, n_instr, spar, px, py, pz, d, cs.hint, cs.move);
        bool moved = false;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const float nx = fmaf(d[k], vx[k], px[k]);
            const float ny = fmaf(d[k], vy[k], py[k]);
            const float nz = fmaf(d[k], vz[k], pz[k]);
            if (early_exit) moved |= (nx != px[k]) | (ny != py[k]) | (nz != pz[k]);
            px[k] = nx; py[k] = ny; pz[k] = nz;
            if (CULL) cs.move[CULL ? k : 0] = fabsf(d[k]) * vlen[CULL ? k : 0];
        }
        // A ray whose position did not change is at a fixed point of the update: all further
        // steps reproduce it bit for bit, so leaving the loop (when the whole warp is there) does
        // not change the result.
        if (early_exit && !__any_sync(0xffffffffu, moved)) break;
    }
}

template <bool FAST, int CAP, bool CULL>
__global__ void __launch_bounds__(BLOCK)
march_kernel(const __grid_constant__ TapeCode<CAP> code, const float* par, int n_instr, int n_par,
             const float* __restrict__ origin, const float* __restrict__ dirs, int64_t n, int steps,
             float* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem[];
    const float4* spar = stage_params(par, n_par, smem);
    const int64_t first = ((int64_t)blockIdx.x * BLOCK + threadIdx.x) * K;
    // no early return for tail threads: the culling interpreter votes with a full-warp mask; lanes
    // past the end march zero rays (load3 pads with zeros) and store nothing
    float px[K], py[K], pz[K], vx[K], vy[K], vz[K];
    load3(origin, first, n, px, py, pz);
    load3(dirs, first, n, vx, vy, vz);
    CullState<CULL ? K : 1> cull;
    cull.reset();
    march_rays<FAST, CAP, CULL>(code, n_instr, spar, steps, false, px, py, pz, vx, vy, vz, cull);
    store3(out, first, n, px, py, pz);
}

// ---- stencil: f at the four tetrahedral offsets, then at p ------------------------------------
// One call site of the interpreter; the samples are folded into the normal / Laplacian on the fly
// (o_0, the running sum, and Ainv (o_j - o_0) accumulated column by column, in the reference's
// summation order) so that no per-sample storage and no sample-dependent branch is needed.
// SFAST: the samples use the march's approximate sqrt / exp.  For frames whose outputs do not amplify
// rounding -- no Laplacian (x 6/eps^2 = 2400), no reported f(p), no G-buffer plane: just shaded pixels from
// the normal -- a 2^-22 relative error on f moves the normal by ~1e-5, a hundred times inside the 1e-3 pixel
// tolerance, and the five samples cost what march steps cost instead of ~1.8x (IEEE expf, rounded sqrt).
// (Folding the samples into the march loop, so that ONE interpreter instance serves both, was measured
// slower: the stencil accumulators stay live across the march and spill at 64 registers --
// profiles/r02/variants.md.)
// With SFAST and a culling launch the samples are culled too, with the same rigorous bound as the steps:
// `cull` arrives from the march (last evaluated point p_{S-1}, move = |p_S - p_{S-1}|); the first sample is at
// most eps further, consecutive samples at most 2 eps apart.  The rounded samples (SFAST off) evaluate
// every child.
template <int CAP, bool CULLFLAGS = false, bool SFAST = false>
__device__ __forceinline__ void stencil_eval(const TapeCode<CAP>& code, int n_instr, const float4* spar,
                                             const Stencil& st, bool need_offsets,
                                             const float (&px)[K], const float (&py)[K],
                                             const float (&pz)[K], float (&f)[K],
                                             float (&nx)[K], float (&ny)[K], float (&nz)[K],
                                             float (&lap)[K],
                                             CullState<(CULLFLAGS && SFAST) ? K : 1>* cull = nullptr) {
    constexpr bool SCULL = CULLFLAGS && SFAST;
    const float eps1 = SCULL ? 1.0001f * sqrt_rn_inline(st.off[0][0] * st.off[0][0] + st.off[0][1] * st.off[0][1] +
                                                        st.off[0][2] * st.off[0][2]) : 0.f;
    float o0[K], sum[K];
#pragma unroll
    for (int k = 0; k < K; ++k) { o0[k] = sum[k] = nx[k] = ny[k] = nz[k] = f[k] = 0.f; }
    for (int j = need_offsets ? 0 : 4; j < 5; ++j) {  // j == 4 is the centre (zero offset / weights)
        float x[K], y[K], z[K], d[K];
        const float dx = st.off[j][0], dy = st.off[j][1], dz = st.off[j][2];
#pragma unroll
        for (int k = 0; k < K; ++k) { x[k] = px[k] + dx; y[k] = py[k] + dy; z[k] = pz[k] + dz; }
        if constexpr (SCULL) {
#pragma unroll
            for (int k = 0; k < K; ++k) cull->move[k] = (j == (need_offsets ? 0 : 4) ? cull->move[k] + eps1 : 2.f * eps1);
            eval_scene<K, true, CAP, 2>(code, n_instr, spar, x, y, z, d, cull->hint, cull->move);
        } else {
            eval_scene<K, SFAST, CAP, CULLFLAGS ? 1 : 0>(code, n_instr, spar, x, y, z, d);
        }
        const float wx = st.w[j][0], wy = st.w[j][1], wz = st.w[j][2];
        const float first = st.first[j], in_sum = st.in_sum[j];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            o0[k] = first != 0.f ? d[k] : o0[k];
            const float dj = d[k] - o0[k];             // o_j - o_0 (ray_marching.py:117)
            nx[k] = fmaf(wx, dj, nx[k]);
            ny[k] = fmaf(wy, dj, ny[k]);
            nz[k] = fmaf(wz, dj, nz[k]);
            sum[k] = in_sum != 0.f ? sum[k] + d[k] : sum[k];
            f[k] = d[k];
        }
    }
#pragma unroll
    for (int k = 0; k < K; ++k) {
        // n = normalize(Ainv d, eps=0): 0/0 = NaN for missed rays, as F.normalize gives (:119)
        const float len = sqrt_rn_inline(nx[k] * nx[k] + ny[k] * ny[k] + nz[k] * nz[k]);
        nx[k] = div_rn_inline(nx[k], len);
        ny[k] = div_rn_inline(ny[k], len);
        nz[k] = div_rn_inline(nz[k], len);
        lap[k] = (f[k] - sum[k] * 0.25f) * st.lap_scale;    // (:120-124)
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

#ifdef RM_CULL_STATS
extern "C" int rm_debug_cull_stats(unsigned long long* out8, int reset) {
    cudaDeviceSynchronize();
    if (out8) cudaMemcpyFromSymbol(out8, rm::g_cull_stats, sizeof(unsigned long long) * 8);
    if (reset) { unsigned long long z[8] = {0}; cudaMemcpyToSymbol(rm::g_cull_stats, z, sizeof(z)); }
    return 0;
}
#endif
