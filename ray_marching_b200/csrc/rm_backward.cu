// Backward kernels of the sphere-tracing path + their C ABI (include/rm_b200.h).
//
//   rm_sdf_vjp    : VJP of scene(points) -- gradients w.r.t. the points and the tape parameters.
//   rm_march_bwd  : VJP of the march p_{i+1} = p_i + f(p_i) v (ray_marching.py:78-84).  The
//                   reference differentiates the unrolled S-step graph; here the trajectory is
//                   replayed: a forward pass stores p_0..p_{S-1}, then for i = S-1..0
//                       g_i = lambda . v,   d par += g_i df/dpar(p_i),   d v += f(p_i) lambda,
//                       lambda += g_i grad f(p_i)
//                   which is exactly the chain rule of the unrolled graph.
//
// Per-parameter accumulation: warp butterfly -> shared-memory atomics per block -> one global
// atomicAdd per parameter per block (rm_sweep.cuh).
#include <cuda_runtime.h>

#include "../../include/rm_b200.h"
#include "rm_host.h"
#include "rm_sweep.cuh"

namespace rm {

constexpr int KB = 2;            // samples per thread in the backward kernels
constexpr int BLOCK_B = 128;

__device__ __forceinline__ void load3k(const float* base, int64_t first, int64_t n, float (&x)[KB],
                                       float (&y)[KB], float (&z)[KB]) {
#pragma unroll
    for (int k = 0; k < KB; ++k) {
        const bool ok = first + k < n;
        x[k] = ok ? __ldg(base + (first + k) * 3 + 0) : 0.f;
        y[k] = ok ? __ldg(base + (first + k) * 3 + 1) : 0.f;
        z[k] = ok ? __ldg(base + (first + k) * 3 + 2) : 0.f;
    }
}
__device__ __forceinline__ void store3k(float* base, int64_t first, int64_t n, const float (&x)[KB],
                                        const float (&y)[KB], const float (&z)[KB]) {
    if (base == nullptr) return;
#pragma unroll
    for (int k = 0; k < KB; ++k)
        if (first + k < n) {
            base[(first + k) * 3 + 0] = x[k];
            base[(first + k) * 3 + 1] = y[k];
            base[(first + k) * 3 + 2] = z[k];
        }
}

// Shared memory: [tape code | tape par | accumulators (n_par floats)]
__device__ __forceinline__ float* stage_accumulators(unsigned char* smem, int n_instr, int n_par) {
    float* sacc = reinterpret_cast<float*>(smem + (size_t)n_instr * sizeof(int4) + (size_t)n_par * sizeof(float));
    for (int i = threadIdx.x; i < n_par; i += blockDim.x) sacc[i] = 0.f;
    return sacc;
}
__device__ __forceinline__ void flush_accumulators(const float* sacc, float* dpar, int n_par) {
    __syncthreads();
    for (int i = threadIdx.x; i < n_par; i += blockDim.x) {
        const float v = sacc[i];
        if (v != 0.f) atomicAdd(dpar + i, v);
    }
}

// Grid-stride over sample groups so the number of blocks (and global atomics) stays bounded.
__global__ void __launch_bounds__(BLOCK_B)
sdf_vjp_kernel(const int32_t* code, const float* par, int n_instr, int n_par,
               const float* __restrict__ pts, const float* __restrict__ gout, int64_t n,
               float* __restrict__ dpts, float* __restrict__ dpar) {
    extern __shared__ __align__(16) unsigned char smem[];
    float* sacc = stage_accumulators(smem, n_instr, n_par);
    const TapeView T = stage_tape(code, par, n_instr, n_par, smem);     // syncs
    const int64_t groups = (n + KB - 1) / KB;
    const int64_t warp_groups = (groups + 31) / 32 * 32;                // whole warps iterate together
    for (int64_t g = (int64_t)blockIdx.x * BLOCK_B + threadIdx.x; g < warp_groups;
         g += (int64_t)gridDim.x * BLOCK_B) {
        const int64_t first = g * KB;
        float x[KB], y[KB], z[KB], a[KB], f[KB], gx[KB], gy[KB], gz[KB];
        load3k(pts, first, n, x, y, z);
#pragma unroll
        for (int k = 0; k < KB; ++k) a[k] = first + k < n ? __ldg(gout + first + k) : 0.f;
        NodeRecords<KB> rec;
        eval_record<KB>(T, x, y, z, rec, f);
        sweep_vjp<KB>(T, x, y, z, a, rec, sacc, gx, gy, gz);
        store3k(dpts, first, n, gx, gy, gz);
    }
    flush_accumulators(sacc, dpar, n_par);
}

__global__ void __launch_bounds__(BLOCK_B)
march_bwd_kernel(const int32_t* code, const float* par, int n_instr, int n_par,
                 const float* __restrict__ origin, const float* __restrict__ dirs,
                 const float* __restrict__ gpos, int64_t n, int steps, float* __restrict__ traj,
                 float* __restrict__ dorigin, float* __restrict__ ddirs, float* __restrict__ dpar) {
    extern __shared__ __align__(16) unsigned char smem[];
    float* sacc = stage_accumulators(smem, n_instr, n_par);
    const TapeView T = stage_tape(code, par, n_instr, n_par, smem);
    const int64_t groups = (n + KB - 1) / KB;
    const int64_t warp_groups = (groups + 31) / 32 * 32;
    for (int64_t g = (int64_t)blockIdx.x * BLOCK_B + threadIdx.x; g < warp_groups;
         g += (int64_t)gridDim.x * BLOCK_B) {
        const int64_t first = g * KB;
        float px[KB], py[KB], pz[KB], vx[KB], vy[KB], vz[KB];
        load3k(origin, first, n, px, py, pz);
        load3k(dirs, first, n, vx, vy, vz);
        // ---- replay forward, storing the position before every step (IEEE math, as the
        // forward op used when autograd is recording) ----
        for (int s = 0; s < steps; ++s) {
            store3k(traj + (size_t)s * n * 3, first, n, px, py, pz);
            float d[KB];
            eval_scene<KB, false>(T, px, py, pz, d);
#pragma unroll
            for (int k = 0; k < KB; ++k) {
                px[k] = fmaf(d[k], vx[k], px[k]); py[k] = fmaf(d[k], vy[k], py[k]); pz[k] = fmaf(d[k], vz[k], pz[k]);
            }
        }
        // ---- reverse ----
        float lx[KB], ly[KB], lz[KB], dvx[KB], dvy[KB], dvz[KB];
        load3k(gpos, first, n, lx, ly, lz);
#pragma unroll
        for (int k = 0; k < KB; ++k) dvx[k] = dvy[k] = dvz[k] = 0.f;
        for (int s = steps - 1; s >= 0; --s) {
            load3k(traj + (size_t)s * n * 3, first, n, px, py, pz);
            float a[KB], f[KB], gx[KB], gy[KB], gz[KB];
#pragma unroll
            for (int k = 0; k < KB; ++k) a[k] = lx[k] * vx[k] + ly[k] * vy[k] + lz[k] * vz[k];
            NodeRecords<KB> rec;
            eval_record<KB>(T, px, py, pz, rec, f);
            sweep_vjp<KB>(T, px, py, pz, a, rec, sacc, gx, gy, gz);
#pragma unroll
            for (int k = 0; k < KB; ++k) {
                dvx[k] = fmaf(f[k], lx[k], dvx[k]); dvy[k] = fmaf(f[k], ly[k], dvy[k]); dvz[k] = fmaf(f[k], lz[k], dvz[k]);
                lx[k] += gx[k]; ly[k] += gy[k]; lz[k] += gz[k];
            }
        }
        store3k(dorigin, first, n, lx, ly, lz);
        store3k(ddirs, first, n, dvx, dvy, dvz);
    }
    flush_accumulators(sacc, dpar, n_par);
}

inline int check_backward_tape(const RmTape* t) {
    if (int s = check_tape(t)) return s;
    if (t->n_nodes > MAX_NODES)
        return fail(RM_ERR_UNSUPPORTED, "backward supports at most %d union nodes (tape has %d)",
                    MAX_NODES, t->n_nodes);
    return RM_OK;
}
inline unsigned backward_grid(int64_t n) {
    const int64_t want = (n + (int64_t)BLOCK_B * KB - 1) / ((int64_t)BLOCK_B * KB);
    const int64_t cap = 148 * 8;                      // bounded: one global atomic per parameter per block
    return (unsigned)(want < cap ? (want > 0 ? want : 1) : cap);
}

}  // namespace rm

using namespace rm;

extern "C" int rm_sdf_vjp(const RmTape* tape, const float* points, const float* grad_sdf, int64_t n,
                          float* dpoints, float* dpar, void* stream) {
    if (int s = check_backward_tape(tape)) return s;
    if (n < 0 || (dpar == nullptr && tape->n_par > 0) || (n > 0 && (points == nullptr || grad_sdf == nullptr)))
        return fail(RM_ERR_INVALID, "rm_sdf_vjp: null buffer or negative count");
    if (tape->n_par > 0)
        RM_CUDA(cudaMemsetAsync(dpar, 0, (size_t)tape->n_par * sizeof(float), (cudaStream_t)stream));
    if (n == 0) return RM_OK;
    const size_t sm = tape_smem_bytes(tape) + (size_t)tape->n_par * sizeof(float);
    if (int s = prepare_smem(sdf_vjp_kernel, sm)) return s;
    sdf_vjp_kernel<<<backward_grid(n), BLOCK_B, sm, (cudaStream_t)stream>>>(
        tape->code, tape->par, tape->n_instr, tape->n_par, points, grad_sdf, n, dpoints, dpar);
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}

extern "C" int rm_march_bwd(const RmTape* tape, const float* origin, const float* dirs,
                            const float* grad_pos, int64_t n, int32_t steps, float* traj_scratch,
                            float* dorigin, float* ddirs, float* dpar, void* stream) {
    if (int s = check_backward_tape(tape)) return s;
    if (n < 0 || steps < 0 || (dpar == nullptr && tape->n_par > 0) ||
        (n > 0 && (!origin || !dirs || !grad_pos || (steps > 0 && !traj_scratch))))
        return fail(RM_ERR_INVALID, "rm_march_bwd: null buffer or negative count");
    if (tape->n_par > 0)
        RM_CUDA(cudaMemsetAsync(dpar, 0, (size_t)tape->n_par * sizeof(float), (cudaStream_t)stream));
    if (n == 0) return RM_OK;
    const size_t sm = tape_smem_bytes(tape) + (size_t)tape->n_par * sizeof(float);
    if (int s = prepare_smem(march_bwd_kernel, sm)) return s;
    march_bwd_kernel<<<backward_grid(n), BLOCK_B, sm, (cudaStream_t)stream>>>(
        tape->code, tape->par, tape->n_instr, tape->n_par, origin, dirs, grad_pos, n, steps,
        traj_scratch, dorigin, ddirs, dpar);
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}
