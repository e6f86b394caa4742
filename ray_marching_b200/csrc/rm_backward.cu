// Backward kernels of the sphere-tracing path + their C ABI (include/rm_b200.h).
//
//   rm_sdf_vjp    : VJP of scene(points) -- gradients w.r.t. the points and the tape parameters.
//   rm_march_bwd  : VJP of the march p_{i+1} = p_i + f(p_i) v (ray_marching.py:78-84).  The
//                   reference differentiates the unrolled S-step graph; here the trajectory is
//                   replayed: a forward pass stores p_0..p_{S-1}, then for i = S-1..0
//                       g_i = lambda . v,   d par += g_i df/dpar(p_i),   d v += f(p_i) lambda,
//                       lambda += g_i grad f(p_i)
//                   which is exactly the chain rule of the unrolled graph.
//   rm_render_bwd : VJP of the whole G-buffer pipeline of RenderLoop.forward (control.py:239-245):
//                   camera rays -> march -> f(p) -> tetrahedral stencil, in ONE kernel: rays are
//                   regenerated from the pixel index, the march is replayed, the stencil and the
//                   march are reversed, and the adjoints of the ray origin / direction are
//                   contracted with d(R(q) b + t)/d(q, t) into camera-pose gradients.
//
// Per-parameter accumulation: warp butterfly -> shared-memory atomics per block -> one global
// atomicAdd per parameter per block (rm_sweep.cuh).  Camera gradients: per-thread registers over
// the grid-stride loop -> warp shuffle reduction -> one global atomicAdd per warp and component.
#include <cuda_runtime.h>

#include "../../include/rm_b200.h"
#include "rm_host.h"
#include "rm_shade.cuh"
#include <cstdlib>
#include "rm_sweep.cuh"

namespace rm {

#ifndef RM_MINB_BWD
#define RM_MINB_BWD 4
#endif

#ifndef RM_ACC_BYTES
#define RM_ACC_BYTES (48 * 1024)   // shared-memory budget of the gradient accumulator columns per block
#endif

#ifndef RM_BWD_EARLY_BREAK
#define RM_BWD_EARLY_BREAK 1
#endif

#ifndef RM_KB
#define RM_KB 1
#endif
#ifndef RM_LOSS_FAST_REPLAY
#define RM_LOSS_FAST_REPLAY 1     // rm_render_loss_bwd replays the march with approximate sqrt / exp (see forward_pass)
#endif
#ifndef RM_LOSS_PATCHES
#define RM_LOSS_PATCHES 1   // rm_render_loss_bwd: 8 x 4-pixel warp tiles when the ray range allows (render_bwd_body)
#endif
#ifndef RM_LOSS_SPLIT
#define RM_LOSS_SPLIT 1       // long tapes: replay in a kernel of its own (loss_replay_kernel), reverse in another
#endif
#ifndef RM_LOSS_TWO_PHASE
#define RM_LOSS_TWO_PHASE 0   // 1: long tapes take the four-ray replay (render_loss_onfly4_kernel); measured slower, see below
#endif
#ifndef RM_LOSS_ONFLY_MIN_INSTR
#define RM_LOSS_ONFLY_MIN_INSTR 24
#endif
#ifndef RM_LOSS_ONFLY
#define RM_LOSS_ONFLY 1           // rm_render_loss_bwd: node records kept from the replay, leaves re-evaluated by the sweep
#endif
constexpr int KB = RM_KB;        // samples per thread in the backward kernels (1, 2 or 4)
constexpr int BLOCK_B = 128;

__device__ __forceinline__ void load3k(const float* base, int64_t first, int64_t n, float (&x)[KB],
                                       float (&y)[KB], float (&z)[KB]) {
#pragma unroll
    for (int k = 0; k < KB; ++k) {
        const bool ok = base != nullptr && first + k < n;
        x[k] = ok ? __ldg(base + (first + k) * 3 + 0) : 0.f;
        y[k] = ok ? __ldg(base + (first + k) * 3 + 1) : 0.f;
        z[k] = ok ? __ldg(base + (first + k) * 3 + 2) : 0.f;
    }
}
__device__ __forceinline__ void load1k(const float* base, int64_t first, int64_t n, float (&x)[KB]) {
#pragma unroll
    for (int k = 0; k < KB; ++k) x[k] = (base != nullptr && first + k < n) ? __ldg(base + first + k) : 0.f;
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

// Trajectory scratch, indexed by THREAD SLOT (not by ray): a thread reuses its slots on every
// iteration of its grid-stride loop, so the footprint is steps x grid x BLOCK_B x KB x 12 bytes
// (<= ~116 MB, L2-sized) however many rays there are.  A thread writes 24 contiguous bytes per
// step, a warp 768 contiguous bytes.
__device__ __forceinline__ float* traj_slot(float* traj, int s) {
    const size_t n_slots = (size_t)gridDim.x * BLOCK_B;
    const size_t slot = (size_t)blockIdx.x * BLOCK_B + threadIdx.x;
    return traj + ((size_t)s * n_slots + slot) * (KB * 3);
}
__device__ __forceinline__ void traj_store(float* traj, int s, const float (&x)[KB], const float (&y)[KB],
                                           const float (&z)[KB]) {
    if constexpr (KB == 1) {
        float* p = traj_slot(traj, s);
        p[0] = x[0]; p[1] = y[0]; p[2] = z[0];
    } else if constexpr (KB == 2) {
        float2* p = reinterpret_cast<float2*>(traj_slot(traj, s));
        p[0] = make_float2(x[0], y[0]); p[1] = make_float2(z[0], x[KB - 1]); p[2] = make_float2(y[KB - 1], z[KB - 1]);
    } else {
        static_assert(KB == 1 || KB == 2 || KB == 4, "KB must be 1, 2 or 4");
        float4* p = reinterpret_cast<float4*>(traj_slot(traj, s));
        p[0] = make_float4(x[0], y[0], z[0], x[1]);
        p[1] = make_float4(y[1], z[1], x[KB - 2], y[KB - 2]);
        p[2] = make_float4(z[KB - 2], x[KB - 1], y[KB - 1], z[KB - 1]);
    }
}
__device__ __forceinline__ void traj_load(float* traj, int s, float (&x)[KB], float (&y)[KB], float (&z)[KB]) {
    if constexpr (KB == 1) {
        const float* p = traj_slot(traj, s);
        x[0] = p[0]; y[0] = p[1]; z[0] = p[2];
    } else if constexpr (KB == 2) {
        const float2* p = reinterpret_cast<const float2*>(traj_slot(traj, s));
        const float2 a = p[0], b = p[1], c = p[2];
        x[0] = a.x; y[0] = a.y; z[0] = b.x; x[KB - 1] = b.y; y[KB - 1] = c.x; z[KB - 1] = c.y;
    } else {
        const float4* p = reinterpret_cast<const float4*>(traj_slot(traj, s));
        const float4 a = p[0], b = p[1], c = p[2];
        x[0] = a.x; y[0] = a.y; z[0] = a.z; x[1] = a.w; y[1] = b.x; z[1] = b.y;
        x[KB - 2] = b.z; y[KB - 2] = b.w; z[KB - 2] = c.x; x[KB - 1] = c.y; y[KB - 1] = c.z; z[KB - 1] = c.w;
    }
}

// Node-record scratch of the fused loss kernel (RM_LOSS_ONFLY), indexed by thread slot like the trajectory:
// per evaluation of the replay (march steps, then the four offset samples of the stencil) f and, per union
// node, its raw value and winning child -- NR = 1 + 2 n_nodes floats, each a plane over the slots so that a
// warp stores / loads 128 contiguous bytes.  The reverse tasks read them back instead of re-evaluating the
// scene in a recording pass; a smooth union's children are re-evaluated, value-only, by the sweep itself.
__device__ __forceinline__ size_t nrec_index(int eval, int i, int k, int nr) {
    const size_t n_slots = (size_t)gridDim.x * BLOCK_B * KB;
    const size_t slot = ((size_t)blockIdx.x * BLOCK_B + threadIdx.x) * KB + k;
    return ((size_t)eval * nr + i) * n_slots + slot;
}
template <class REC>
__device__ __forceinline__ void nrec_store(float* nrec, int n_nodes, int eval, const REC& rec,
                                           const float (&f)[KB]) {
    const int nr = 1 + 2 * n_nodes;
#pragma unroll
    for (int k = 0; k < KB; ++k) {
        nrec[nrec_index(eval, 0, k, nr)] = f[k];
        for (int node = 0; node < n_nodes; ++node) {
            nrec[nrec_index(eval, 1 + 2 * node, k, nr)] = rec.d[node][k];
            nrec[nrec_index(eval, 2 + 2 * node, k, nr)] = __int_as_float(rec.win[node][k]);
        }
    }
}
// (kf, ksel): the two-phase kernel (render_loss_onfly4_kernel) keeps kf rays per thread slot; field i of ray
// ksel is plane (i * kf + ksel) of a record of nr * kf planes.
template <class REC>
__device__ __forceinline__ void nrec_load(const float* nrec, int n_nodes, const int (&eval)[KB],
                                          REC& rec, float (&f)[KB], int kf = 1, int ksel = 0) {
    const int nr = (1 + 2 * n_nodes) * kf;
#pragma unroll
    for (int k = 0; k < KB; ++k) {
        f[k] = nrec[nrec_index(eval[k], ksel, k, nr)];
        for (int node = 0; node < n_nodes; ++node) {
            rec.d[node][k] = nrec[nrec_index(eval[k], (1 + 2 * node) * kf + ksel, k, nr)];
            rec.win[node][k] = __float_as_int(nrec[nrec_index(eval[k], (2 + 2 * node) * kf + ksel, k, nr)]);
        }
    }
}
// K-generic stores of the two-phase kernel's replay (KB == 1 slots): trajectory [(s * KF + k)][slot][3],
// records [eval][field * KF + k][slot] -- what traj_slot / nrec_load address with (kf, ksel).
template <int KF>
__device__ __forceinline__ void traj_store_k(float* traj, int s, const float (&x)[KF], const float (&y)[KF],
                                             const float (&z)[KF]) {
    const size_t n_slots = (size_t)gridDim.x * BLOCK_B;
    const size_t slot = (size_t)blockIdx.x * BLOCK_B + threadIdx.x;
#pragma unroll
    for (int k = 0; k < KF; ++k) {
        float* p = traj + ((size_t)(s * KF + k) * n_slots + slot) * 3;
        p[0] = x[k]; p[1] = y[k]; p[2] = z[k];
    }
}
template <int KF, class REC>
__device__ __forceinline__ void nrec_store_k(float* nrec, int n_nodes, int eval, const REC& rec,
                                             const float (&f)[KF]) {
    const size_t n_slots = (size_t)gridDim.x * BLOCK_B;
    const size_t slot = (size_t)blockIdx.x * BLOCK_B + threadIdx.x;
    const size_t base = (size_t)eval * (1 + 2 * n_nodes) * KF;
#pragma unroll
    for (int k = 0; k < KF; ++k) {
        nrec[(base + k) * n_slots + slot] = f[k];
        for (int node = 0; node < n_nodes; ++node) {
            nrec[(base + (size_t)(1 + 2 * node) * KF + k) * n_slots + slot] = rec.d[node][k];
            nrec[(base + (size_t)(2 + 2 * node) * KF + k) * n_slots + slot] = __int_as_float(rec.win[node][k]);
        }
    }
}

// Shared memory: [tape par (n_par floats) | accumulator columns], see GradAcc (rm_sweep.cuh).
template <int MODE>
__device__ __forceinline__ GradAcc<MODE> stage_accumulators(unsigned char* smem, int n_par, int acc_shared) {
    GradAcc<MODE> acc;
    acc.base = reinterpret_cast<float*>(smem + (size_t)n_par * sizeof(float));
    acc.n_par = n_par;
    acc.shared = acc_shared != 0;
    const int cols = MODE == ACC_LANES ? BLOCK_B : (acc.shared ? 1 : BLOCK_B / 32);
    for (int i = threadIdx.x; i < n_par * cols; i += BLOCK_B) acc.base[i] = 0.f;
    return acc;
}
// One pass per block: column sums, then one global atomic per parameter (ACC_WARPS) or per
// parameter and warp (ACC_LANES).
template <int MODE>
__device__ __forceinline__ void flush_accumulators(const GradAcc<MODE> acc, float* dpar, int n_par) {
    __syncthreads();
    if (MODE == ACC_LANES) {
        for (int r = 0; r < n_par; ++r) {
            const float v = warp_sum(acc.base[r * BLOCK_B + threadIdx.x]);
            if ((threadIdx.x & 31) == 0 && v != 0.f) atomicAdd(dpar + r, v);
        }
    } else {
        const int cols = acc.shared ? 1 : BLOCK_B / 32;
        for (int i = threadIdx.x; i < n_par; i += BLOCK_B) {
            float v = 0.f;
            for (int w = 0; w < cols; ++w) v += acc.base[w * n_par + i];
            if (v != 0.f) atomicAdd(dpar + i, v);
        }
    }
}

// ---- the two passes every backward kernel is made of ------------------------------------------
// Each pass has exactly ONE call site of the interpreter (forward: eval_scene; reverse:
// eval_record + sweep_vjp), driven by a warp-uniform task counter, so the kernel holds one copy of
// each and its hot code stays instruction-cache resident.

struct StencilAdjoint {          // per-sample inputs of the stencil's reverse tasks
    float gux[KB], guy[KB], guz[KB];   // adjoint of u = Ainv (o[1..3] - o_0)
    float gm[KB];                      // g_lap * 6/eps^2
    float gs[KB];                      // g_sdf
};

// Forward pass: replays the march with IEEE math (storing the position before every step in the
// slot scratch), then -- if `stencil` -- samples the four tetrahedral offsets at p_S and folds them
// into u = Ainv (o_j - o_0) like the forward kernel does.
// conv[k] = first step from which every step length stayed within tol * (1 + |p|_inf) (= steps when
// the ray never settles, or when tol == 0): the tail [conv, steps) is a stationary point of the
// iteration to first order in tol.
// FASTM: approximate sqrt / exp in the replay.  rm_render_loss_bwd replays with it -- the kernel IS the
// forward of its loss, so there is no separately computed trajectory to reproduce, and a 2^-22 relative
// change of the evaluation points moves the gradient by ~1e-6 of its size.  rm_render_bwd / rm_march_bwd keep
// the IEEE replay (it mirrors the IEEE forward kernel that produced the planes).
template <int CAP, bool FASTM = false, bool ONFLY = false, class REC = NodeRecords<KB, CAP>>
__device__ __forceinline__ void forward_pass(const TapeCode<CAP>& code, int n_instr, const float4* spar,
                                             int steps, float tol, const Stencil* st, float* traj,
                                             float (&px)[KB], float (&py)[KB], float (&pz)[KB],
                                             const float (&vx)[KB], const float (&vy)[KB],
                                             const float (&vz)[KB], int (&conv)[KB],
                                             float (&ux)[KB], float (&uy)[KB], float (&uz)[KB],
                                             REC* rec = nullptr, float* nrec = nullptr,
                                             int n_nodes = 0) {
    float o0[KB], prev[KB];                                    // prev: last step length while settled
#pragma unroll
    for (int k = 0; k < KB; ++k) { conv[k] = 0; ux[k] = uy[k] = uz[k] = o0[k] = 0.f; prev[k] = -1.f; }
    const int n_tasks = steps + (st != nullptr ? 4 : 0);
    for (int task = 0; task < n_tasks; ++task) {
        const bool marching = task < steps;
        const int j = marching ? 4 : task - steps;            // stencil sample (4 = zero offset)
        float x[KB], y[KB], z[KB], d[KB];
        float dx = 0.f, dy = 0.f, dz = 0.f;
        if (!marching) { dx = st->off[j][0]; dy = st->off[j][1]; dz = st->off[j][2]; }
        else traj_store(traj, task, px, py, pz);
#pragma unroll
        for (int k = 0; k < KB; ++k) { x[k] = px[k] + dx; y[k] = py[k] + dy; z[k] = pz[k] + dz; }
        // ONE call site for steps and samples (a second, correctly rounded one for the samples costs c4b
        // 32.9 -> 35.3 ms in instruction-cache misses and changes no gradient digit: profiles/r02/variants.md)
        if constexpr (ONFLY) {
            // ... recording the union nodes (rec->c, constant per tape, stays in *rec for the reverse pass)
            eval_scene<KB, FASTM, CAP, 0, 2, REC>(code, n_instr, spar, x, y, z, d, nullptr, nullptr, rec);
            if (nrec != nullptr) nrec_store(nrec, n_nodes, task, *rec, d);
        } else {
            eval_scene<KB, FASTM, CAP>(code, n_instr, spar, x, y, z, d);
        }
        if (marching) {
            bool at_rest = true;
#pragma unroll
            for (int k = 0; k < KB; ++k) {
                const float scale = 1.f + fmaxf(fabsf(px[k]), fmaxf(fabsf(py[k]), fabsf(pz[k])));
                const float ad = fabsf(d[k]);
                const bool within = ad <= tol * scale;                       // false for NaN
                if (!within) conv[k] = task + 1;
                // settled for two steps in a row without growing: a contracting fixed point
                at_rest = at_rest && within && prev[k] >= ad;
                prev[k] = within ? ad : -1.f;
                px[k] = fmaf(d[k], vx[k], px[k]); py[k] = fmaf(d[k], vy[k], py[k]); pz[k] = fmaf(d[k], vz[k], pz[k]);
            }
            // every ray of the warp at rest: the remaining steps stay within the tolerance (they are
            // part of the converged tail by definition), so the replay jumps to the stencil.  With
            // tol = 0 this triggers only when f == 0 exactly twice, where skipping is exact.
#if RM_BWD_EARLY_BREAK
            if (__all_sync(0xffffffffu, at_rest)) task = steps - 1;
#endif
        } else {
            const float wx = st->w[j][0], wy = st->w[j][1], wz = st->w[j][2], first = st->first[j];
#pragma unroll
            for (int k = 0; k < KB; ++k) {
                o0[k] = first != 0.f ? d[k] : o0[k];
                const float dj = d[k] - o0[k];
                ux[k] = fmaf(wx, dj, ux[k]); uy[k] = fmaf(wy, dj, uy[k]); uz[k] = fmaf(wz, dj, uz[k]);
            }
        }
    }
}

// Reverse pass.  On entry l* = upstream adjoint of p_S and dv* = upstream adjoint of the direction;
// on exit l* = adjoint of p_0, dv* = adjoint of the direction.  Tasks, in order:
//   0..4   stencil samples j (offset j, centre for j = 4) with adjoint
//          wb[j].g_u + lap_w[j] g_m + sdf_w[j] g_s            (skipped when sa == nullptr)
//   5, 6   converged tail: the steps [conv, S) all sit at the stationary point q = p_conv, where the
//          recurrence  a = l.v; dpar += a df/dpar; dv += f l; l += a grad f  has the closed form
//          a_j = a_0 c^j, c = 1 + grad f(q).v.  Task 5 sweeps with a_0 (which also yields c), task 6
//          with a_0 (S_n - 1), S_n = sum_{j<n} c^j, n = S - conv: the implicit-surface adjoint of the
//          hit point truncated to the n steps the reference unrolls; O(tol) from the step-by-step sum.
//   7..    exact steps s = max_warp(conv) - 1 .. 0 of the unrolled graph.
// ONFLY (the fused loss kernel): no recording pass -- the union-node records of every replayed evaluation
// come back from the scratch (nrec_load) and the sweep evaluates leaf values itself; the centre sample
// (task 4) is skipped, its adjoint is identically zero for this loss (no f(p), no Laplacian term).
template <int CAP, int MODE, bool FASTM = false, bool ONFLY = false, class REC = NodeRecords<KB, CAP>>
__device__ __forceinline__ void reverse_pass(const TapeCode<CAP>& code, int n_instr, const float4* spar,
                                             int steps, const Stencil* st, const StencilAdjoint* sa,
                                             float* traj, const GradAcc<MODE> sacc, const int (&conv)[KB],
                                             const float (&psx)[KB], const float (&psy)[KB],
                                             const float (&psz)[KB],
                                             const float (&vx)[KB], const float (&vy)[KB],
                                             const float (&vz)[KB], float (&lx)[KB], float (&ly)[KB],
                                             float (&lz)[KB], float (&dvx)[KB], float (&dvy)[KB],
                                             float (&dvz)[KB],
                                             REC* frec = nullptr, const float* nrec = nullptr,
                                             int n_nodes = 0, int kf = 1, int ksel = 0) {
    bool any_tail = false;
    int first_exact = 0;
#pragma unroll
    for (int k = 0; k < KB; ++k) { any_tail |= conv[k] < steps; first_exact = max(first_exact, conv[k]); }
    const bool tail = __any_sync(0xffffffffu, any_tail);
    // warp max through REDUX: the result lives in a uniform register, which keeps the task loop
    // (and the interpreters' decode inside it) on the uniform datapath
    first_exact = __reduce_max_sync(0xffffffffu, first_exact);

    float g1x[KB], g1y[KB], g1z[KB], a0[KB], rest[KB], tsum[KB];   // carried from task 5 to task 6
#pragma unroll
    for (int k = 0; k < KB; ++k) g1x[k] = g1y[k] = g1z[k] = a0[k] = rest[k] = tsum[k] = 0.f;
    int task = sa != nullptr ? 0 : 5;
    if (task == 5 && !tail) task = 7;
    const int n_tasks = 7 + first_exact;
    while (task < n_tasks) {
        float x[KB], y[KB], z[KB], a[KB], f[KB], gx[KB], gy[KB], gz[KB];
        const int s = first_exact - 1 - (task - 7);           // exact step of tasks >= 7
        if (task < 5) {
            const float ox = st->off[task][0], oy = st->off[task][1], oz = st->off[task][2];
            const float w0 = st->wb[task][0], w1 = st->wb[task][1], w2 = st->wb[task][2];
            const float wl = st->lap_w[task], ws = st->sdf_w[task];
#pragma unroll
            for (int k = 0; k < KB; ++k) {
                x[k] = psx[k] + ox; y[k] = psy[k] + oy; z[k] = psz[k] + oz;
                a[k] = fmaf(w0, sa->gux[k], fmaf(w1, sa->guy[k], w2 * sa->guz[k])) + wl * sa->gm[k] + ws * sa->gs[k];
            }
        } else if (task < 7) {
#pragma unroll
            for (int k = 0; k < KB; ++k) {
                // stationary point q = p_conv, read from this lane's own slot (unused when conv == steps)
                const float* slot = traj_slot(traj, (conv[k] < steps ? conv[k] : 0) * kf + ksel) + 3 * k;
                x[k] = slot[0]; y[k] = slot[1]; z[k] = slot[2];
                if (task == 5) a0[k] = conv[k] < steps ? lx[k] * vx[k] + ly[k] * vy[k] + lz[k] * vz[k] : 0.f;
                a[k] = task == 5 ? a0[k] : rest[k];
            }
        } else {
            traj_load(traj, s * kf + ksel, x, y, z);
#pragma unroll
            for (int k = 0; k < KB; ++k)
                a[k] = s < conv[k] ? lx[k] * vx[k] + ly[k] * vy[k] + lz[k] * vz[k] : 0.f;
        }
        if constexpr (ONFLY) {
            int ev[KB];                                        // which replayed evaluation this task revisits
#pragma unroll
            for (int k = 0; k < KB; ++k)
                ev[k] = task < 5 ? steps + task : (task < 7 ? (conv[k] < steps ? conv[k] : 0) : s);
            nrec_load(nrec, n_nodes, ev, *frec, f, kf, ksel);
            sweep_vjp<KB, CAP, MODE, true, FASTM, REC>(code, n_instr, spar, x, y, z, a, *frec, sacc, gx, gy, gz);
        } else {
            NodeRecords<KB, CAP> rec;
            eval_record<KB, CAP, FASTM>(code, n_instr, spar, x, y, z, rec, f);
            sweep_vjp<KB, CAP, MODE>(code, n_instr, spar, x, y, z, a, rec, sacc, gx, gy, gz);
        }
        if (task < 5) {
#pragma unroll
            for (int k = 0; k < KB; ++k) { lx[k] += gx[k]; ly[k] += gy[k]; lz[k] += gz[k]; }
            if (ONFLY && task == 3) task = 4;                  // the centre sample carries no adjoint
            if (task == 4 && !tail) task = 6;                  // no lane has a tail: skip tasks 5, 6
        } else if (task == 5) {
#pragma unroll
            for (int k = 0; k < KB; ++k) {
                const int n = steps - conv[k];
                // c from the sweep: (a0 grad f).v / a0
                const float gv = a0[k] != 0.f ? div_rn_inline(gx[k] * vx[k] + gy[k] * vy[k] + gz[k] * vz[k], a0[k]) : 0.f;
                const float c = 1.f + gv;
                float cn = 1.f, sn = 0.f, tn = 0.f;            // c^j, S_j, T_j = sum_{i<j} S_i
                for (int j = 0; j < n; ++j) { tn += sn; sn += cn; cn *= c; }
                rest[k] = a0[k] != 0.f ? a0[k] * (sn - 1.f) : 0.f;
                tsum[k] = tn;
                g1x[k] = gx[k]; g1y[k] = gy[k]; g1z[k] = gz[k];
            }
        } else if (task == 6) {
#pragma unroll
            for (int k = 0; k < KB; ++k) {
                if (conv[k] < steps) {
                    const float n = (float)(steps - conv[k]);
                    // dv += f(q) * sum_j l_j,  l_j = l_0 + S_j (a0 grad f)
                    dvx[k] += f[k] * fmaf(n, lx[k], tsum[k] * g1x[k]);
                    dvy[k] += f[k] * fmaf(n, ly[k], tsum[k] * g1y[k]);
                    dvz[k] += f[k] * fmaf(n, lz[k], tsum[k] * g1z[k]);
                    lx[k] += g1x[k] + gx[k]; ly[k] += g1y[k] + gy[k]; lz[k] += g1z[k] + gz[k];
                }
            }
        } else {
#pragma unroll
            for (int k = 0; k < KB; ++k) {
                if (s < conv[k]) {
                    dvx[k] = fmaf(f[k], lx[k], dvx[k]); dvy[k] = fmaf(f[k], ly[k], dvy[k]); dvz[k] = fmaf(f[k], lz[k], dvz[k]);
                    lx[k] += gx[k]; ly[k] += gy[k]; lz[k] += gz[k];
                }
            }
        }
        ++task;
    }
}

// Grid-stride over sample groups so the number of blocks (and global atomics) stays bounded.
template <int CAP, int MODE>
__global__ void __launch_bounds__(BLOCK_B, RM_MINB_BWD)
sdf_vjp_kernel(const __grid_constant__ TapeCode<CAP> code, const float* par, int n_instr, int n_par,
               int acc_shared, int iters, const float* __restrict__ pts, const float* __restrict__ gout, int64_t n,
               float* __restrict__ dpts, float* __restrict__ dpar) {
    extern __shared__ __align__(16) unsigned char smem[];
    const GradAcc<MODE> sacc = stage_accumulators<MODE>(smem, n_par, acc_shared);
    const float4* spar = stage_params(par, n_par, smem);                // syncs
    for (int it = 0; it < iters; ++it) {                    // block-uniform trip count (see backward_grid)
        const int64_t g = ((int64_t)it * gridDim.x + blockIdx.x) * BLOCK_B + threadIdx.x;
        const int64_t first = g * KB;
        float x[KB], y[KB], z[KB], a[KB], f[KB], gx[KB], gy[KB], gz[KB];
        load3k(pts, first, n, x, y, z);
        load1k(gout, first, n, a);
        NodeRecords<KB, CAP> rec;
        eval_record<KB, CAP>(code, n_instr, spar, x, y, z, rec, f);
        sweep_vjp<KB, CAP, MODE>(code, n_instr, spar, x, y, z, a, rec, sacc, gx, gy, gz);
        store3k(dpts, first, n, gx, gy, gz);
    }
    flush_accumulators(sacc, dpar, n_par);
}

template <int CAP, int MODE>
__global__ void __launch_bounds__(BLOCK_B, RM_MINB_BWD)
march_bwd_kernel(const __grid_constant__ TapeCode<CAP> code, const float* par, int n_instr, int n_par,
                 int acc_shared, int iters, const float* __restrict__ origin, const float* __restrict__ dirs,
                 const float* __restrict__ gpos, int64_t n, int steps, float tol, float* __restrict__ traj,
                 float* __restrict__ dorigin, float* __restrict__ ddirs, float* __restrict__ dpar) {
    extern __shared__ __align__(16) unsigned char smem[];
    const GradAcc<MODE> sacc = stage_accumulators<MODE>(smem, n_par, acc_shared);
    const float4* spar = stage_params(par, n_par, smem);
    for (int it = 0; it < iters; ++it) {                    // block-uniform trip count (see backward_grid)
        const int64_t g = ((int64_t)it * gridDim.x + blockIdx.x) * BLOCK_B + threadIdx.x;
        const int64_t first = g * KB;
        float px[KB], py[KB], pz[KB], vx[KB], vy[KB], vz[KB];
        load3k(origin, first, n, px, py, pz);
        load3k(dirs, first, n, vx, vy, vz);
        int conv[KB];
        float ux[KB], uy[KB], uz[KB];
        forward_pass<CAP>(code, n_instr, spar, steps, tol, nullptr, traj, px, py, pz, vx, vy, vz, conv, ux, uy, uz);
        float lx[KB], ly[KB], lz[KB], dvx[KB], dvy[KB], dvz[KB];
        load3k(gpos, first, n, lx, ly, lz);
#pragma unroll
        for (int k = 0; k < KB; ++k) dvx[k] = dvy[k] = dvz[k] = 0.f;
        reverse_pass<CAP, MODE>(code, n_instr, spar, steps, nullptr, nullptr, traj, sacc, conv, px, py, pz,
                          vx, vy, vz, lx, ly, lz, dvx, dvy, dvz);
        store3k(dorigin, first, n, lx, ly, lz);
        store3k(ddirs, first, n, dvx, dvy, dvz);
    }
    flush_accumulators(sacc, dpar, n_par);
}

// ---- fused G-buffer backward ------------------------------------------------------------------
struct RenderBwdParams {
    int acc_shared;
    int iters;
    CamParams cam;
    Stencil st;
    int64_t ray_begin, n_rays, strip_rays, strip_stride;
    int steps;
    float tol;
    RmRenderGrads g;
    float* traj;
    float* dpar;
    float* dposes;
    // LOSS instantiation (rm_render_loss_bwd): the upstream gradients are not read from planes but
    // derived in registers from the residuals of the shaded pixel against the target planes
    const float* target_lam;
    const float* target_depth;
    float w_lam, w_depth;
    float* loss_out;        // [2]: weighted sums of squared residuals (lambertian, depth)
    int want_grad;          // 0: loss only (no reverse pass)
    float* nrec;            // node-record scratch (RM_LOSS_ONFLY; see nrec_index)
    int n_nodes;
    int patch_cols;         // > 0: a warp owns an 8 x 4 pixel patch (patch_cols = width / 8), see render_bwd_body
    float* aux;             // split route: per-ray planes (conv, u[3], p_S[3]) + c per node, see loss_replay_kernel
    int rev_grid;           // split route: blocks of the reverse kernel (the replay kernel addresses ITS slots)
};

// Adjoint of w = R(q) b w.r.t. q for an upstream adjoint l of w (R = rotation polynomial):
//   d/dw   = 2 l . (v x b)
//   d/dv   = 2 w (b x l) + 2 [ (v.b) l + (l.v) b - 2 (l.b) v ]          (v = vector part of q)
__device__ __forceinline__ void rotation_vjp(const float (&q)[4], const float (&b)[3], const float (&l)[3],
                                             float (&dq)[4]) {
    const float w = q[0], vx = q[1], vy = q[2], vz = q[3];
    const float cx = vy * b[2] - vz * b[1], cy = vz * b[0] - vx * b[2], cz = vx * b[1] - vy * b[0];   // v x b
    dq[0] += 2.f * (l[0] * cx + l[1] * cy + l[2] * cz);
    const float bx = b[1] * l[2] - b[2] * l[1], by = b[2] * l[0] - b[0] * l[2], bz = b[0] * l[1] - b[1] * l[0]; // b x l
    const float vb = vx * b[0] + vy * b[1] + vz * b[2];
    const float lv = l[0] * vx + l[1] * vy + l[2] * vz;
    const float lb = l[0] * b[0] + l[1] * b[1] + l[2] * b[2];
    dq[1] += 2.f * (w * bx + vb * l[0] + lv * b[0] - 2.f * lb * vx);
    dq[2] += 2.f * (w * by + vb * l[1] + lv * b[1] - 2.f * lb * vy);
    dq[3] += 2.f * (w * bz + vb * l[2] + lv * b[2] - 2.f * lb * vz);
}

// Called by whole, converged warps at kernel end.  Lanes may hold different views only across a
// view boundary, so the loop below runs once in almost every warp.
__device__ __forceinline__ void flush_pose_grad(float* dposes, int view, const float (&acc)[7]) {
    const int lane = threadIdx.x & 31;
    unsigned todo = 0xffffffffu;
    while (todo) {
        const int leader = __ffs(todo) - 1;
        const int v = __shfl_sync(0xffffffffu, view, leader);
        const bool mine = view == v;
#pragma unroll
        for (int i = 0; i < 7; ++i) {
            const float x = warp_sum(mine ? acc[i] : 0.f);
            if (lane == 0 && v >= 0 && x != 0.f) atomicAdd(dposes + (size_t)v * POSE_STRIDE + i, x);
        }
        todo &= ~__ballot_sync(0xffffffffu, mine);
    }
}

// LOSS = false: rm_render_bwd (upstream gradients of the G-buffer planes come from memory).
// LOSS = true : rm_render_loss_bwd -- forward AND backward of the inverse-rendering loss of SURVEY.md
// 8(d) in one kernel: after the replay the thread holds p_S, the stencil vector u and the ray, so it
// shades the pixel (LambertianShader, shader.py:16-20; raw depth |p_S - origin|), forms the residuals
// against the target planes, adds their weighted squares to the loss and seeds the reverse pass with
//     g_lam = 2 w_lam r_lam [0 <= -v.n <= 1],   g_depth = 2 w_depth r_depth
//     d/dn = -g_lam v,  d/dv = -g_lam n,  d/dp_S = g_depth e/|e|,  d/do = -g_depth e/|e|   (e = p_S - o)
// (clamp passes its gradient on the closed interval, the norm has sub-gradient 0 at 0: torch's
// conventions).  No G-buffer plane, no gradient plane and no elementwise kernel exists on this route.
template <int CAP, int MODE, bool LOSS, bool ONFLY = false, bool SPLIT = false>
__device__ __forceinline__ void render_bwd_body(const TapeCode<CAP>& code, const float* par, int n_instr, int n_par,
                                                const float* __restrict__ poses, const RenderBwdParams& P) {
    extern __shared__ __align__(16) unsigned char smem[];
    const GradAcc<MODE> sacc = stage_accumulators<MODE>(smem, n_par, P.acc_shared);
    const float4* spar = stage_params(par, n_par, smem);
    const int64_t n = P.n_rays;
    const bool need_stencil = LOSS ? (P.target_lam != nullptr)
                                   : (P.g.normals != nullptr || P.g.laplacian != nullptr);
    const int64_t per_view = (int64_t)P.cam.width * P.cam.height;

    float cam_acc[7];                       // dq[4], dt[3] of `acc_view`, summed over this thread's rays
#pragma unroll
    for (int i = 0; i < 7; ++i) cam_acc[i] = 0.f;
    int acc_view = -1;
    float loss_l = 0.f, loss_d = 0.f;       // LOSS: this thread's sums of squared residuals

    const int iters = P.iters;
    for (int it = 0; it < iters; ++it) {                    // block-uniform trip count (see backward_grid)
        const int64_t g = ((int64_t)it * gridDim.x + blockIdx.x) * BLOCK_B + threadIdx.x;
        int64_t first = g * KB;
        if (KB == 1 && P.patch_cols > 0) {
            // 2-D warp tiles: the 32 rays of a warp are an 8 x 4 pixel patch instead of 32 pixels of one row, so
            // that what the warp decides together (settled rays, exact steps to reverse, dead children of a
            // smooth union) is decided for neighbouring rays.  Whole rows, width % 8 == 0, rows % 4 == 0 (host).
            const unsigned t = (unsigned)(g >> 5), lane = threadIdx.x & 31u;
            const unsigned ty = t / (unsigned)P.patch_cols, tx = t - ty * (unsigned)P.patch_cols;
            first = (int64_t)(ty * 4u + (lane >> 3)) * P.cam.width + (tx * 8u + (lane & 7u));
            if (first >= n) first = n;                      // (grid rounding: past the end, nothing is stored)
        }
        // ---- rays from the pixel index (PinholeCamera.forward) ----
        float px[KB], py[KB], pz[KB], vx[KB], vy[KB], vz[KB];
        Pose pose;
        int cur_view = -1;
#pragma unroll
        for (int k = 0; k < KB; ++k) {
            float ob[3], vb[3], ow[3], vw[3];
            world_ray(P.cam, poses, shard_ray(P.ray_begin, P.strip_rays, P.strip_stride, first + k), pose, cur_view, ob, vb, ow, vw);
            px[k] = ow[0]; py[k] = ow[1]; pz[k] = ow[2];
            vx[k] = vw[0]; vy[k] = vw[1]; vz[k] = vw[2];
        }
        // ---- forward pass: replay the march, sample the stencil offsets ----
        int conv[KB];
        float ux[KB], uy[KB], uz[KB];
        using NodeRec = NodeRecords<KB, ONFLY ? 1 : CAP>;  // ONFLY: union nodes only (no per-instruction values)
        NodeRec frec_store;                                // ... shared by the replay and the sweeps
        NodeRec* frec = nullptr;
        if constexpr (ONFLY) frec = &frec_store;
        if constexpr (SPLIT) {
            // the replay ran in loss_replay_kernel: p_S, u, conv of this thread's ray and c per node come back
            // from its planes; column = (iteration, slot), as reverse_pass addresses it with (kf, ksel) = (iters, it)
            const size_t n_slots = (size_t)gridDim.x * BLOCK_B, ncols = n_slots * (size_t)P.iters;
            const size_t col = (size_t)it * n_slots + (size_t)blockIdx.x * BLOCK_B + threadIdx.x;
            static_assert(!SPLIT || KB == 1, "the split route reverses one ray per thread");
            const bool mine = first < n;                     // threads past the end own no column: inert values
            conv[0] = mine ? __float_as_int(P.aux[0 * ncols + col]) : 0;
            ux[0] = mine ? P.aux[1 * ncols + col] : 0.f; uy[0] = mine ? P.aux[2 * ncols + col] : 0.f;
            uz[0] = mine ? P.aux[3 * ncols + col] : 1.f;
            px[0] = mine ? P.aux[4 * ncols + col] : 0.f; py[0] = mine ? P.aux[5 * ncols + col] : 0.f;
            pz[0] = mine ? P.aux[6 * ncols + col] : 0.f;
            for (int node = 0; node < P.n_nodes; ++node) frec->c[node] = P.aux[7 * ncols + node];
        } else {
            forward_pass<CAP, LOSS && RM_LOSS_FAST_REPLAY, ONFLY, NodeRec>(code, n_instr, spar, P.steps, P.tol,
                                                                  need_stencil ? &P.st : nullptr, P.traj, px, py, pz,
                                                                  vx, vy, vz, conv, ux, uy, uz, frec,
                                                                  P.want_grad ? P.nrec : nullptr, P.n_nodes);
        }

        // ---- adjoints of the stencil outputs (SDFNormals.forward, ray_marching.py:115-125), of
        //      f(p_S) (control.py:244), of p_S, the direction and the origin ----
        float lx[KB], ly[KB], lz[KB], dvx[KB], dvy[KB], dvz[KB], gox[KB], goy[KB], goz[KB];
        StencilAdjoint sa;
        if constexpr (LOSS) {
#pragma unroll
            for (int k = 0; k < KB; ++k) {
                const int64_t r = shard_ray(P.ray_begin, P.strip_rays, P.strip_stride, first + k);
                const bool valid = first + k < n;
                float ob[3], vb[3], ow[3], vw[3];
                world_ray(P.cam, poses, r, pose, cur_view, ob, vb, ow, vw);      // origin, regenerated
                const bool has_n = P.target_lam != nullptr;     // no lambertian term: the stencil was not sampled
                const float len = sqrt_rn_inline(ux[k] * ux[k] + uy[k] * uy[k] + uz[k] * uz[k]);
                const float il = has_n ? div_rn_inline(1.f, len) : 0.f;
                const float nx = ux[k] * il, ny = uy[k] * il, nz = uz[k] * il;
                float g_lam = 0.f, g_dep = 0.f;
                if (has_n && valid) {
                    const float nv = -(vx[k] * nx + vy[k] * ny + vz[k] * nz);
                    const float res = clamp_nan(nv, 0.f, 1.f) - __ldg(P.target_lam + first + k);
                    loss_l = fmaf(res, res, loss_l);
                    // clamp passes the gradient on [0, 1] inclusive; a NaN pixel poisons the gradient
                    // exactly as it poisons the reference's (NaN * 0 = NaN there, too)
                    g_lam = (!(nv < 0.f) && !(nv > 1.f)) ? 2.f * P.w_lam * res : 0.f;
                }
                const float ex = px[k] - ow[0], ey = py[k] - ow[1], ez = pz[k] - ow[2];
                float hx = 0.f, hy = 0.f, hz = 0.f;
                if (P.target_depth != nullptr && valid) {
                    const float depth = sqrt_rn_inline(ex * ex + ey * ey + ez * ez);
                    const float res = depth - __ldg(P.target_depth + first + k);
                    loss_d = fmaf(res, res, loss_d);
                    g_dep = 2.f * P.w_depth * res;
                    const float id = rcp_or_zero(depth);
                    hx = g_dep * ex * id; hy = g_dep * ey * id; hz = g_dep * ez * id;
                }
                lx[k] = hx; ly[k] = hy; lz[k] = hz;
                gox[k] = -hx; goy[k] = -hy; goz[k] = -hz;
                dvx[k] = -g_lam * nx; dvy[k] = -g_lam * ny; dvz[k] = -g_lam * nz;
                // d nrm / d u = (I - nrm nrm^T) / |u| applied to g_n = -g_lam v
                const float gnx = -g_lam * vx[k], gny = -g_lam * vy[k], gnz = -g_lam * vz[k];
                const float gn = gnx * nx + gny * ny + gnz * nz;
                sa.gux[k] = has_n ? (gnx - gn * nx) * il : 0.f;
                sa.guy[k] = has_n ? (gny - gn * ny) * il : 0.f;
                sa.guz[k] = has_n ? (gnz - gn * nz) * il : 0.f;
                sa.gm[k] = 0.f; sa.gs[k] = 0.f;
            }
        } else {
            load3k(P.g.pos, first, n, lx, ly, lz);
            load1k(P.g.sdf, first, n, sa.gs);
            float gnx[KB], gny[KB], gnz[KB], glap[KB];
            load3k(P.g.normals, first, n, gnx, gny, gnz);
            load1k(P.g.laplacian, first, n, glap);
#pragma unroll
            for (int k = 0; k < KB; ++k) {
                // forward: nrm = u / |u|; d nrm / d u = (I - nrm nrm^T) / |u|
                const float len = sqrt_rn_inline(ux[k] * ux[k] + uy[k] * uy[k] + uz[k] * uz[k]);
                const float il = div_rn_inline(1.f, len);
                const float nx = ux[k] * il, ny = uy[k] * il, nz = uz[k] * il;
                const float gn = gnx[k] * nx + gny[k] * ny + gnz[k] * nz;
                const bool has_n = P.g.normals != nullptr;
                sa.gux[k] = has_n ? (gnx[k] - gn * nx) * il : 0.f;
                sa.guy[k] = has_n ? (gny[k] - gn * ny) * il : 0.f;
                sa.guz[k] = has_n ? (gnz[k] - gn * nz) * il : 0.f;
                sa.gm[k] = glap[k] * P.st.lap_scale;
            }
            load3k(P.g.dirs, first, n, dvx, dvy, dvz);
            load3k(P.g.origin, first, n, gox, goy, goz);
        }
        if (LOSS && !P.want_grad) continue;                 // loss evaluation only

        // ---- reverse pass: stencil samples, converged tail, exact steps ----
        reverse_pass<CAP, MODE, LOSS && RM_LOSS_FAST_REPLAY, ONFLY, NodeRec>(code, n_instr, spar, P.steps, &P.st,
                          (need_stencil || (!LOSS && P.g.sdf != nullptr)) ? &sa : nullptr, P.traj, sacc, conv,
                          px, py, pz, vx, vy, vz, lx, ly, lz, dvx, dvy, dvz, frec, P.nrec, P.n_nodes,
                          SPLIT ? P.iters : 1, SPLIT ? it : 0);

        // ---- camera pose: origin = R(q) o_b + t, dir = R(q) v_b ----
#pragma unroll
        for (int k = 0; k < KB; ++k) {
            const int64_t r = shard_ray(P.ray_begin, P.strip_rays, P.strip_stride, first + k);
            const bool valid = first + k < n;
            const int view = valid ? (int)(r / per_view) : acc_view;
            if (view != acc_view) {                           // rare: view boundary inside a thread's stream
                if (acc_view >= 0) {
#pragma unroll
                    for (int i = 0; i < 7; ++i)
                        if (cam_acc[i] != 0.f) atomicAdd(P.dposes + (size_t)acc_view * POSE_STRIDE + i, cam_acc[i]);
                }
#pragma unroll
                for (int i = 0; i < 7; ++i) cam_acc[i] = 0.f;
                acc_view = view;
            }
            if (valid) {
                // camera-frame ray regenerated from the pixel index (not kept live across the passes)
                float ob[3], vb[3], ow[3], vw[3];
                world_ray(P.cam, poses, r, pose, cur_view, ob, vb, ow, vw);
                const Pose& ps = pose;
                const float lo[3] = {lx[k] + gox[k], ly[k] + goy[k], lz[k] + goz[k]};
                const float lv[3] = {dvx[k], dvy[k], dvz[k]};
                float dq[4] = {0.f, 0.f, 0.f, 0.f};
                rotation_vjp(ps.q, ob, lo, dq);
                rotation_vjp(ps.q, vb, lv, dq);
                cam_acc[0] += dq[0]; cam_acc[1] += dq[1]; cam_acc[2] += dq[2]; cam_acc[3] += dq[3];
                cam_acc[4] += lo[0]; cam_acc[5] += lo[1]; cam_acc[6] += lo[2];
            }
        }
    }
    if constexpr (LOSS) {
        const float sl = warp_sum(loss_l) * P.w_lam, sd = warp_sum(loss_d) * P.w_depth;
        if ((threadIdx.x & 31) == 0) {
            if (P.target_lam != nullptr) atomicAdd(P.loss_out + 0, sl);
            if (P.target_depth != nullptr) atomicAdd(P.loss_out + 1, sd);
        }
        if (!P.want_grad) return;
    }
    flush_pose_grad(P.dposes, acc_view, cam_acc);
    flush_accumulators(sacc, P.dpar, n_par);
}

template <int CAP, int MODE>
__global__ void __launch_bounds__(BLOCK_B, RM_MINB_BWD)
render_bwd_kernel(const __grid_constant__ TapeCode<CAP> code, const float* par, int n_instr, int n_par,
                  const float* __restrict__ poses, const __grid_constant__ RenderBwdParams P) {
    render_bwd_body<CAP, MODE, false>(code, par, n_instr, n_par, poses, P);
}
template <int CAP, int MODE>
__global__ void __launch_bounds__(BLOCK_B, RM_MINB_BWD)
render_loss_bwd_kernel(const __grid_constant__ TapeCode<CAP> code, const float* par, int n_instr, int n_par,
                       const float* __restrict__ poses, const __grid_constant__ RenderBwdParams P) {
    render_bwd_body<CAP, MODE, true>(code, par, n_instr, n_par, poses, P);
}
// ... with the union-node records kept from the replay and the leaves re-evaluated by the sweep (long tapes)
template <int CAP, int MODE>
__global__ void __launch_bounds__(BLOCK_B, RM_MINB_BWD)
render_loss_onfly_kernel(const __grid_constant__ TapeCode<CAP> code, const float* par, int n_instr, int n_par,
                         const float* __restrict__ poses, const __grid_constant__ RenderBwdParams P) {
    render_bwd_body<CAP, MODE, true, true>(code, par, n_instr, n_par, poses, P);
}
// ---- two-phase loss kernel for long tapes (render_loss_onfly4_kernel) ---------------------------------
// One ray per thread leaves the replay's interpreter decode unamortised (~100 warp instructions per tape
// instruction against ~35 at four rays per thread).  Here a thread REPLAYS four rays at once (phase 1:
// replay_record, node records and trajectory to the slot scratch, planes [.][k][slot]) and then REVERSES them
// one after the other (phase 2: the one-ray reverse_pass with (kf, ksel) addressing).  The four rays of a
// thread are the same lane position of four different warp tiles -- tiles t0 + 4 k + warp -- so that in phase
// 2, pass k, the 32 lanes of a warp hold one whole tile (an 8 x 4 pixel patch, or 32 rays of a row): what the
// warp decides together is still decided for neighbouring rays, and every scratch element is read by the
// thread that wrote it (no block barrier between the phases).
// Measured (c4b, 1024^2): 23.6 ms against 21.4 ms for the one-ray route -- the replay executes a third of the
// instructions, but 512 rays per block iteration quantise 2048 iterations onto 592 block slots (1.73 waves,
// 13 % idle) and the four-ray recording interpreter spills inside the 128-register kernel.  Kept as an opt-in
// route (RM_B200_LOSS_ROUTE=onfly4, -DRM_LOSS_TWO_PHASE=1) with the same tests; a per-warp tile queue would
// remove the quantisation.
constexpr int KF = 4;

template <int CAP>
__device__ __forceinline__ void replay_record(const TapeCode<CAP>& code, int n_instr, const float4* spar,
                                              int steps, float tol, const Stencil* st, float* traj,
                                              float (&px)[KF], float (&py)[KF], float (&pz)[KF],
                                              const float (&vx)[KF], const float (&vy)[KF],
                                              const float (&vz)[KF], int (&conv)[KF],
                                              float (&ux)[KF], float (&uy)[KF], float (&uz)[KF],
                                              NodeRecords<KF, 1>& rec, float* nrec, int n_nodes) {
    float o0[KF], prev[KF];
#pragma unroll
    for (int k = 0; k < KF; ++k) { conv[k] = 0; ux[k] = uy[k] = uz[k] = o0[k] = 0.f; prev[k] = -1.f; }
    const int n_tasks = steps + (st != nullptr ? 4 : 0);
    for (int task = 0; task < n_tasks; ++task) {               // forward_pass, four rays wide
        const bool marching = task < steps;
        const int j = marching ? 4 : task - steps;
        float x[KF], y[KF], z[KF], d[KF];
        float dx = 0.f, dy = 0.f, dz = 0.f;
        if (!marching) { dx = st->off[j][0]; dy = st->off[j][1]; dz = st->off[j][2]; }
        else if (nrec != nullptr) traj_store_k<KF>(traj, task, px, py, pz);
#pragma unroll
        for (int k = 0; k < KF; ++k) { x[k] = px[k] + dx; y[k] = py[k] + dy; z[k] = pz[k] + dz; }
        eval_scene<KF, true, CAP, 0, 2, NodeRecords<KF, 1>>(code, n_instr, spar, x, y, z, d, nullptr, nullptr, &rec);
        if (nrec != nullptr) nrec_store_k<KF>(nrec, n_nodes, task, rec, d);
        if (marching) {
            bool at_rest = true;
#pragma unroll
            for (int k = 0; k < KF; ++k) {
                const float scale = 1.f + fmaxf(fabsf(px[k]), fmaxf(fabsf(py[k]), fabsf(pz[k])));
                const float ad = fabsf(d[k]);
                const bool within = ad <= tol * scale;                       // false for NaN
                if (!within) conv[k] = task + 1;
                at_rest = at_rest && within && prev[k] >= ad;
                prev[k] = within ? ad : -1.f;
                px[k] = fmaf(d[k], vx[k], px[k]); py[k] = fmaf(d[k], vy[k], py[k]); pz[k] = fmaf(d[k], vz[k], pz[k]);
            }
#if RM_BWD_EARLY_BREAK
            if (__all_sync(0xffffffffu, at_rest)) task = steps - 1;
#endif
        } else {
            const float wx = st->w[j][0], wy = st->w[j][1], wz = st->w[j][2], first = st->first[j];
#pragma unroll
            for (int k = 0; k < KF; ++k) {
                o0[k] = first != 0.f ? d[k] : o0[k];
                const float dj = d[k] - o0[k];
                ux[k] = fmaf(wx, dj, ux[k]); uy[k] = fmaf(wy, dj, uy[k]); uz[k] = fmaf(wz, dj, uz[k]);
            }
        }
    }
}

__device__ __forceinline__ float pick(const float (&a)[KF], int j) {      // a[j] without a local-memory array
    return j == 0 ? a[0] : (j == 1 ? a[1] : (j == 2 ? a[2] : a[3]));
}

template <int CAP, int MODE>
__global__ void __launch_bounds__(BLOCK_B, RM_MINB_BWD)
render_loss_onfly4_kernel(const __grid_constant__ TapeCode<CAP> code, const float* par, int n_instr, int n_par,
                          const float* __restrict__ poses, const __grid_constant__ RenderBwdParams P) {
    static_assert(KB == 1, "the two-phase loss kernel reverses one ray per thread");
    extern __shared__ __align__(16) unsigned char smem[];
    const GradAcc<MODE> sacc = stage_accumulators<MODE>(smem, n_par, P.acc_shared);
    const float4* spar = stage_params(par, n_par, smem);
    const int64_t n = P.n_rays;
    const bool need_stencil = P.target_lam != nullptr;
    const int64_t per_view = (int64_t)P.cam.width * P.cam.height;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    float cam_acc[7];
#pragma unroll
    for (int i = 0; i < 7; ++i) cam_acc[i] = 0.f;
    int acc_view = -1;
    float loss_l = 0.f, loss_d = 0.f;

    for (int it = 0; it < P.iters; ++it) {                     // block-uniform trip count
        // ---- the thread's four rays: lane `lane` of tiles t0 + 4 k + warp ----
        const int64_t t0 = ((int64_t)it * gridDim.x + blockIdx.x) * (KF * (BLOCK_B / 32));
        int64_t first[KF];
        float px[KF], py[KF], pz[KF], vx[KF], vy[KF], vz[KF];
        Pose pose;
        int cur_view = -1;
#pragma unroll
        for (int k = 0; k < KF; ++k) {
            const int64_t t = t0 + k * (BLOCK_B / 32) + warp;
            int64_t f = t * 32 + lane;
            if (P.patch_cols > 0) {
                const unsigned ty = (unsigned)t / (unsigned)P.patch_cols, tx = (unsigned)t - ty * (unsigned)P.patch_cols;
                f = (int64_t)(ty * 4u + ((unsigned)lane >> 3)) * P.cam.width + (tx * 8u + ((unsigned)lane & 7u));
            }
            first[k] = f < n ? f : n;                          // past the end: marched (clamped duplicate), never used
            float ob[3], vb[3], ow[3], vw[3];
            world_ray(P.cam, poses, shard_ray(P.ray_begin, P.strip_rays, P.strip_stride, first[k]), pose, cur_view, ob, vb, ow, vw);
            px[k] = ow[0]; py[k] = ow[1]; pz[k] = ow[2];
            vx[k] = vw[0]; vy[k] = vw[1]; vz[k] = vw[2];
        }
        // ---- phase 1: replay four rays, records to the scratch ----
        int conv[KF];
        float ux[KF], uy[KF], uz[KF];
        NodeRecords<KF, 1> rec4;
        replay_record<CAP>(code, n_instr, spar, P.steps, P.tol, need_stencil ? &P.st : nullptr, P.traj,
                           px, py, pz, vx, vy, vz, conv, ux, uy, uz, rec4, P.want_grad ? P.nrec : nullptr, P.n_nodes);
        NodeRecords<1, 1> rec1;                                // c (constant per tape) for the sweeps
        for (int node = 0; node < P.n_nodes; ++node) rec1.c[node] = rec4.c[node];

        // ---- phase 2: loss, seeds and reverse pass, one ray at a time ----
#pragma unroll 1
        for (int j = 0; j < KF; ++j) {
            const int64_t fj = j == 0 ? first[0] : (j == 1 ? first[1] : (j == 2 ? first[2] : first[3]));
            const bool valid = fj < n;
            const int64_t r = shard_ray(P.ray_begin, P.strip_rays, P.strip_stride, fj);
            const float psx[1] = {pick(px, j)}, psy[1] = {pick(py, j)}, psz[1] = {pick(pz, j)};
            const float v1x[1] = {pick(vx, j)}, v1y[1] = {pick(vy, j)}, v1z[1] = {pick(vz, j)};
            const float u0 = pick(ux, j), u1 = pick(uy, j), u2 = pick(uz, j);
            const int conv1[1] = {j == 0 ? conv[0] : (j == 1 ? conv[1] : (j == 2 ? conv[2] : conv[3]))};
            float ob[3], vb[3], ow[3], vw[3];
            world_ray(P.cam, poses, r, pose, cur_view, ob, vb, ow, vw);          // origin, regenerated
            // residuals and adjoint seeds (the LOSS branch of render_bwd_body)
            const bool has_n = P.target_lam != nullptr;
            const float len = sqrt_rn_inline(u0 * u0 + u1 * u1 + u2 * u2);
            const float il = has_n ? div_rn_inline(1.f, len) : 0.f;
            const float nx = u0 * il, ny = u1 * il, nz = u2 * il;
            float g_lam = 0.f, g_dep = 0.f;
            if (has_n && valid) {
                const float nv = -(v1x[0] * nx + v1y[0] * ny + v1z[0] * nz);
                const float res = clamp_nan(nv, 0.f, 1.f) - __ldg(P.target_lam + fj);
                loss_l = fmaf(res, res, loss_l);
                g_lam = (!(nv < 0.f) && !(nv > 1.f)) ? 2.f * P.w_lam * res : 0.f;
            }
            const float ex = psx[0] - ow[0], ey = psy[0] - ow[1], ez = psz[0] - ow[2];
            float hx = 0.f, hy = 0.f, hz = 0.f;
            if (P.target_depth != nullptr && valid) {
                const float depth = sqrt_rn_inline(ex * ex + ey * ey + ez * ez);
                const float res = depth - __ldg(P.target_depth + fj);
                loss_d = fmaf(res, res, loss_d);
                g_dep = 2.f * P.w_depth * res;
                const float id = rcp_or_zero(depth);
                hx = g_dep * ex * id; hy = g_dep * ey * id; hz = g_dep * ez * id;
            }
            if (!P.want_grad) continue;                        // loss evaluation only
            float lx[1] = {hx}, ly[1] = {hy}, lz[1] = {hz};
            const float gox = -hx, goy = -hy, goz = -hz;
            float dvx[1] = {-g_lam * nx}, dvy[1] = {-g_lam * ny}, dvz[1] = {-g_lam * nz};
            StencilAdjoint sa;
            {
                const float gnx = -g_lam * v1x[0], gny = -g_lam * v1y[0], gnz = -g_lam * v1z[0];
                const float gn = gnx * nx + gny * ny + gnz * nz;
                sa.gux[0] = has_n ? (gnx - gn * nx) * il : 0.f;
                sa.guy[0] = has_n ? (gny - gn * ny) * il : 0.f;
                sa.guz[0] = has_n ? (gnz - gn * nz) * il : 0.f;
                sa.gm[0] = 0.f; sa.gs[0] = 0.f;
            }
            reverse_pass<CAP, MODE, true, true, NodeRecords<1, 1>>(code, n_instr, spar, P.steps, &P.st, need_stencil ? &sa : nullptr,
                                                P.traj, sacc, conv1, psx, psy, psz, v1x, v1y, v1z, lx, ly, lz,
                                                dvx, dvy, dvz, &rec1, P.nrec, P.n_nodes, KF, j);
            // camera pose: origin = R(q) o_b + t, dir = R(q) v_b
            const int view = valid ? (int)(r / per_view) : acc_view;
            if (view != acc_view) {
                if (acc_view >= 0) {
#pragma unroll
                    for (int i = 0; i < 7; ++i)
                        if (cam_acc[i] != 0.f) atomicAdd(P.dposes + (size_t)acc_view * POSE_STRIDE + i, cam_acc[i]);
                }
#pragma unroll
                for (int i = 0; i < 7; ++i) cam_acc[i] = 0.f;
                acc_view = view;
            }
            if (valid) {
                const float lo[3] = {lx[0] + gox, ly[0] + goy, lz[0] + goz};
                const float lv[3] = {dvx[0], dvy[0], dvz[0]};
                float dq[4] = {0.f, 0.f, 0.f, 0.f};
                rotation_vjp(pose.q, ob, lo, dq);
                rotation_vjp(pose.q, vb, lv, dq);
                cam_acc[0] += dq[0]; cam_acc[1] += dq[1]; cam_acc[2] += dq[2]; cam_acc[3] += dq[3];
                cam_acc[4] += lo[0]; cam_acc[5] += lo[1]; cam_acc[6] += lo[2];
            }
        }
    }
    const float sl = warp_sum(loss_l) * P.w_lam, sd = warp_sum(loss_d) * P.w_depth;
    if ((threadIdx.x & 31) == 0) {
        if (P.target_lam != nullptr) atomicAdd(P.loss_out + 0, sl);
        if (P.target_depth != nullptr) atomicAdd(P.loss_out + 1, sd);
    }
    if (!P.want_grad) return;
    flush_pose_grad(P.dposes, acc_view, cam_acc);
    flush_accumulators(sacc, P.dpar, n_par);
}
// Which of the two: the records pay for themselves once a recording walk over the tape costs more than their
// traffic and the slower replay (measured: c4b, 69 instructions, 27.3 -> 23.3 ms; c4, 6 instructions,
// 0.41 -> 0.52 ms).  A tape-specialised library replays straight-line code and never uses them.
// ---- split route for long tapes: loss_replay_kernel + render_loss_reverse_kernel ------------------------
// Inside the 128-register loss kernel (16 warps per SM) the replay runs far below what the forward kernel
// reaches on the same interpreter.  Here it is a kernel of its own, shaped like render_kernel -- four rays per
// thread, 256 threads, 64 registers, 32 warps per SM -- that writes every ray's trajectory, node records and
// (conv, u, p_S) to per-ray planes; the reverse kernel is the one-ray body of the node-record route without its
// replay.  The planes are addressed the way the REVERSE kernel numbers its threads: column = (iteration,
// slot), so reverse_pass finds them with (kf, ksel) = (iterations, iteration) and nothing else changes.
__device__ __forceinline__ size_t reverse_column(const RenderBwdParams& P, int64_t f) {
    // global thread index g of the reverse kernel that owns local ray f (the inverse of its ray mapping)
    int64_t g = f;
    if (P.patch_cols > 0) {
        const unsigned row = (unsigned)(f / P.cam.width), colp = (unsigned)(f - (int64_t)row * P.cam.width);
        const unsigned t = (row >> 2) * (unsigned)P.patch_cols + (colp >> 3);
        g = (int64_t)t * 32 + ((row & 3u) << 3) + (colp & 7u);
    }
    return (size_t)g;          // = iteration * n_slots + slot, because slot = g mod n_slots and iteration = g / n_slots
}

constexpr int BLOCK_A = 256;      // replay kernel: the forward kernel's launch shape (256 threads x 4 rays, 64 registers)
template <int CAP>
__global__ void __launch_bounds__(BLOCK_A, 4)
loss_replay_kernel(const __grid_constant__ TapeCode<CAP> code, const float* par, int n_instr, int n_par,
                   const float* __restrict__ poses, const __grid_constant__ RenderBwdParams P) {
    extern __shared__ __align__(16) unsigned char smem[];
    const float4* spar = stage_params(par, n_par, smem);
    const int64_t n = P.n_rays;
    const size_t ncols = (size_t)P.rev_grid * BLOCK_B * (size_t)P.iters;
    const int nr = 1 + 2 * P.n_nodes;
    const int64_t first = ((int64_t)blockIdx.x * BLOCK_A + threadIdx.x) * KF;
    // no early return: the replay votes with full-warp masks; rays past the end are clamped duplicates, not stored
    float px[KF], py[KF], pz[KF], vx[KF], vy[KF], vz[KF];
    size_t col[KF];
    bool valid[KF];
    Pose pose;
    int cur_view = -1;
#pragma unroll
    for (int k = 0; k < KF; ++k) {
        valid[k] = first + k < n;
        col[k] = reverse_column(P, valid[k] ? first + k : 0);
        float ob[3], vb[3], ow[3], vw[3];
        world_ray(P.cam, poses, shard_ray(P.ray_begin, P.strip_rays, P.strip_stride, valid[k] ? first + k : n - 1), pose, cur_view, ob, vb, ow, vw);
        px[k] = ow[0]; py[k] = ow[1]; pz[k] = ow[2];
        vx[k] = vw[0]; vy[k] = vw[1]; vz[k] = vw[2];
    }
    const Stencil* st = P.target_lam != nullptr ? &P.st : nullptr;
    int conv[KF];
    float ux[KF], uy[KF], uz[KF], o0[KF], prev[KF];
#pragma unroll
    for (int k = 0; k < KF; ++k) { conv[k] = 0; ux[k] = uy[k] = uz[k] = o0[k] = 0.f; prev[k] = -1.f; }
    NodeRecords<KF, 1> rec;
    const int n_tasks = P.steps + (st != nullptr ? 4 : 0);
    for (int task = 0; task < n_tasks; ++task) {               // forward_pass, four rays wide, per-ray planes
        const bool marching = task < P.steps;
        const int j = marching ? 4 : task - P.steps;
        float x[KF], y[KF], z[KF], d[KF];
        float dx = 0.f, dy = 0.f, dz = 0.f;
        if (!marching) { dx = st->off[j][0]; dy = st->off[j][1]; dz = st->off[j][2]; }
        else {
#pragma unroll
            for (int k = 0; k < KF; ++k)
                if (valid[k]) {
                    float* p = P.traj + ((size_t)task * ncols + col[k]) * 3;
                    p[0] = px[k]; p[1] = py[k]; p[2] = pz[k];
                }
        }
#pragma unroll
        for (int k = 0; k < KF; ++k) { x[k] = px[k] + dx; y[k] = py[k] + dy; z[k] = pz[k] + dz; }
        eval_scene<KF, true, CAP, 0, 2, NodeRecords<KF, 1>>(code, n_instr, spar, x, y, z, d, nullptr, nullptr, &rec);
#pragma unroll
        for (int k = 0; k < KF; ++k)
            if (valid[k]) {
                float* q = P.nrec + (size_t)task * nr * ncols + col[k];
                q[0] = d[k];
                for (int node = 0; node < P.n_nodes; ++node) {
                    q[(size_t)(1 + 2 * node) * ncols] = rec.d[node][k];
                    q[(size_t)(2 + 2 * node) * ncols] = __int_as_float(rec.win[node][k]);
                }
            }
        if (marching) {
            bool at_rest = true;
#pragma unroll
            for (int k = 0; k < KF; ++k) {
                const float scale = 1.f + fmaxf(fabsf(px[k]), fmaxf(fabsf(py[k]), fabsf(pz[k])));
                const float ad = fabsf(d[k]);
                const bool within = ad <= P.tol * scale;                     // false for NaN
                if (!within) conv[k] = task + 1;
                at_rest = at_rest && within && prev[k] >= ad;
                prev[k] = within ? ad : -1.f;
                px[k] = fmaf(d[k], vx[k], px[k]); py[k] = fmaf(d[k], vy[k], py[k]); pz[k] = fmaf(d[k], vz[k], pz[k]);
            }
#if RM_BWD_EARLY_BREAK
            if (__all_sync(0xffffffffu, at_rest)) task = P.steps - 1;
#endif
        } else {
            const float wx = st->w[j][0], wy = st->w[j][1], wz = st->w[j][2], firstw = st->first[j];
#pragma unroll
            for (int k = 0; k < KF; ++k) {
                o0[k] = firstw != 0.f ? d[k] : o0[k];
                const float dj = d[k] - o0[k];
                ux[k] = fmaf(wx, dj, ux[k]); uy[k] = fmaf(wy, dj, uy[k]); uz[k] = fmaf(wz, dj, uz[k]);
            }
        }
    }
#pragma unroll
    for (int k = 0; k < KF; ++k)
        if (valid[k]) {
            float* a = P.aux + col[k];
            a[0] = __int_as_float(conv[k]);
            a[1 * ncols] = ux[k]; a[2 * ncols] = uy[k]; a[3 * ncols] = uz[k];
            a[4 * ncols] = px[k]; a[5 * ncols] = py[k]; a[6 * ncols] = pz[k];
        }
    if (blockIdx.x == 0 && threadIdx.x == 0)
        for (int node = 0; node < P.n_nodes; ++node) P.aux[7 * ncols + node] = rec.c[node];
}

template <int CAP, int MODE>
__global__ void __launch_bounds__(BLOCK_B, RM_MINB_BWD)
render_loss_reverse_kernel(const __grid_constant__ TapeCode<CAP> code, const float* par, int n_instr, int n_par,
                           const float* __restrict__ poses, const __grid_constant__ RenderBwdParams P) {
    render_bwd_body<CAP, MODE, true, true, true>(code, par, n_instr, n_par, poses, P);
}

// Route of rm_render_loss_bwd: 0 = recording walk per reverse task (render_loss_bwd_kernel), 1 = node records,
// one ray per thread throughout (render_loss_onfly_kernel), 2 = node records with the four-ray replay
// (render_loss_onfly4_kernel), 3 = node records written by a replay kernel of its own (loss_replay_kernel +
// render_loss_reverse_kernel).  RM_B200_LOSS_ROUTE = record | onfly | onfly4 | split forces one (tests).
inline int loss_route(const RmTape* t) {
#if defined(RM_SPEC_INC) || !RM_LOSS_ONFLY
    (void)t;
    return 0;
#else
    if (const char* e = getenv("RM_B200_LOSS_ROUTE")) {
        if (e[0] == 'r') return 0;
        if (e[0] == 's') return 3;
        if (e[0] == 'o') return (e[1] == 'n' && e[2] == 'f' && e[3] == 'l' && e[4] == 'y' && e[5] == '4') ? 2 : 1;
    }
    return t->n_instr >= RM_LOSS_ONFLY_MIN_INSTR ? (RM_LOSS_SPLIT ? 3 : (RM_LOSS_TWO_PHASE ? 2 : 1)) : 0;
#endif
}
// Launch geometry and scratch of a route: blocks, block-uniform iterations, floats of trajectory and records.
struct LossLaunch { int route; unsigned grid; int iters; int64_t traj_floats, nrec_floats, aux_floats; };
inline LossLaunch loss_launch(const RmTape* t, int64_t n_rays, int steps, bool want_grad = true) {
    LossLaunch L;
    L.route = loss_route(t);
    if (L.route == 3 && !want_grad) L.route = 1;          // loss only: one kernel, nothing to hand over
    L.aux_floats = 0;
    const int kf = L.route == 2 ? KF : KB;
    const int64_t want = (n_rays + (int64_t)BLOCK_B * kf - 1) / ((int64_t)BLOCK_B * kf);
    const int64_t cap = 148 * 8;
    L.iters = (int)((want + cap - 1) / cap > 0 ? (want + cap - 1) / cap : 1);
    const int64_t blocks = (want + L.iters - 1) / L.iters;
    L.grid = (unsigned)(blocks > 0 ? blocks : 1);
    const int64_t slots = (int64_t)L.grid * BLOCK_B;
    L.traj_floats = (int64_t)steps * kf * slots * 3;
    L.nrec_floats = L.route == 0 ? 0 : (int64_t)(steps + 4) * (1 + 2 * (int64_t)t->n_nodes) * kf * slots;
    if (L.route == 3) {                                    // per-ray planes: a column per (iteration, slot)
        const int64_t ncols = slots * L.iters;
        L.traj_floats = (int64_t)steps * ncols * 3;
        L.nrec_floats = (int64_t)(steps + 4) * (1 + 2 * (int64_t)t->n_nodes) * ncols;
        L.aux_floats = 7 * ncols + MAX_NODES;
        if (L.traj_floats + L.nrec_floats + L.aux_floats > (int64_t)1 << 31) {   // > 8 GB of planes: one-kernel route
            L.route = 1;
            L.aux_floats = 0;
            L.traj_floats = (int64_t)steps * kf * slots * 3;
            L.nrec_floats = (int64_t)(steps + 4) * (1 + 2 * (int64_t)t->n_nodes) * kf * slots;
        }
    }
    return L;
}

inline int check_backward_tape(const RmTape* t) {
    if (int s = check_tape(t)) return s;
    if (t->n_nodes > MAX_NODES)
        return fail(RM_ERR_UNSUPPORTED, "backward supports at most %d union nodes (tape has %d)",
                    MAX_NODES, t->n_nodes);
#ifdef RM_SPEC_INC
    // specialised library: the tape is compiled in; refuses any other structure
    if (t->n_instr != RM_SPEC_N || tape_code_hash(t) != RM_SPEC_HASH)
        return fail(RM_ERR_INVALID, "this library is specialised for another tape (hash %08x)", (unsigned)RM_SPEC_HASH);
#endif
    return RM_OK;
}
// The kernels loop `iters` times over a grid of `blocks`; both are chosen so that every block runs
// the same (block-uniform) number of iterations and only the last block has idle slots.  Bounded
// grid: one global atomic per parameter per block, and the trajectory scratch is sized by it.
inline int backward_iters(int64_t n) {
    const int64_t want = (n + (int64_t)BLOCK_B * KB - 1) / ((int64_t)BLOCK_B * KB);
    const int64_t cap = 148 * 8;
    return (int)((want + cap - 1) / cap > 0 ? (want + cap - 1) / cap : 1);
}
inline unsigned backward_grid(int64_t n) {
    const int64_t want = (n + (int64_t)BLOCK_B * KB - 1) / ((int64_t)BLOCK_B * KB);
    const int64_t it = backward_iters(n);
    const int64_t blocks = (want + it - 1) / it;
    return (unsigned)(blocks > 0 ? blocks : 1);
}
// Accumulator layout for a tape (GradAcc): lane columns when n_par x BLOCK_B floats fit the budget,
// else warp columns, shared by all warps when even those do not fit.
inline int backward_acc_mode(const RmTape* t) {
    return (size_t)t->n_par * BLOCK_B * sizeof(float) <= RM_ACC_BYTES ? ACC_LANES : ACC_WARPS;
}
inline int backward_acc_shared(const RmTape* t) {
    return backward_acc_mode(t) == ACC_WARPS && (size_t)t->n_par * (BLOCK_B / 32) * sizeof(float) > RM_ACC_BYTES;
}
inline size_t backward_smem(const RmTape* t) {
    const size_t cols = backward_acc_mode(t) == ACC_LANES ? BLOCK_B : (backward_acc_shared(t) ? 1 : BLOCK_B / 32);
    return (size_t)t->n_par * sizeof(float) * (1 + cols);
}

// Picks the kernel instantiation: lane columns exist only for the small code capacity (a tape of
// <= 96 parameter floats never has more than 96 instructions).
#ifdef RM_SPEC_INC
// specialised library: one code capacity (the kernels read the compiled-in tape, not `code`)
#define RM_BWD_LAUNCH(tape, KERNEL, grid, sm, stream, ...)                                              \
    do {                                                                                                \
        rm::TapeCode<rm::CODE_CAP_SMALL> small;                                                         \
        rm::expand_code((tape)->code_host, (tape)->n_instr, small);                                     \
        if (rm::backward_acc_mode(tape) == rm::ACC_LANES) {                                             \
            if (int s_ = rm::prepare_smem(KERNEL<rm::CODE_CAP_SMALL, rm::ACC_LANES>, sm)) return s_;    \
            KERNEL<rm::CODE_CAP_SMALL, rm::ACC_LANES><<<grid, rm::BLOCK_B, sm, (cudaStream_t)stream>>>( \
                small, __VA_ARGS__);                                                                    \
        } else {                                                                                        \
            if (int s_ = rm::prepare_smem(KERNEL<rm::CODE_CAP_SMALL, rm::ACC_WARPS>, sm)) return s_;    \
            KERNEL<rm::CODE_CAP_SMALL, rm::ACC_WARPS><<<grid, rm::BLOCK_B, sm, (cudaStream_t)stream>>>( \
                small, __VA_ARGS__);                                                                    \
        }                                                                                               \
    } while (0)
#else
#define RM_BWD_LAUNCH(tape, KERNEL, grid, sm, stream, ...)                                              \
    RM_WITH_CODE(tape, {                                                                                \
        if (CAP == rm::CODE_CAP_SMALL && rm::backward_acc_mode(tape) == rm::ACC_LANES) {                \
            rm::TapeCode<rm::CODE_CAP_SMALL> small;                                                     \
            rm::expand_code((tape)->code_host, (tape)->n_instr, small);                             \
            if (int s_ = rm::prepare_smem(KERNEL<rm::CODE_CAP_SMALL, rm::ACC_LANES>, sm)) return s_;    \
            KERNEL<rm::CODE_CAP_SMALL, rm::ACC_LANES><<<grid, rm::BLOCK_B, sm, (cudaStream_t)stream>>>( \
                small, __VA_ARGS__);                                                                    \
        } else {                                                                                        \
            if (int s_ = rm::prepare_smem(KERNEL<CAP, rm::ACC_WARPS>, sm)) return s_;                   \
            KERNEL<CAP, rm::ACC_WARPS><<<grid, rm::BLOCK_B, sm, (cudaStream_t)stream>>>(code, __VA_ARGS__); \
        }                                                                                               \
    })
#endif

}  // namespace rm

using namespace rm;

extern "C" int64_t rm_bwd_scratch_floats(int64_t n_rays, int32_t steps) {
    if (n_rays <= 0 || steps <= 0) return 0;
    return (int64_t)steps * backward_grid(n_rays) * BLOCK_B * KB * 3;
}

// rm_render_loss_bwd: the trajectory plus (RM_LOSS_ONFLY) the node records of steps + 4 evaluations.
extern "C" int64_t rm_loss_scratch_floats(const RmTape* tape, int64_t n_rays, int32_t steps) {
    if (tape == nullptr || n_rays <= 0 || steps < 0) return 0;
    const LossLaunch L = loss_launch(tape, n_rays, steps);
    return L.traj_floats + L.nrec_floats + L.aux_floats;
}

extern "C" int rm_sdf_vjp(const RmTape* tape, const float* points, const float* grad_sdf, int64_t n,
                          float* dpoints, float* dpar, void* stream) {
    if (int s = check_backward_tape(tape)) return s;
    if (n < 0 || (dpar == nullptr && tape->n_par > 0) || (n > 0 && (points == nullptr || grad_sdf == nullptr)))
        return fail(RM_ERR_INVALID, "rm_sdf_vjp: null buffer or negative count");
    if (tape->n_par > 0)
        RM_CUDA(cudaMemsetAsync(dpar, 0, (size_t)tape->n_par * sizeof(float), (cudaStream_t)stream));
    if (n == 0) return RM_OK;
    const size_t sm = backward_smem(tape);
    RM_BWD_LAUNCH(tape, sdf_vjp_kernel, backward_grid(n), sm, stream, tape->par, tape->n_instr, tape->n_par,
                  backward_acc_shared(tape), backward_iters(n), points, grad_sdf, n, dpoints, dpar);
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}

extern "C" int rm_march_bwd(const RmTape* tape, const float* origin, const float* dirs,
                            const float* grad_pos, int64_t n, int32_t steps, float converge_tol,
                            float* traj_scratch, float* dorigin, float* ddirs, float* dpar, void* stream) {
    if (int s = check_backward_tape(tape)) return s;
    if (n < 0 || steps < 0 || !(converge_tol >= 0.f) || (dpar == nullptr && tape->n_par > 0) ||
        (n > 0 && (!origin || !dirs || !grad_pos || (steps > 0 && !traj_scratch))))
        return fail(RM_ERR_INVALID, "rm_march_bwd: null buffer or negative count");
    if (tape->n_par > 0)
        RM_CUDA(cudaMemsetAsync(dpar, 0, (size_t)tape->n_par * sizeof(float), (cudaStream_t)stream));
    if (n == 0) return RM_OK;
    const size_t sm = backward_smem(tape);
    RM_BWD_LAUNCH(tape, march_bwd_kernel, backward_grid(n), sm, stream, tape->par, tape->n_instr, tape->n_par,
                  backward_acc_shared(tape), backward_iters(n), origin, dirs, grad_pos, n, steps, converge_tol,
                  traj_scratch, dorigin, ddirs, dpar);
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}

// Argument checks and kernel constants shared by rm_render_bwd and rm_render_loss_bwd.
static int prepare_render_bwd(const char* who, const RmTape* tape, const RmCamera* cam, const float* poses,
                              const RmRenderArgs* a, float* traj_scratch, float* dpar, float* dposes,
                              bool want_grad, void* stream, RenderBwdParams* P) {
    if (int s = check_backward_tape(tape)) return s;
    if (int s = make_cam(cam, &P->cam)) return s;
    if (a == nullptr || poses == nullptr || (want_grad && (dposes == nullptr || (dpar == nullptr && tape->n_par > 0))))
        return fail(RM_ERR_INVALID, "%s: null argument", who);
    const int64_t total = (int64_t)cam->n_views * cam->width * cam->height;
    if (const char* why = check_shard(a->ray_begin, a->n_rays, a->strip_rays, a->strip_stride, total))
        return fail(RM_ERR_INVALID, "%s: %s", who, why);
    if (a->steps < 0 || !(a->normals_eps > 0.f) || !(a->bwd_converge_tol >= 0.f))
        return fail(RM_ERR_INVALID, "%s: negative steps, eps <= 0 or negative tolerance", who);
    if (a->bounces != 0)
        return fail(RM_ERR_UNSUPPORTED, "%s: reflected segments are not differentiable here", who);
    if (a->n_rays > 0 && a->steps > 0 && traj_scratch == nullptr)
        return fail(RM_ERR_INVALID, "%s: trajectory scratch is null", who);
    if (want_grad) {
        if (tape->n_par > 0)
            RM_CUDA(cudaMemsetAsync(dpar, 0, (size_t)tape->n_par * sizeof(float), (cudaStream_t)stream));
        RM_CUDA(cudaMemsetAsync(dposes, 0, (size_t)cam->n_views * RM_POSE_STRIDE * sizeof(float), (cudaStream_t)stream));
    }
    make_stencil(a->normals_eps, &P->st);
    P->ray_begin = a->ray_begin; P->n_rays = a->n_rays; P->steps = a->steps;
    P->strip_rays = a->strip_rays; P->strip_stride = a->strip_stride;
    P->acc_shared = backward_acc_shared(tape);
    P->iters = backward_iters(a->n_rays);
    P->tol = a->bwd_converge_tol;
    P->g = RmRenderGrads{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    P->traj = traj_scratch; P->dpar = dpar; P->dposes = dposes;
    P->target_lam = P->target_depth = nullptr; P->w_lam = P->w_depth = 0.f; P->loss_out = nullptr;
    P->want_grad = want_grad ? 1 : 0;
    P->nrec = nullptr; P->n_nodes = tape->n_nodes; P->patch_cols = 0; P->aux = nullptr; P->rev_grid = 0;
    return RM_OK;
}

extern "C" int rm_render_bwd(const RmTape* tape, const RmCamera* cam, const float* poses,
                             const RmRenderArgs* a, const RmRenderGrads* grads, float* traj_scratch,
                             float* dpar, float* dposes, void* stream) {
    RenderBwdParams P;
    if (grads == nullptr) return fail(RM_ERR_INVALID, "rm_render_bwd: null argument");
    if (int s = prepare_render_bwd("rm_render_bwd", tape, cam, poses, a, traj_scratch, dpar, dposes, true, stream, &P))
        return s;
    if (a->n_rays == 0) return RM_OK;
    P.g = *grads;
    const size_t sm = backward_smem(tape);
    RM_BWD_LAUNCH(tape, render_bwd_kernel, backward_grid(a->n_rays), sm, stream, tape->par, tape->n_instr,
                  tape->n_par, poses, P);
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}

extern "C" int rm_render_loss_bwd(const RmTape* tape, const RmCamera* cam, const float* poses,
                                  const RmRenderArgs* a, const RmLossArgs* loss, float* traj_scratch,
                                  float* dpar, float* dposes, void* stream) {
    RenderBwdParams P;
    if (loss == nullptr || loss->loss_out == nullptr)
        return fail(RM_ERR_INVALID, "rm_render_loss_bwd: null loss block or loss_out");
    if (a != nullptr && a->n_rays > 0 && loss->target_lambertian == nullptr && loss->target_depth == nullptr)
        return fail(RM_ERR_INVALID, "rm_render_loss_bwd: no target plane given");
    const bool want_grad = dpar != nullptr || dposes != nullptr;
    if (int s = prepare_render_bwd("rm_render_loss_bwd", tape, cam, poses, a, traj_scratch, dpar, dposes,
                                   want_grad, stream, &P))
        return s;
    RM_CUDA(cudaMemsetAsync(loss->loss_out, 0, 2 * sizeof(float), (cudaStream_t)stream));
    if (a->n_rays == 0) return RM_OK;
    const LossLaunch L = loss_launch(tape, a->n_rays, a->steps, want_grad);
    if (loss->scratch_floats < L.traj_floats + L.nrec_floats + L.aux_floats)
        return fail(RM_ERR_INVALID, "rm_render_loss_bwd: scratch of %lld floats, rm_loss_scratch_floats asks for %lld",
                    (long long)loss->scratch_floats, (long long)(L.traj_floats + L.nrec_floats + L.aux_floats));
    P.nrec = traj_scratch + L.traj_floats;
    P.aux = P.nrec + L.nrec_floats;
    P.iters = L.iters;
    P.rev_grid = (int)L.grid;
#if RM_LOSS_PATCHES
    {
        const int64_t Wd = cam->width;
        if (KB == 1 && Wd % 8 == 0 && a->ray_begin % Wd == 0 && a->n_rays % (4 * Wd) == 0 &&
            (a->strip_rays == 0 || a->strip_rays % (4 * Wd) == 0) && a->n_rays / 32 <= 0x7fffffff)
            P.patch_cols = (int)(Wd / 8);
    }
#endif
    P.target_lam = loss->target_lambertian; P.target_depth = loss->target_depth;
    P.w_lam = loss->weight_lambertian; P.w_depth = loss->weight_depth; P.loss_out = loss->loss_out;
    const size_t sm = backward_smem(tape);
    if (L.route == 3) {
        // replay: four rays per thread, the forward kernel's launch shape; then the reverse kernel
        const size_t sm_a = tape_smem_bytes(tape);
        const unsigned grid_a = (unsigned)((a->n_rays + (int64_t)BLOCK_A * KF - 1) / ((int64_t)BLOCK_A * KF));
        RM_WITH_CODE(tape, {
            if (int s_ = prepare_smem(loss_replay_kernel<CAP>, sm_a)) return s_;
            loss_replay_kernel<CAP><<<grid_a, BLOCK_A, sm_a, (cudaStream_t)stream>>>(code, tape->par, tape->n_instr,
                                                                                   tape->n_par, poses, P);
        });
        RM_CUDA(cudaGetLastError());
        RM_BWD_LAUNCH(tape, render_loss_reverse_kernel, L.grid, sm, stream, tape->par, tape->n_instr, tape->n_par, poses, P);
    } else if (L.route == 2) {
        RM_BWD_LAUNCH(tape, render_loss_onfly4_kernel, L.grid, sm, stream, tape->par, tape->n_instr, tape->n_par, poses, P);
    } else if (L.route == 1) {
        RM_BWD_LAUNCH(tape, render_loss_onfly_kernel, L.grid, sm, stream, tape->par, tape->n_instr, tape->n_par, poses, P);
    } else {
        RM_BWD_LAUNCH(tape, render_loss_bwd_kernel, L.grid, sm, stream, tape->par, tape->n_instr, tape->n_par, poses, P);
    }
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}

#ifndef RM_SPEC_INC      // the general library only: a tape-specialised library would hold a second copy of the array
// ---- rm_march_bwd_static: rm_march_bwd without any caller- or library-side allocation ---------------
// No counterpart in the reference; added for callers that cannot
// call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
// The trajectory scratch of the replay lives in a file-scope device array of a fixed, modest size
// (RM_MARCH_BWD_STATIC_MAX_FLOATS floats = 32 MiB; every GPU that loads this library holds it until the
// process exits).  In HOST code the array's name is only a stand-in for the symbol, not a device address:
// the address is fetched once per device, with that device current, through cudaGetSymbolAddress, and that
// pointer is what the kernel receives.  Every call on a device shares the array, which keeps the last
// call's contents: nothing in it has to start at zero (the replay writes every slot before the reverse pass
// reads it; the outputs dpar are cleared by a stream-ordered memset), but calls that OVERLAP on one GPU --
// different streams, or different host threads -- race on it: serialise them, or use rm_march_bwd.
#include <mutex>

namespace rm {
static __device__ __align__(256) float g_static_traj[RM_MARCH_BWD_STATIC_MAX_FLOATS];

constexpr int STATIC_MAX_DEVICES = 64;
static float* g_static_traj_ptr[STATIC_MAX_DEVICES];     // device address of g_static_traj, per device
static std::mutex g_static_traj_mutex;

// Set-up, once per device: the device address of the static scratch on the CURRENT device.
static int static_traj_pointer(float** out) {
    int dev = -1;
    RM_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= STATIC_MAX_DEVICES)
        return fail(RM_ERR_UNSUPPORTED, "rm_march_bwd_static: device ordinal %d outside 0..%d", dev, STATIC_MAX_DEVICES - 1);
    std::lock_guard<std::mutex> lock(g_static_traj_mutex);
    if (g_static_traj_ptr[dev] == nullptr) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_traj);
        if (e != cudaSuccess || p == nullptr)
            return fail(RM_ERR_CUDA, "rm_march_bwd_static: cudaGetSymbolAddress(g_static_traj) failed on device %d: %s",
                        dev, cudaGetErrorString(e));
        g_static_traj_ptr[dev] = static_cast<float*>(p);
    }
    *out = g_static_traj_ptr[dev];
    return RM_OK;
}
}  // namespace rm

extern "C" int rm_march_bwd_static(const RmTape* tape, const float* origin, const float* dirs,
                                   const float* grad_pos, int64_t n, int32_t steps, float converge_tol,
                                   float* dorigin, float* ddirs, float* dpar, void* stream) {
    if (int s = check_backward_tape(tape)) return s;
    if (n < 0 || steps < 0 || !(converge_tol >= 0.f) || (dpar == nullptr && tape->n_par > 0) ||
        (n > 0 && (!origin || !dirs || !grad_pos)))
        return fail(RM_ERR_INVALID, "rm_march_bwd_static: null buffer or negative count");
    // same launch geometry as rm_march_bwd, so the scratch need is rm_bwd_scratch_floats(n, steps)
    const int64_t need = n > 0 && steps > 0 ? (int64_t)steps * backward_grid(n) * BLOCK_B * KB * 3 : 0;
    if (need > (int64_t)RM_MARCH_BWD_STATIC_MAX_FLOATS)
        return fail(RM_ERR_UNSUPPORTED,
                    "rm_march_bwd_static: %lld rays x %d steps need %lld floats of trajectory scratch, the static "
                    "scratch holds %lld (RM_MARCH_BWD_STATIC_MAX_FLOATS); use rm_march_bwd with caller-provided scratch",
                    (long long)n, (int)steps, (long long)need, (long long)RM_MARCH_BWD_STATIC_MAX_FLOATS);
    float* traj = nullptr;
    if (int s = static_traj_pointer(&traj)) return s;       // set-up on the first call per device; no allocation
    if (tape->n_par > 0)
        RM_CUDA(cudaMemsetAsync(dpar, 0, (size_t)tape->n_par * sizeof(float), (cudaStream_t)stream));
    if (n == 0) return RM_OK;
    const size_t sm = backward_smem(tape);
    RM_BWD_LAUNCH(tape, march_bwd_kernel, backward_grid(n), sm, stream, tape->par, tape->n_instr, tape->n_par,
                  backward_acc_shared(tape), backward_iters(n), origin, dirs, grad_pos, n, steps, converge_tol,
                  traj, dorigin, ddirs, dpar);
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}
#endif  // RM_SPEC_INC
