// Register-resident tape interpreter: evaluates the scene SDF at K points per thread.
//
// Every lane of a warp executes the same tape instruction with the same parameter address
// (the reference evaluates the whole tree for every ray, transformations.py:69,92), so
// instruction fetch and parameter loads are shared-memory broadcasts and there is no opcode
// divergence; decoding one instruction is amortised over the K points a thread owns.
#pragma once
#include "rm_prims.cuh"

namespace rm {

// One accumulator per open union: hard union keeps the running min in a0; smooth union keeps
// an online log-sum-exp of x_i = c * f_i (c = -k, or -k*log2(e) in FAST mode): a0 = running
// max, a1 = sum of exp(x_i - a0).  Equals the reference's max-shifted logsumexp
// (transformations.py:69-70) up to rounding.
template <int K, bool FAST>
__device__ __forceinline__ void eval_scene(const TapeView& T, const float (&px)[K],
                                           const float (&py)[K], const float (&pz)[K],
                                           float (&out)[K]) {
    using M = Math<FAST>;
    float qx[K], qy[K], qz[K];
    float a0[K], a1[K], d[K];
    float c = 0.f;                                   // scale of the innermost open smooth union
    float stk[MAX_DEPTH * (2 * K + 1)];              // spilled accumulators of enclosing unions
    int sp = 0;
#pragma unroll
    for (int k = 0; k < K; ++k) { qx[k] = px[k]; qy[k] = py[k]; qz[k] = pz[k]; a0[k] = 0.f; a1[k] = 0.f; d[k] = 0.f; }

    for (int pc = 0; pc < T.n_instr; ++pc) {
        const int4 ins = T.code[pc];
        const int head = ins.x;
        const float4* P = T.par + ins.y;
        const int op = head_op(head);
        if (op < OP_OPEN_MIN) {
            const int fm = head_frame(head);
            if (fm == FRAME_NEW) {                   // q = A p + b  (composed affine chain)
                const float4 r0 = P[0], r1 = P[1], r2 = P[2];
                P += 3;
#pragma unroll
                for (int k = 0; k < K; ++k) {
                    qx[k] = fmaf(r0.x, px[k], fmaf(r0.y, py[k], fmaf(r0.z, pz[k], r0.w)));
                    qy[k] = fmaf(r1.x, px[k], fmaf(r1.y, py[k], fmaf(r1.z, pz[k], r1.w)));
                    qz[k] = fmaf(r2.x, px[k], fmaf(r2.y, py[k], fmaf(r2.z, pz[k], r2.w)));
                }
            } else if (fm == FRAME_WORLD) {
#pragma unroll
                for (int k = 0; k < K; ++k) { qx[k] = px[k]; qy[k] = py[k]; qz[k] = pz[k]; }
            }
            if (op == OP_SPHERE) {
                const float4 p0 = P[0];
#pragma unroll
                for (int k = 0; k < K; ++k) d[k] = prim_sphere<FAST>(p0, qx[k], qy[k], qz[k]);
            } else if (op == OP_BOX) {
                const float4 p0 = P[0];
#pragma unroll
                for (int k = 0; k < K; ++k) d[k] = prim_box<FAST>(p0, qx[k], qy[k], qz[k]);
            } else if (op == OP_TORUS) {
                const float4 p0 = P[0];
#pragma unroll
                for (int k = 0; k < K; ++k) d[k] = prim_torus<FAST>(p0, qx[k], qy[k], qz[k]);
            } else if (op == OP_LINE) {
                const float4 p0 = P[0], p1 = P[1], p2 = P[2];
#pragma unroll
                for (int k = 0; k < K; ++k) d[k] = prim_line<FAST>(p0, p1, p2, qx[k], qy[k], qz[k]);
            } else if (op == OP_DISK) {
                const float4 p0 = P[0];
#pragma unroll
                for (int k = 0; k < K; ++k) d[k] = prim_disk<FAST>(p0, qx[k], qy[k], qz[k]);
            } else {                                  // OP_PLANE: q.x   (primitives.py:40-41)
#pragma unroll
                for (int k = 0; k < K; ++k) d[k] = qx[k];
            }
            P += prim_f4(op);
        } else if (op == OP_OPEN_MIN || op == OP_OPEN_SMOOTH) {
            if (!(head & HEAD_ROOT)) {
                float* s = stk + sp * (2 * K + 1);
#pragma unroll
                for (int k = 0; k < K; ++k) { s[2 * k] = a0[k]; s[2 * k + 1] = a1[k]; }
                s[2 * K] = c;
                ++sp;
            }
            if (op == OP_OPEN_MIN) {
#pragma unroll
                for (int k = 0; k < K; ++k) a0[k] = CUDART_INF_F;
            } else {
                c = M::exp_scale(P[0].x);
#pragma unroll
                for (int k = 0; k < K; ++k) { a0[k] = -CUDART_INF_F; a1[k] = 0.f; }
            }
            continue;
        } else {                                      // CLOSE: finish this union, restore parent
            if (op == OP_CLOSE_MIN) {
#pragma unroll
                for (int k = 0; k < K; ++k) d[k] = a0[k];
            } else {
#pragma unroll
                for (int k = 0; k < K; ++k) d[k] = M::div(a0[k] + M::log(a1[k]), c);
            }
            if (!(head & HEAD_ROOT)) {
                --sp;
                const float* s = stk + sp * (2 * K + 1);
#pragma unroll
                for (int k = 0; k < K; ++k) { a0[k] = s[2 * k]; a1[k] = s[2 * k + 1]; }
                c = s[2 * K];
            }
        }
        // unary value ops attached to this instruction (rounding / onion)
        const int np = head_npost(head);
        if (np) {
            const float4 pp = P[0];
            const float pv[3] = {pp.x, pp.y, pp.z};
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                if (i < np) {
                    const bool onion = head_post(head, i) == POST_ONION;
#pragma unroll
                    for (int k = 0; k < K; ++k) d[k] = (onion ? fabsf(d[k]) : d[k]) - pv[i];
                }
            }
        }
        const int fold = head_fold(head);
        if (fold == FOLD_MIN) {
#pragma unroll
            for (int k = 0; k < K; ++k) a0[k] = min_nan(a0[k], d[k]);
        } else if (fold == FOLD_SMOOTH) {
#pragma unroll
            for (int k = 0; k < K; ++k) {
                const float x = c * d[k];
                const float e = M::exp(-fabsf(x - a0[k]));
                a1[k] = (x > a0[k]) ? fmaf(a1[k], e, 1.f) : a1[k] + e;
                a0[k] = fmaxf(a0[k], x);
            }
        }
    }
#pragma unroll
    for (int k = 0; k < K; ++k) out[k] = d[k];
}

// Cooperative copy of the tape from global into dynamic shared memory.
__device__ __forceinline__ TapeView stage_tape(const int32_t* code, const float* par, int n_instr,
                                               int n_par, unsigned char* smem) {
    int4* scode = reinterpret_cast<int4*>(smem);
    float4* spar = reinterpret_cast<float4*>(smem + (size_t)n_instr * sizeof(int4));
    const int4* gcode = reinterpret_cast<const int4*>(code);
    const float4* gpar = reinterpret_cast<const float4*>(par);
    for (int i = threadIdx.x; i < n_instr; i += blockDim.x) scode[i] = __ldg(gcode + i);
    for (int i = threadIdx.x; i < n_par / 4; i += blockDim.x) spar[i] = __ldg(gpar + i);
    __syncthreads();
    TapeView T;
    T.code = scode;
    T.par = spar;
    T.n_instr = n_instr;
    return T;
}

}  // namespace rm
