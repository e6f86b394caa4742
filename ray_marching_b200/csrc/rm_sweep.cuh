// Pass 2 of the scene VJP: reverse sweep over the tape (see rm_grad.cuh for the overview).
#pragma once
#include "rm_grad.cuh"

namespace rm {

// State of the union whose children are currently being visited (reverse order).
template <int K>
struct OpenNode {
    float a[K];       // adjoint of the union's raw value
    float d[K];       // raw value (pass-1 record)
    int win[K];       // winning ordinal (hard union)
    float s[K];       // sum_c a * w_c * y_c over visited children (for d/dk of a smooth union)
    float c;          // -k
};

// Applies the post-op chain forward (keeping the stage inputs' signs), then maps the adjoint of
// the post-op'd value back to the raw value and accumulates -adjoint into the post parameters.
template <int K>
__device__ __forceinline__ void posts_forward(int head, const float (&pv)[3], const float (&raw)[K],
                                              float (&y)[K], float (&sg)[3][K]) {
    const int np = head_npost(head);
#pragma unroll
    for (int k = 0; k < K; ++k) y[k] = raw[k];
#pragma unroll
    for (int i = 0; i < 3; ++i)
        if (i < np) {
            const bool onion = head_post(head, i) == POST_ONION;
#pragma unroll
            for (int k = 0; k < K; ++k) {
                sg[i][k] = onion ? sgn0(y[k]) : 1.f;
                y[k] = (onion ? fabsf(y[k]) : y[k]) - pv[i];
            }
        }
}
template <int K>
__device__ __forceinline__ void posts_backward(int head, const float (&sg)[3][K], float (&a)[K],
                                               float (&gpost)[3]) {
    const int np = head_npost(head);
#pragma unroll
    for (int i = 2; i >= 0; --i)
        if (i < np) {
#pragma unroll
            for (int k = 0; k < K; ++k) { gpost[i] -= a[k]; a[k] *= sg[i][k]; }
        }
}

// Adjoint of a child's (post-op'd) value y given the enclosing union's state; also feeds the
// d/dk accumulator of a smooth parent.
template <int K>
__device__ __forceinline__ void child_adjoint(int fold, int ordinal, OpenNode<K>& par,
                                              const float (&root_abar)[K], const float (&y)[K],
                                              float (&a)[K]) {
#pragma unroll
    for (int k = 0; k < K; ++k) {
        if (fold == FOLD_NONE) a[k] = root_abar[k];
        else if (fold == FOLD_MIN) a[k] = (par.win[k] == ordinal) ? par.a[k] : 0.f;
        else {
            // softmax weight exp(-k (y - d_U)); 0 * inf guards keep inactive samples inert
            const float w = expf(par.c * (y[k] - par.d[k]));
            a[k] = par.a[k] != 0.f ? par.a[k] * w : 0.f;
            par.s[k] = fmaf(a[k], y[k], par.s[k]);
        }
    }
}

// The reverse sweep.  `sacc` = per-block shared accumulators, one float per tape parameter.
// Outputs g{x,y,z}[k] = abar[k] * grad_p f(p_k).  All 32 lanes of the warp must call this together.
template <int K, int CAP>
__device__ __forceinline__ void sweep_vjp(const TapeCode<CAP>& C, const int n_instr,
                                          const float4* __restrict__ spar, const float (&px)[K],
                                          const float (&py)[K], const float (&pz)[K],
                                          const float (&abar)[K], const NodeRecords<K>& rec,
                                          float* sacc, float (&gx)[K], float (&gy)[K],
                                          float (&gz)[K]) {
    const int lane = threadIdx.x & 31;
    OpenNode<K> cur;
    OpenNode<K> stk[MAX_DEPTH];
    int sp = 0;
#pragma unroll
    for (int k = 0; k < K; ++k) { gx[k] = gy[k] = gz[k] = 0.f; cur.a[k] = cur.d[k] = cur.s[k] = 0.f; cur.win[k] = -1; }
    cur.c = 0.f;

    for (int pc = n_instr - 1; pc >= 0; --pc) {
        const int2 ins = C.ins[pc];
        const int head = ins.x, op = head_op(head);
        const int fold = head_fold(head), np = head_npost(head);
        const int off = word_par(ins.y), node = word_node(ins.y), ordinal = head_ordinal(head);
        if (op == OP_OPEN_MIN || op == OP_OPEN_SMOOTH) {
            // all children of this union have been visited: finish d/dk, return to the parent
            if (op == OP_OPEN_SMOOTH) {
                float g = 0.f;
#pragma unroll
                for (int k = 0; k < K; ++k) g += cur.a[k] != 0.f ? __fdiv_rn(cur.s[k] - cur.a[k] * cur.d[k], cur.c) : 0.f;
                g = warp_sum(g);
                if (lane == 0 && g != 0.f) atomicAdd(sacc + off * 4, g);
            }
            if (!(head & HEAD_ROOT)) cur = stk[--sp];
            continue;
        }
        float raw[K], y[K], a[K], sg[3][K];
        float gpost[3] = {0.f, 0.f, 0.f};
        if (op >= OP_CLOSE_MIN) {
            // ---- union close: value from the records, descend into the node ----
            const float4 pp = np ? spar[off] : make_float4(0.f, 0.f, 0.f, 0.f);
            const float pv[3] = {pp.x, pp.y, pp.z};
#pragma unroll
            for (int k = 0; k < K; ++k) raw[k] = rec.d[node][k];
            posts_forward<K>(head, pv, raw, y, sg);
            child_adjoint<K>(fold, ordinal, cur, abar, y, a);
            posts_backward<K>(head, sg, a, gpost);
            if (np) {
#pragma unroll
                for (int i = 0; i < 3; ++i)
                    if (i < np) {
                        const float g = warp_sum(gpost[i]);
                        if (lane == 0 && g != 0.f) atomicAdd(sacc + off * 4 + i, g);
                    }
            }
            if (!(head & HEAD_ROOT)) stk[sp++] = cur;
#pragma unroll
            for (int k = 0; k < K; ++k) {
                cur.a[k] = a[k]; cur.d[k] = raw[k]; cur.s[k] = 0.f;
                cur.win[k] = op == OP_CLOSE_MIN ? rec.win[node][k] : -1;
            }
            cur.c = op == OP_CLOSE_SMOOTH ? rec.c[node] : 0.f;
            continue;
        }
        // ---- leaf: recompute q, value and local gradients --------------------------------
        const int frame_rec = (head & HEAD_FRAME) ? off : -1;
        const int prim_rec = off + ((head & HEAD_FRAME) ? 3 : 0);
        const int post_rec = prim_rec + prim_f4(op);
        float4 r0 = make_float4(1.f, 0.f, 0.f, 0.f), r1 = make_float4(0.f, 1.f, 0.f, 0.f),
               r2 = make_float4(0.f, 0.f, 1.f, 0.f);
        if (frame_rec >= 0) { r0 = spar[frame_rec]; r1 = spar[frame_rec + 1]; r2 = spar[frame_rec + 2]; }
        const float4 pp = np ? spar[post_rec] : make_float4(0.f, 0.f, 0.f, 0.f);
        const float pv[3] = {pp.x, pp.y, pp.z};
        float gq[K][3], gp[K][12];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const float qx = fmaf(r0.x, px[k], fmaf(r0.y, py[k], fmaf(r0.z, pz[k], r0.w)));
            const float qy = fmaf(r1.x, px[k], fmaf(r1.y, py[k], fmaf(r1.z, pz[k], r1.w)));
            const float qz = fmaf(r2.x, px[k], fmaf(r2.y, py[k], fmaf(r2.z, pz[k], r2.w)));
            raw[k] = prim_grad(op, spar + prim_rec, qx, qy, qz, gq[k], gp[k]);
        }
        posts_forward<K>(head, pv, raw, y, sg);
        child_adjoint<K>(fold, ordinal, cur, abar, y, a);
        posts_backward<K>(head, sg, a, gpost);
        // a[k] is now the adjoint of the leaf's raw value.  Record layout of the 32 partials:
        // [0,12) frame rows (A_r0 A_r1 A_r2 b_r), [12,24) primitive record, [24,27) post params.
        float acc[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) acc[i] = 0.f;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const float ak = a[k];
            const float g0 = ak * gq[k][0], g1 = ak * gq[k][1], g2 = ak * gq[k][2];
            // world-space gradient: A^T (a gq)
            gx[k] += fmaf(r0.x, g0, fmaf(r1.x, g1, r2.x * g2));
            gy[k] += fmaf(r0.y, g0, fmaf(r1.y, g1, r2.y * g2));
            gz[k] += fmaf(r0.z, g0, fmaf(r1.z, g1, r2.z * g2));
            acc[0] = fmaf(g0, px[k], acc[0]); acc[1] = fmaf(g0, py[k], acc[1]); acc[2] = fmaf(g0, pz[k], acc[2]); acc[3] += g0;
            acc[4] = fmaf(g1, px[k], acc[4]); acc[5] = fmaf(g1, py[k], acc[5]); acc[6] = fmaf(g1, pz[k], acc[6]); acc[7] += g1;
            acc[8] = fmaf(g2, px[k], acc[8]); acc[9] = fmaf(g2, py[k], acc[9]); acc[10] = fmaf(g2, pz[k], acc[10]); acc[11] += g2;
#pragma unroll
            for (int i = 0; i < 12; ++i) acc[12 + i] = fmaf(ak, gp[k][i], acc[12 + i]);
        }
        acc[24] = gpost[0]; acc[25] = gpost[1]; acc[26] = gpost[2];
        const float tot = warp_transpose_sum(acc, lane);
        int addr = -1;
        if (lane < 12) addr = frame_rec >= 0 ? frame_rec * 4 + lane : -1;
        else if (lane < 24) addr = (lane - 12) < 4 * prim_f4(op) ? prim_rec * 4 + (lane - 12) : -1;
        else if (lane < 27) addr = (lane - 24) < np ? post_rec * 4 + (lane - 24) : -1;
        if (addr >= 0 && tot != 0.f) atomicAdd(sacc + addr, tot);
    }
}

}  // namespace rm
