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

// Shared-memory gradient accumulators.  Shared-memory float atomicAdd is a compare-and-swap loop
// on this architecture (LDS, FADD, ATOMS.CAST.SPIN, BRA), so both layouts avoid atomics:
//   ACC_LANES  one column per THREAD of the block: element (addr, thread).  A lane adds with a
//              plain read-modify-write into its own word (conflict-free banks, no reduction at
//              all).  n_par x block floats: tapes up to 96 parameter floats (the registry scenes).
//   ACC_WARPS  one column per WARP: the 27 partials of a leaf (12 frame, <= 12 primitive, <= 3
//              post-op) are summed across the warp by a transposing butterfly (31 shuffles; lane L
//              ends up with the total of value L) and lane L adds it to row addr(L) of its warp's
//              column.  n_par x warps floats; tapes too large even for that share ONE column and
//              lane L uses atomicAdd (`shared`).
// flush_accumulators (rm_backward.cu) sums the columns once per block.
enum : int { ACC_LANES = 0, ACC_WARPS = 1 };
template <int MODE>
struct GradAcc {
    float* base;
    int n_par;
    bool shared;      // ACC_WARPS: all warps add into one column with atomics

    __device__ __forceinline__ void put(int addr, float v) const {       // ACC_WARPS, one lane per row
        if (shared) atomicAdd(base + addr, v);
        else base[(threadIdx.x >> 5) * n_par + addr] += v;
    }
    // One value per lane.  ACC_WARPS: every lane of the warp must call.
    __device__ __forceinline__ void add1(int addr, float v) const {
        if (MODE == ACC_LANES) {
            if (v != 0.f) base[addr * blockDim.x + threadIdx.x] += v;
        } else {
            v = warp_sum(v);
            if ((threadIdx.x & 31) == 0 && v != 0.f) put(addr, v);
        }
    }
    // The partials of one leaf: frame rows at `frame_row` (when has_frame), n_prim primitive floats at
    // `prim_row`, np post-op floats at `post_row` (rows in floats).  ACC_WARPS: every lane must call.
    __device__ __forceinline__ void add_leaf(bool has_frame, int frame_row, int n_prim, int prim_row, int np,
                                             int post_row, const float (&fr)[12], const float (&pr)[12],
                                             const float (&gpost)[3]) const {
        if (MODE == ACC_LANES) {
            float* col = base + threadIdx.x;
            const int stride = blockDim.x;
            if (has_frame) {
#pragma unroll
                for (int i = 0; i < 12; ++i)
                    if (fr[i] != 0.f) col[(frame_row + i) * stride] += fr[i];
            }
#pragma unroll
            for (int i = 0; i < 12; ++i)
                if (i < n_prim && pr[i] != 0.f) col[(prim_row + i) * stride] += pr[i];
#pragma unroll
            for (int i = 0; i < 3; ++i)
                if (i < np && gpost[i] != 0.f) col[(post_row + i) * stride] += gpost[i];
        } else {
            const int lane = threadIdx.x & 31;
            float v[32];
#pragma unroll
            for (int i = 0; i < 12; ++i) { v[i] = fr[i]; v[12 + i] = pr[i]; }
#pragma unroll
            for (int i = 0; i < 3; ++i) v[24 + i] = gpost[i];
#pragma unroll
            for (int i = 27; i < 32; ++i) v[i] = 0.f;
            const float tot = warp_transpose_sum(v, lane);
            const int j = lane < 12 ? lane : (lane < 24 ? lane - 12 : lane - 24);
            const bool valid = lane < 12 ? has_frame : (lane < 24 ? j < n_prim : (lane < 27 && j < np));
            const int row = lane < 12 ? frame_row : (lane < 24 ? prim_row : post_row);
            if (valid && tot != 0.f) put(row + j, tot);
        }
    }
};

// Softmax weights below this are treated as zero (smooth-union children far behind the surface);
// a 512-child union then loses at most 512e-9 of the gradient, far inside the 1e-3 tolerance.
#ifndef RM_SWEEP_DEAD_SKIP
#define RM_SWEEP_DEAD_SKIP 1
#endif
#ifndef RM_LIVE_INLINE
#define RM_LIVE_INLINE __noinline__
#endif
constexpr float CULL_EPS = 1e-9f;
constexpr float LOG_CULL_EPS = -20.723266f;        // ln(1e-9)

// ONFLY: is this leaf child of a smooth union live for any sample of this thread?  The child's value is
// evaluated here, value-only (its own copies of the frame: nothing is kept for the gradient path).
// Kept OUT OF LINE on purpose.  Inlined into the sweep, nvcc 12.9 / ptxas (-O3, -O2 and -O1 alike, depending on
// unrelated code around it) produced a loss kernel whose far children of a smooth union received large
// gradients -- deterministic, caught by the reference fixtures (tests/test_gpu_backward.py), gone with a
// printf in the loop or with this call boundary; the cause was not found in the SASS.  The call costs ~2 % of
// the sweep.  Do not change this without running the gradient fixtures on the GPU.
template <int K, bool FASTM>
__device__ RM_LIVE_INLINE bool smooth_leaf_live(const float4* __restrict__ spar, int head, int op, int np, int off,
                                                bool has_frame, int prim_rec, int post_rec, const float (&px)[K],
                                                const float (&py)[K], const float (&pz)[K], const OpenNode<K>& cur) {
    float4 f0 = make_float4(1.f, 0.f, 0.f, 0.f), f1 = make_float4(0.f, 1.f, 0.f, 0.f), f2 = make_float4(0.f, 0.f, 1.f, 0.f);
    if (has_frame) { f0 = spar[off]; f1 = spar[off + 1]; f2 = spar[off + 2]; }
    const float4 pq = np ? spar[post_rec] : make_float4(0.f, 0.f, 0.f, 0.f);
    const float pvv[3] = {pq.x, pq.y, pq.z};
    bool live = false;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        const float qx = fmaf(f0.x, px[k], fmaf(f0.y, py[k], fmaf(f0.z, pz[k], f0.w)));
        const float qy = fmaf(f1.x, px[k], fmaf(f1.y, py[k], fmaf(f1.z, pz[k], f1.w)));
        const float qz = fmaf(f2.x, px[k], fmaf(f2.y, py[k], fmaf(f2.z, pz[k], f2.w)));
        float yk = prim_value<FASTM>(op, spar + prim_rec, qx, qy, qz);
        if (np) {                                          // (uniform; most leaves have no post-op)
#pragma unroll
            for (int i = 0; i < 3; ++i)
                if (i < np) yk = (head_post(head, i) == POST_ONION ? fabsf(yk) : yk) - pvv[i];
        }
        live |= cur.a[k] != 0.f && cur.c * (yk - cur.d[k]) >= LOG_CULL_EPS;
    }
    return live;
}

// The reverse sweep.  Outputs g{x,y,z}[k] = abar[k] * grad_p f(p_k) and adds d f / d par into
// `acc`.  All 32 lanes of the warp must call this together (warp votes skip leaves whose adjoint
// is zero in every lane: the losing children of a hard union, culled children of a smooth one).
// ONFLY: the records hold the union nodes only (d, win, c); a smooth-union child's value is evaluated here,
// value-only (FASTM = the math policy of the evaluation that produced the node records), for the liveness test
// -- one walk over the tape per reverse task instead of a recording walk plus this one.
template <int K, int CAP, int MODE, bool ONFLY = false, bool FASTM = false, class REC = NodeRecords<K, CAP>>
__device__ __forceinline__ void sweep_vjp(const TapeCode<CAP>& C, const int n_instr,
                                          const float4* __restrict__ spar, const float (&px)[K],
                                          const float (&py)[K], const float (&pz)[K],
                                          const float (&abar)[K], const REC& rec,
                                          const GradAcc<MODE> acc, float (&gx)[K], float (&gy)[K],
                                          float (&gz)[K]) {
    OpenNode<K> cur;
    OpenNode<K> stk[MAX_DEPTH];
    int sp = 0;
    // ONFLY, inside the culled union (tape.py: sphere table): children that are dead for EVERY ray of the warp
    // are found 32 at a time from the table and jumped over without being decoded.  All warp-uniform.
    bool dead_on = false;
    unsigned dead_mask = 0u;
    int mask_top = 0x40000000, tbl_pc0 = 0;
    const float4* tbl = spar;
    float cwx = 0.f, cwy = 0.f, cwz = 0.f, rw = 0.f, dead_cs = 0.f, dead_thr = 0.f;
#pragma unroll
    for (int k = 0; k < K; ++k) { gx[k] = gy[k] = gz[k] = 0.f; cur.a[k] = cur.d[k] = cur.s[k] = 0.f; cur.win[k] = -1; }
    cur.c = 0.f;

    RM_TAPE_LOOP
    for (int pc = RM_TAPE_LEN(n_instr) - 1; pc >= 0; --pc) {
        if (ONFLY && RM_SWEEP_DEAD_SKIP && dead_on) {
            int off_m = mask_top - pc;
            if (off_m >= 32) {
                // lane L tests body instruction pc - L: outside its sphere by t, its value is >= s_min t, its softmax
                // weight <= exp(c (s_min t - d_U)) for every ray with a live adjoint (dead_thr holds the rest)
                const int q = pc - (int)(threadIdx.x & 31);
                bool dead = false;
                if (q >= tbl_pc0) {
                    const float4 cr = tbl[q - tbl_pc0];
                    const float dx = cwx - cr.x, dy = cwy - cr.y, dz = cwz - cr.z;
                    const float t = sqrt_rn_inline(fmaf(dx, dx, fmaf(dy, dy, dz * dz))) - cr.w - rw;
                    dead = t > 0.f && dead_cs * t < dead_thr;                         // false for NaN
                }
                dead_mask = __reduce_or_sync(0xffffffffu, dead ? (1u << (threadIdx.x & 31)) : 0u);
                mask_top = pc;
                off_m = 0;
            }
            const unsigned m = dead_mask >> off_m;
            if (m & 1u) {
                const int run = (~m) != 0u ? __ffs((int)~m) - 1 : 32;
                pc -= run - 1;                                                        // (the loop takes the last one)
                continue;
            }
        }
        const int4 ins = tape_ins(C, pc);
        const int head = ins.x, op = head_op(head);
        const int fold = head_fold(head), np = head_npost(head);
        const int off = word_par(ins.y), node = word_node(ins.y), ordinal = head_ordinal(head);
        if (op == OP_OPEN_MIN || op == OP_OPEN_SMOOTH) {
            // all children of this union have been visited: finish d/dk, return to the parent
            if (ONFLY && op == OP_OPEN_SMOOTH && word_cull(ins.y)) dead_on = false;
            if (op == OP_OPEN_SMOOTH) {
                float g = 0.f;
#pragma unroll
                for (int k = 0; k < K; ++k) g += cur.a[k] != 0.f ? div_rn_inline(cur.s[k] - cur.a[k] * cur.d[k], cur.c) : 0.f;
                acc.add1(off * 4, g);
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
#pragma unroll
            for (int i = 0; i < 3; ++i)
                if (i < np) acc.add1(off * 4 + i, gpost[i]);
            if (!(head & HEAD_ROOT)) stk[sp++] = cur;
#pragma unroll
            for (int k = 0; k < K; ++k) {
                cur.a[k] = a[k]; cur.d[k] = raw[k]; cur.s[k] = 0.f;
                cur.win[k] = op == OP_CLOSE_MIN ? rec.win[node][k] : -1;
            }
            cur.c = op == OP_CLOSE_SMOOTH ? rec.c[node] : 0.f;
            if constexpr (ONFLY) {
                if (RM_SWEEP_DEAD_SKIP && op == OP_CLOSE_SMOOTH && word_cull(ins.y)) {
                    const int open_pc = ins.z;                         // rm_tape.h: expand_instructions
                    const int4 oi = tape_ins(C, open_pc);
                    const float4 u = spar[word_par(oi.y)];             // (-k, s_min, L_max, log2 n)
                    // the warp's bounding sphere and the largest union value among the rays with a live adjoint
                    cwx = __shfl_sync(0xffffffffu, px[0], 16);
                    cwy = __shfl_sync(0xffffffffu, py[0], 16);
                    cwz = __shfl_sync(0xffffffffu, pz[0], 16);
                    float r2 = 0.f, dmax = -CUDART_INF_F;
#pragma unroll
                    for (int k = 0; k < K; ++k) {
                        const float dx = px[k] - cwx, dy = py[k] - cwy, dz = pz[k] - cwz;
                        const float q2 = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
                        r2 = q2 > r2 ? q2 : (q2 == q2 ? r2 : CUDART_INF_F);          // NaN point: nothing is skipped
                        if (cur.a[k] != 0.f) dmax = cur.d[k] > dmax ? cur.d[k] : (cur.d[k] == cur.d[k] ? dmax : CUDART_INF_F);
                    }
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) {
                        r2 = fmaxf(r2, __shfl_xor_sync(0xffffffffu, r2, o));
                        dmax = fmaxf(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
                    }
                    rw = sqrt_rn_inline(r2) * 1.0001f;
                    // dead for the warp:  c (s_min t - d_max) < ln CULL_EPS  <=>  (c s_min) t < ln CULL_EPS + c d_max,
                    // with 0.01 of slack for the rounding of both sides (the per-lane test below stays the rule)
                    dead_cs = cur.c * u.y;
                    dead_thr = LOG_CULL_EPS - 0.01f + cur.c * dmax;
                    tbl = spar + head_ordinal(oi.x);
                    tbl_pc0 = open_pc + 1;
                    mask_top = 0x40000000;                              // first child: compute a mask
                    const bool usable = u.w > 0.f && u.y > 0.f && rw < CUDART_INF_F && dmax < CUDART_INF_F;
                    dead_on = __reduce_and_sync(0xffffffffu, usable ? 1u : 0u) != 0u;
                }
            }
            continue;
        }
        // ---- leaf ------------------------------------------------------------------------
        // cheap pre-test FIRST (most leaves of a large smooth union are dead: nothing is loaded for them):
        // is any lane's adjoint of this child non-zero?
        const bool has_frame = (head & HEAD_FRAME) != 0;
        const int prim_rec = off + (has_frame ? 3 : 0);
        const int post_rec = prim_rec + prim_f4(op);
        bool live = false;
        if (fold == FOLD_MIN) {
#pragma unroll
            for (int k = 0; k < K; ++k) live |= (cur.win[k] == ordinal) && cur.a[k] != 0.f;
        } else if (fold == FOLD_SMOOTH) {
            // softmax weight exp(c (y - d_U)) >= CULL_EPS  <=>  c (y - d_U) >= ln CULL_EPS (no exp)
            if constexpr (ONFLY) {
                live = smooth_leaf_live<K, FASTM>(spar, head, op, np, off, has_frame, prim_rec, post_rec, px, py, pz, cur);
            } else {
                // ... from the child's value recorded by pass 1 (no re-evaluation)
#pragma unroll
                for (int k = 0; k < K; ++k)
                    live |= cur.a[k] != 0.f && cur.c * (rec.y[pc][k] - cur.d[k]) >= LOG_CULL_EPS;
            }
        } else {
#pragma unroll
            for (int k = 0; k < K; ++k) live |= abar[k] != 0.f;
        }
        if (!__any_sync(0xffffffffu, live)) continue;

        float4 r0 = make_float4(1.f, 0.f, 0.f, 0.f), r1 = make_float4(0.f, 1.f, 0.f, 0.f),
               r2 = make_float4(0.f, 0.f, 1.f, 0.f);
        if (has_frame) { r0 = spar[off]; r1 = spar[off + 1]; r2 = spar[off + 2]; }
        const float4 pp = np ? spar[post_rec] : make_float4(0.f, 0.f, 0.f, 0.f);
        const float pv[3] = {pp.x, pp.y, pp.z};

        // value + local gradients, one sample at a time (keeps the 15 local partials of a sample
        // short-lived); per-thread sums over the K samples of every parameter partial
        float fr[12];                         // d/d(frame rows): A_r0 b_0 | A_r1 b_1 | A_r2 b_2
        float pr[12];                         // d/d(primitive record)
#pragma unroll
        for (int i = 0; i < 12; ++i) { fr[i] = 0.f; pr[i] = 0.f; }
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const float qx = fmaf(r0.x, px[k], fmaf(r0.y, py[k], fmaf(r0.z, pz[k], r0.w)));
            const float qy = fmaf(r1.x, px[k], fmaf(r1.y, py[k], fmaf(r1.z, pz[k], r1.w)));
            const float qz = fmaf(r2.x, px[k], fmaf(r2.y, py[k], fmaf(r2.z, pz[k], r2.w)));
            float gq[3], gp[12];
            const float rawk = prim_grad(op, spar + prim_rec, qx, qy, qz, gq, gp);
            // post-ops forward (stage signs), adjoint of the child's value, post-ops backward
            float yk = rawk, sgk[3] = {1.f, 1.f, 1.f};
#pragma unroll
            for (int i = 0; i < 3; ++i)
                if (i < np) {
                    const bool onion = head_post(head, i) == POST_ONION;
                    sgk[i] = onion ? sgn0(yk) : 1.f;
                    yk = (onion ? fabsf(yk) : yk) - pv[i];
                }
            float ak;
            if (fold == FOLD_NONE) ak = abar[k];
            else if (fold == FOLD_MIN) ak = (cur.win[k] == ordinal) ? cur.a[k] : 0.f;
            else {
                const float w = expf(cur.c * (yk - cur.d[k]));
                ak = (cur.a[k] != 0.f && w >= CULL_EPS) ? cur.a[k] * w : 0.f;
                cur.s[k] = fmaf(ak, yk, cur.s[k]);
            }
#pragma unroll
            for (int i = 2; i >= 0; --i)
                if (i < np) { gpost[i] -= ak; ak *= sgk[i]; }
            const float g0 = ak * gq[0], g1 = ak * gq[1], g2 = ak * gq[2];
            // world-space gradient: A^T (a gq)
            gx[k] += fmaf(r0.x, g0, fmaf(r1.x, g1, r2.x * g2));
            gy[k] += fmaf(r0.y, g0, fmaf(r1.y, g1, r2.y * g2));
            gz[k] += fmaf(r0.z, g0, fmaf(r1.z, g1, r2.z * g2));
            fr[0] = fmaf(g0, px[k], fr[0]); fr[1] = fmaf(g0, py[k], fr[1]); fr[2] = fmaf(g0, pz[k], fr[2]); fr[3] += g0;
            fr[4] = fmaf(g1, px[k], fr[4]); fr[5] = fmaf(g1, py[k], fr[5]); fr[6] = fmaf(g1, pz[k], fr[6]); fr[7] += g1;
            fr[8] = fmaf(g2, px[k], fr[8]); fr[9] = fmaf(g2, py[k], fr[9]); fr[10] = fmaf(g2, pz[k], fr[10]); fr[11] += g2;
#pragma unroll
            for (int i = 0; i < 12; ++i) pr[i] = fmaf(ak, gp[i], pr[i]);
        }
        acc.add_leaf(has_frame, off * 4, 4 * prim_f4(op), prim_rec * 4, np, post_rec * 4, fr, pr, gpost);
    }
}

}  // namespace rm
