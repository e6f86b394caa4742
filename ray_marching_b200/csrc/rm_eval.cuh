// Register-resident tape interpreter: evaluates the scene SDF at K points per thread.
//
// The reference evaluates the whole tree for every ray (transformations.py:69,92), so every lane
// of a warp executes the same tape instruction.  The instruction words come from the kernel
// parameter block (rm_tape.h): the program counter, the decode and every branch below are
// warp-uniform and run on the uniform datapath; parameter records are shared-memory broadcasts;
// the cost of one decode is amortised over the K points a thread owns.
#pragma once
#include "rm_prims.cuh"

namespace rm {

// Leaf distance at the frame-local point (x, y, z); `op` is warp-uniform.
template <bool FAST>
__device__ __forceinline__ float prim_eval(int op, const float4* P, float x, float y, float z) {
    switch (op) {
        case OP_SPHERE: return prim_sphere<FAST>(P[0], x, y, z);
        case OP_BOX: return prim_box<FAST>(P[0], x, y, z);
        case OP_TORUS: return prim_torus<FAST>(P[0], x, y, z);
        case OP_LINE: return prim_line<FAST>(P[0], P[1], P[2], x, y, z);
        case OP_DISK: return prim_disk<FAST>(P[0], x, y, z);
        default: return x;                                 // OP_PLANE   (primitives.py:40-41)
    }
}

// One accumulator per open union: hard union keeps the running min in a0; smooth union keeps
// an online log-sum-exp of x_i = c * f_i (c = -k, or -k*log2(e) in FAST mode): a0 = running
// max, a1 = sum of exp(x_i - a0).  Equals the reference's max-shifted logsumexp
// (transformations.py:69-70) up to rounding.
//
// Decode per instruction: one 16-byte constant load (prefetched one instruction ahead), one add
// for the record address, a switch on the dispatch id, a switch on the tail selector.
//
// CULL (march steps only): the children of one designated smooth union carry a
// bounding sphere (world centre c, radius R with f_i(p) >= s_min (|p - c| - R) outside it).  A child
// is skipped when, for every ray of the warp, its term is below 2^-30 of a LOWER bound of the
// union's largest term: thr = max(running max, c_k (h + L m) - log2 n), where h is the union's value
// at the previous march position, m the distance moved since and L the Lipschitz bound of the scene
// (max term >= logsumexp - log n, and f_U(p') <= f_U(p) + L |p' - p|).  `hint` carries h in and
// the union's value at this point out; `move` is m (+inf disables the seed).
constexpr float CULL_LOG2 = -30.0f;

// RECORD (backward kernels): per union node, its raw value (before post-ops) and -- hard unions --
// the winning child ordinal (first index on ties, as torch.min(dim)), kept in local memory and
// touched only at union boundaries; the reverse sweep (rm_sweep.cuh) consumes them.
constexpr int MAX_NODES = 16;      // union nodes supported by the backward records
template <int K, int CAP>
struct NodeRecords {
    float d[MAX_NODES][K];     // raw value of the union (before its post-ops)
    int win[MAX_NODES][K];     // winning child ordinal (hard unions)
    float c[MAX_NODES];        // -k of smooth unions
    float y[CAP][K];           // per instruction: value a child of a smooth union fed into the blend
};

// Tape-specialised build (ray_marching_b200/specialise.py): the translation unit is compiled once
// per scene STRUCTURE with RM_SPEC_INC naming a generated header that defines
//     template <int K, bool FAST> eval_spec(spar, px, py, pz, out)
// -- the straight-line sequence of exactly the statements the interpreter below executes for that
// tape (same primitive bodies, same fold order, constant record offsets) -- and the non-recording
// interpreter call sites use it instead of decoding.  Parameters still come from the staged block.
//
// The RECORDING interpreter and the reverse sweep (backward kernels) are specialised by the compiler
// instead: the generated header also carries the tape words as a switch (spec_word), the tape loops
// below fetch through tape_ins() and are fully unrolled, so every decode, dispatch and stack index
// constant-folds and what remains is the straight-line code of that tape.
#ifdef RM_SPEC_INC
}  // namespace rm
#include RM_SPEC_INC
namespace rm {
template <int CAP>
__device__ __forceinline__ int4 tape_ins(const TapeCode<CAP>&, int pc) {
    return expand_one(spec_word(pc, 0), spec_word(pc, 1), false);
}
#define RM_TAPE_LEN(n_instr) RM_SPEC_N
#define RM_TAPE_LOOP _Pragma("unroll")
#else
template <int CAP>
__device__ __forceinline__ int4 tape_ins(const TapeCode<CAP>& C, int pc) { return C.ins[pc]; }
#define RM_TAPE_LEN(n_instr) (n_instr)
#define RM_TAPE_LOOP
#endif

// CULLMODE: 0 = the code carries no culling flags; 1 = flags present (a culling launch) but this
// instance evaluates every child (IEEE stencil samples); 2 = cull.
template <int K, bool FAST, int CAP, int CULLMODE = 0, bool RECORD = false>
__device__ __forceinline__ void eval_scene(const TapeCode<CAP>& C, const int n_instr,
                                           const float4* __restrict__ spar, const float (&px)[K],
                                           const float (&py)[K], const float (&pz)[K],
                                           float (&out)[K], float* hint = nullptr,
                                           const float* move = nullptr,
                                           NodeRecords<K, CAP>* rec = nullptr) {
#ifdef RM_SPEC_INC
    if constexpr (!RECORD && CULLMODE != 2) {
        eval_spec<K, FAST>(spar, px, py, pz, out);
        return;
    }
#endif
    using M = Math<FAST>;
    constexpr bool CULL = CULLMODE == 2;
    float seed[CULL ? K : 1];                        // lower bound of the culled union's largest term
    float cs = 0.f;                                  // c * s_min of the culled union (<= 0)
    if (CULL) {
#pragma unroll
        for (int k = 0; k < (CULL ? K : 1); ++k) seed[k] = -CUDART_INF_F;
    }
    float a0[K], a1[K], d[K];
    float c = 0.f;                                   // scale of the innermost open smooth union
    float stk[MAX_DEPTH * (2 * K + 1)];              // spilled accumulators of enclosing unions
    int sp = 0;
    const char* sbytes = reinterpret_cast<const char*>(spar);
    // root union without post-ops: start inside it (see TapeCode::hdr); record mode walks everything
    int pc0 = 0, pc1 = RM_TAPE_LEN(n_instr), rootm = 0;
    if (!RECORD) {
        const int4 h = C.hdr;
        pc0 = h.x; pc1 = h.y; rootm = h.z;
        if (rootm == 2) c = M::exp_scale(*reinterpret_cast<const float*>(sbytes + h.w));
    }
    const float a_init = rootm == 1 ? CUDART_INF_F : (rootm == 2 ? -CUDART_INF_F : 0.f);
#pragma unroll
    for (int k = 0; k < K; ++k) { a0[k] = a_init; a1[k] = 0.f; d[k] = 0.f; }

    int4 nxt = tape_ins(C, pc0);
    RM_TAPE_LOOP
    for (int pc = pc0; pc < pc1; ++pc) {
        const int4 ins = nxt;
        nxt = tape_ins(C, pc + 1);                    // prefetch (pad entry after the last instruction)
        const float4* P = reinterpret_cast<const float4*>(sbytes + ins.z);
        const int disp_raw = ins_disp(ins.w);
        const int disp = CULLMODE ? (disp_raw & ~DISP_CULL) : disp_raw;  // flags are only set for culling launches
        if (CULL) {
            if ((disp_raw & DISP_CULL) && disp < DISP_OPEN_MIN) {
                const float4 cr = P[-1];             // bounding sphere of this leaf
                bool skip = true;
#pragma unroll
                for (int k = 0; k < K; ++k) {
                    const float dx = px[k] - cr.x, dy = py[k] - cr.y, dz = pz[k] - cr.z;
                    const float t = M::sqrt(fmaf(dx, dx, fmaf(dy, dy, dz * dz))) - cr.w;
                    skip = skip && t > 0.f && (cs * t - fmaxf(a0[k], seed[k]) < CULL_LOG2);
                }
                if (__all_sync(0xffffffffu, skip)) continue;
            }
        }

#define RM_LEAF_FRAME(EXPR)                                                                     \
    {                                                                                           \
        const float4 r0 = P[0], r1 = P[1], r2 = P[2];                                           \
        const float4 p0 = P[3];                                                                 \
        (void)p0;                                                                               \
        _Pragma("unroll") for (int k = 0; k < K; ++k) {                                         \
            const float x = fmaf(r0.x, px[k], fmaf(r0.y, py[k], fmaf(r0.z, pz[k], r0.w)));      \
            const float y = fmaf(r1.x, px[k], fmaf(r1.y, py[k], fmaf(r1.z, pz[k], r1.w)));      \
            const float z = fmaf(r2.x, px[k], fmaf(r2.y, py[k], fmaf(r2.z, pz[k], r2.w)));      \
            (void)y; (void)z;                                                                   \
            d[k] = EXPR;                                                                        \
        }                                                                                       \
    }                                                                                           \
    break;
#define RM_LEAF_WORLD(EXPR)                                                                     \
    {                                                                                           \
        const float4 p0 = P[0];                                                                 \
        (void)p0;                                                                               \
        _Pragma("unroll") for (int k = 0; k < K; ++k) {                                         \
            const float x = px[k], y = py[k], z = pz[k];                                        \
            (void)y; (void)z;                                                                   \
            d[k] = EXPR;                                                                        \
        }                                                                                       \
    }                                                                                           \
    break;
        switch (disp) {
            case OP_SPHERE: RM_LEAF_WORLD(prim_sphere<FAST>(p0, x, y, z))
            case OP_BOX: RM_LEAF_WORLD(prim_box<FAST>(p0, x, y, z))
            case OP_TORUS: RM_LEAF_WORLD(prim_torus<FAST>(p0, x, y, z))
            case OP_LINE: RM_LEAF_WORLD(prim_line<FAST>(p0, P[1], P[2], x, y, z))
            case OP_DISK: RM_LEAF_WORLD(prim_disk<FAST>(p0, x, y, z))
            case OP_PLANE: RM_LEAF_WORLD(x)                               // primitives.py:40-41
            case OP_SPHERE | DISP_FRAME: RM_LEAF_FRAME(prim_sphere<FAST>(p0, x, y, z))
            case OP_BOX | DISP_FRAME: RM_LEAF_FRAME(prim_box<FAST>(p0, x, y, z))
            case OP_TORUS | DISP_FRAME: RM_LEAF_FRAME(prim_torus<FAST>(p0, x, y, z))
            case OP_LINE | DISP_FRAME: RM_LEAF_FRAME(prim_line<FAST>(p0, P[4], P[5], x, y, z))
            case OP_DISK | DISP_FRAME: RM_LEAF_FRAME(prim_disk<FAST>(p0, x, y, z))
            case OP_PLANE | DISP_FRAME: RM_LEAF_FRAME(x)
            case DISP_OPEN_MIN:
            case DISP_OPEN_SMOOTH: {
                if (!(ins.x & HEAD_ROOT)) {
                    float* s = stk + sp * (2 * K + 1);
#pragma unroll
                    for (int k = 0; k < K; ++k) { s[2 * k] = a0[k]; s[2 * k + 1] = a1[k]; }
                    s[2 * K] = c;
                    ++sp;
                }
                if (disp == DISP_OPEN_MIN) {
#pragma unroll
                    for (int k = 0; k < K; ++k) { a0[k] = CUDART_INF_F; if (RECORD) a1[k] = __int_as_float(0); }
                } else {
                    const float4 u = P[0];           // (-k, s_min, L_max, log2 n)
                    c = M::exp_scale(u.x);
                    if (RECORD) rec->c[word_node(ins.y)] = u.x;      // natural -k (c is base-2 scaled under FAST)
#pragma unroll
                    for (int k = 0; k < K; ++k) { a0[k] = -CUDART_INF_F; a1[k] = 0.f; }
                    if (CULL && (disp_raw & DISP_CULL) && u.w > 0.f) {
                        cs = c * u.y;
#pragma unroll
                        for (int k = 0; k < (CULL ? K : 1); ++k)
                            seed[k] = fmaf(c, fmaf(u.z, move[k], hint[k]), -u.w);
                    }
                }
                continue;
            }
            default: {                                // CLOSE: finish this union, restore parent
                if (disp == DISP_CLOSE_MIN) {
#pragma unroll
                    for (int k = 0; k < K; ++k) {
                        d[k] = a0[k];
                        if (RECORD) rec->win[word_node(ins.y)][k] = __float_as_int(a1[k]);
                    }
                } else {
#pragma unroll
                    for (int k = 0; k < K; ++k) d[k] = M::div(a0[k] + M::log(a1[k]), c);
                    if (CULL && (disp_raw & DISP_CULL)) {
#pragma unroll
                        for (int k = 0; k < (CULL ? K : 1); ++k) hint[k] = d[k];
                        cs = 0.f;
                    }
                }
                if (RECORD) {
#pragma unroll
                    for (int k = 0; k < K; ++k) rec->d[word_node(ins.y)][k] = d[k];
                }
                if (!(ins.x & HEAD_ROOT)) {
                    --sp;
                    const float* s = stk + sp * (2 * K + 1);
#pragma unroll
                    for (int k = 0; k < K; ++k) { a0[k] = s[2 * k]; a1[k] = s[2 * k + 1]; }
                    c = s[2 * K];
                }
                break;
            }
        }
#undef RM_LEAF_FRAME
#undef RM_LEAF_WORLD

        // tail = fold * 4 + post case: unary value ops (rounding / onion), then the parent's fold
        const int tail = ins_tail(ins.w);
        const int post = tail & 3;
        if (post != TAIL_POST_NONE) {
            const float4 pp = *reinterpret_cast<const float4*>(sbytes + ins_post_off(ins.w));
            if (post == TAIL_POST_ONION) {
#pragma unroll
                for (int k = 0; k < K; ++k) d[k] = fabsf(d[k]) - pp.x;
            } else if (post == TAIL_POST_ROUND) {
#pragma unroll
                for (int k = 0; k < K; ++k) d[k] = d[k] - pp.x;
            } else {
                const float pv[3] = {pp.x, pp.y, pp.z};
                const int np = head_npost(ins.x);
#pragma unroll
                for (int i = 0; i < 3; ++i) {
                    if (i < np) {
                        const bool onion = head_post(ins.x, i) == POST_ONION;
#pragma unroll
                        for (int k = 0; k < K; ++k) d[k] = (onion ? fabsf(d[k]) : d[k]) - pv[i];
                    }
                }
            }
        }
        // the parent's fold
        if (tail & (FOLD_SMOOTH * 4)) {
#pragma unroll
            for (int k = 0; k < K; ++k) {
                if (RECORD) rec->y[pc][k] = d[k];
                const float x = c * d[k];
                const float e = M::exp(-fabsf(x - a0[k]));
                a1[k] = (x > a0[k]) ? fmaf(a1[k], e, 1.f) : a1[k] + e;
                a0[k] = fmaxf(a0[k], x);
            }
        } else if (tail & (FOLD_MIN * 4)) {
            if (RECORD) {
                const int ordinal = head_ordinal(ins.x);
#pragma unroll
                for (int k = 0; k < K; ++k)
                    if (d[k] < a0[k]) { a0[k] = d[k]; a1[k] = __int_as_float(ordinal); }   // strict: first wins ties
            } else {
#pragma unroll
                for (int k = 0; k < K; ++k) a0[k] = min_nan(a0[k], d[k]);
            }
        }                                             // else FOLD_NONE: the root (value stays in d)
    }
    if (rootm == 1) {
#pragma unroll
        for (int k = 0; k < K; ++k) out[k] = a0[k];
    } else if (rootm == 2) {
#pragma unroll
        for (int k = 0; k < K; ++k) out[k] = M::div(a0[k] + M::log(a1[k]), c);
    } else {
#pragma unroll
        for (int k = 0; k < K; ++k) out[k] = d[k];
    }
}

// Cooperative copy of the tape parameters from global into dynamic shared memory.
__device__ __forceinline__ const float4* stage_params(const float* par, int n_par,
                                                      unsigned char* smem) {
    float4* spar = reinterpret_cast<float4*>(smem);
    const float4* gpar = reinterpret_cast<const float4*>(par);
    for (int i = threadIdx.x; i < n_par / 4; i += blockDim.x) spar[i] = __ldg(gpar + i);
    __syncthreads();
    return spar;
}

}  // namespace rm
