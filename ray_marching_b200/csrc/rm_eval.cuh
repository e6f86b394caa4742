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
// CULL: the children of one designated smooth union carry a bounding sphere (world
// centre c_i, radius R_i with f_i(p) >= s_min (|p - c_i| - R_i) outside it).  A child is skipped when, for
// EVERY ray of the warp, its term x_i = c_k f_i (base-2 scaled, c_k < 0) lies 30 below a LOWER bound of the
// union's largest term.  That bound comes from the previous evaluation: `hint` carries the running maximum
// a0 = max_i x_i of the culled union at the previous point in, and at this point out; every term is
// Lipschitz, |x_i(p') - x_i(p)| <= |c_k| L |p' - p|, so max_i x_i(p') >= hint + c_k L m with m = `move`, the
// distance moved since (hint = -inf: no information, nothing is skipped).
// The test is WARP-PARALLEL: the 32 lanes test 32 consecutive instructions at once, each against the warp's
// bounding sphere (centre = one lane's point, radius R_w = the farthest ray of the warp from it):
// |p - c_i| >= |c_w - c_i| - R_w for every ray, so one ballot yields a skip mask for 32 children and the
// interpreter jumps over masked runs.  A skipped child costs 1/32 of a dozen instructions instead of half a
// leaf, which is what makes culling pay on 512-primitive scenes (profiles/r02/variants.md).
// Cut-off: a skipped term is below 2^-24 = 6e-8 of the largest one (SURVEY 8(f)-1 proposes 1e-7): the sum of n
// children changes by at most n 2^-24 relative (512 children: 3e-5, i.e. 4.4e-5 in the base-2 logarithm, i.e.
// 1.4e-6 distance units at k = 22 -- seventy times inside the 1e-4 relative depth tolerance at depth 4).
constexpr float CULL_LOG2 = -24.0f;
#ifdef RM_CULL_STATS
// counters for tools/cull_stats.py (a -DRM_CULL_STATS build only): [0] evals, [1] evals with a usable bound, [2] whole-union skips, [3] leaves evaluated inside the
// culled union, [4] leaves skipped by the mask (per warp)
__device__ unsigned long long g_cull_stats[8];
#define RM_STAT(i, n) do { if ((threadIdx.x & 31) == 0) atomicAdd(&g_cull_stats[i], (unsigned long long)(n)); } while (0)
#else
#define RM_STAT(i, n) do { } while (0)
#endif

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
// RECORD: 0 = value only; 1 = fill the node records and every smooth-union child's value (pass 1 of the
// scene VJP); 2 = node records only (the replay of the fused loss kernel, whose sweep re-evaluates leaves):
// REC may then be NodeRecords<K, 1> -- the per-instruction array y is never touched.
//
// Culling (one smooth union per tape, tape.py: the largest one).  A child whose term 2^(c f_i) is below
// 2^CULL_LOG2 of the union's largest term cannot change a float32 logsumexp by more than 2^CULL_LOG2 relative,
// and is skipped.  Everything is decided per WARP, so that the instruction stream stays warp-uniform:
//   * at OPEN the warp bounds its 32 K points by a sphere (centre c_w, radius r_w) and the union's largest term
//     from below by seed_min = max(carried bound, geometric bound): `hint` is the largest term at the previous
//     evaluation of each ray and decays by c L |move| (every child is L-Lipschitz); the geometric bound needs
//     no history -- some child's surface lies inside its sphere, so its value is <= L (|c_w - c_j| + R_j + r_w);
//   * the union's sphere table (one float4 per body instruction, shared memory) is tested 32 instructions at a
//     time, one per lane: child j is at least t_j = |c_w - c_j| - R_j - r_w outside its sphere, its term is
//     <= c s_min t_j, and it is masked when that is 2^CULL_LOG2 below seed_min.  The mask is OR-reduced into
//     a uniform register and the interpreter jumps over runs of masked instructions;
//   * when the union is a plain child of a hard union, the same table gives f_U >= (c s_min t_min + log2 n)/c;
//     if that exceeds the parent's running minimum for every ray of the warp, the whole union is skipped --
//     exactly (min(parent, f_U) = parent).
// render_kernel gives a culling warp a 16 x 8 pixel patch instead of 128 pixels of one row: r_w is what
// decides how many children survive.
template <int K, bool FAST, int CAP, int CULLMODE = 0, int RECORD = 0, class REC = NodeRecords<K, CAP>>
__device__ __forceinline__ void eval_scene(const TapeCode<CAP>& C, const int n_instr,
                                           const float4* __restrict__ spar, const float (&px)[K],
                                           const float (&py)[K], const float (&pz)[K],
                                           float (&out)[K], float* hint = nullptr,
                                           const float* move = nullptr,
                                           REC* rec = nullptr) {
#ifdef RM_SPEC_INC
    if constexpr (!RECORD && CULLMODE != 2) {
        eval_spec<K, FAST>(spar, px, py, pz, out);
        return;
    }
#endif
    using M = Math<FAST>;
    constexpr bool CULL = CULLMODE == 2;
    // culling state, all warp-uniform: skip mask of the 32 instructions from mask_base on
    unsigned skip_mask = 0u;
    int mask_base = -0x40000000;
    bool cull_on = false;                            // inside the culled union, with a usable bound
    float cs = 0.f, seed_min = 0.f;                  // c * s_min of the culled union (<= 0); min over the warp of the seed
    float cwx = 0.f, cwy = 0.f, cwz = 0.f, rw = 0.f; // the warp's bounding sphere
    const float4* tbl = spar;                        // sphere table of the culled union, for pcs [tbl_pc0, tbl_pc1)
    int tbl_pc0 = 0, tbl_pc1 = 0;
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
        if (CULL && cull_on) {
            int off = pc - mask_base;
            if (off >= 32) {
                // lane L tests instruction pc + L against the union's sphere table (one float4 per instruction of
                // the body, consecutive in shared memory; radius = 3e38 for an instruction without a sphere)
                const int q = pc + (int)(threadIdx.x & 31);
                bool sk = false;
                if (q < tbl_pc1) {
                    const float4 cr = tbl[q - tbl_pc0];
                    const float dx = cwx - cr.x, dy = cwy - cr.y, dz = cwz - cr.z;
                    const float t = M::sqrt(fmaf(dx, dx, fmaf(dy, dy, dz * dz))) - cr.w - rw;
                    sk = t > 0.f && (cs * t - seed_min < CULL_LOG2);                      // false for NaN
                }
                // REDUX (not VOTE): the result lands in a uniform register, so pc and every branch of the
                // interpreter stay on the uniform datapath
                skip_mask = __reduce_or_sync(0xffffffffu, sk ? (1u << (threadIdx.x & 31)) : 0u);
                mask_base = pc;
                off = 0;
            }
            const unsigned m = skip_mask >> off;
            if (m & 1u) {                                              // jump over the whole masked run
                const int run = (~m) != 0u ? __ffs((int)~m) - 1 : 32;  // consecutive masked instructions from pc
                RM_STAT(4, run);
                pc += run - 1;                                         // (the loop adds the last one)
                nxt = tape_ins(C, pc + 1);
                continue;
            }
        }
        if (CULL && pc >= tbl_pc0 && pc < tbl_pc1) RM_STAT(3, 1);
        const int4 ins = nxt;
        nxt = tape_ins(C, pc + 1);                    // prefetch (pad entry after the last instruction)
        const float4* P = reinterpret_cast<const float4*>(sbytes + ins.z);
        const int disp_raw = ins_disp(ins.w);
        const int disp = CULLMODE ? (disp_raw & ~DISP_CULL) : disp_raw;  // flags are only set for culling launches

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
                float s_parent[CULL ? K : 1], s_parent1[CULL ? K : 1];      // the parent's accumulators (culling)
                const float c_parent = c;
                if (CULL) {
#pragma unroll
                    for (int k = 0; k < (CULL ? K : 1); ++k) { s_parent[k] = a0[k]; s_parent1[k] = a1[k]; }
                }
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
                        // lower bound of this union's largest term at these points, minimum over the warp
                        float sm = CUDART_INF_F;
#pragma unroll
                        for (int k = 0; k < (CULL ? K : 1); ++k) sm = fminf(sm, fmaf(c * u.z, move[k], hint[k]));
                        // the warp's bounding sphere: centre = lane 16's first point
                        cwx = __shfl_sync(0xffffffffu, px[0], 16);
                        cwy = __shfl_sync(0xffffffffu, py[0], 16);
                        cwz = __shfl_sync(0xffffffffu, pz[0], 16);
                        float r2 = 0.f;
#pragma unroll
                        for (int k = 0; k < K; ++k) {
                            const float dx = px[k] - cwx, dy = py[k] - cwy, dz = pz[k] - cwz;
                            const float q2 = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
                            r2 = q2 > r2 ? q2 : (q2 == q2 ? r2 : CUDART_INF_F);      // NaN point: no culling
                        }
#pragma unroll
                        for (int o = 16; o > 0; o >>= 1) {
                            sm = fminf(sm, __shfl_xor_sync(0xffffffffu, sm, o));
                            r2 = fmaxf(r2, __shfl_xor_sync(0xffffffffu, r2, o));
                        }
                        rw = sqrt_rn_inline(r2) * 1.0001f;
                        RM_STAT(0, 1);
                        const int close_pc = ins_post_off(ins.w) & 0xFFFF;        // pc of this union's CLOSE
                        const bool plain_min_child = (ins.w & (1 << 16)) != 0;     // ... of a hard union, no post-op
                        tbl = spar + head_ordinal(ins.x);                          // (ordinal field of OPEN = table offset)
                        tbl_pc0 = pc + 1;
                        tbl_pc1 = close_pc;
                        // One lane-parallel pass over the sphere table: t_min = the least distance from the warp's
                        // sphere to a child's sphere (every child's value >= smin t_i), u_min = the least distance
                        // to the FAR side of a child's sphere (that child's value <= L (u_i + rw): its surface has a
                        // point inside its sphere and its value is L-Lipschitz).
                        float tmin = CUDART_INF_F, umin = CUDART_INF_F;
                        for (int base = pc + 1; base < close_pc; base += 32) {
                            const int q = base + (int)(threadIdx.x & 31);
                            if (q < close_pc) {
                                const float4 cr = tbl[q - tbl_pc0];
                                const float dx = cwx - cr.x, dy = cwy - cr.y, dz = cwz - cr.z;
                                const float dist = M::sqrt(fmaf(dx, dx, fmaf(dy, dy, dz * dz)));
                                const float t = dist - cr.w - rw;
                                tmin = t < tmin ? t : (t == t ? tmin : -CUDART_INF_F);   // no sphere: t = -3e38
                                umin = fminf(umin, dist + fabsf(cr.w));                  // no sphere: 3e38, ignored
                            }
                        }
#pragma unroll
                        for (int o = 16; o > 0; o >>= 1) {
                            tmin = fminf(tmin, __shfl_xor_sync(0xffffffffu, tmin, o));
                            umin = fminf(umin, __shfl_xor_sync(0xffffffffu, umin, o));
                        }
                        // the largest term is >= c L (u_min + rw) (c < 0) whatever the history; a bound carried from
                        // the previous evaluation is usually tighter.  Moved down 1e-5 relative for rounding.
                        const float geo0 = (c * u.z) * (umin + rw);
                        const float geo = geo0 - 1e-5f * fabsf(geo0);
                        if (!(sm >= geo)) sm = geo;                            // (sm NaN: take geo)
                        seed_min = sm;
                        // Skip the WHOLE union when it cannot win its parent's min: every child is outside its
                        // sphere by t_i >= t_min > 0, so every term <= cs t_min, logsumexp <= cs t_min + log2 n
                        // and f_U >= (cs t_min + log2 n) / c.  If that exceeds the parent's running minimum
                        // (still in a0 here) for every ray of the warp, min(parent, f_U) = parent: exact.
                        if (plain_min_child) {                          // warp-uniform (rw = inf makes lb = -inf below)
                            // lower bound of the union's value (u.w = log2 n), moved DOWN by 1e-6 relative for rounding
                            const float lb0 = M::div(fmaf(cs, tmin, u.w), c);
                            const float lb = tmin > 0.f ? lb0 - 1e-6f * fabsf(lb0) : -CUDART_INF_F;
                            bool lose = true;
#pragma unroll
                            for (int k = 0; k < K; ++k) lose = lose && lb > s_parent[k];
                            if (__reduce_and_sync(0xffffffffu, lose ? 1u : 0u) != 0u) {
                                // leave the parent's accumulators as they are and continue after the CLOSE
#pragma unroll
                                for (int k = 0; k < K; ++k) { a0[k] = s_parent[k]; a1[k] = s_parent1[k]; }
                                const float c_union = c;
                                c = c_parent;
                                // nothing was evaluated here: the bound survives, decayed by the distance moved
                                // (every term is Lipschitz), so the next evaluation can still cull
#pragma unroll
                                for (int k = 0; k < (CULL ? K : 1); ++k) hint[k] = fmaf(c_union * u.z, move[k], hint[k]);
                                if (!(ins.x & HEAD_ROOT)) --sp;
                                cull_on = false;
                                RM_STAT(2, 1);
                                pc = close_pc;
                                nxt = tape_ins(C, pc + 1);
                                continue;
                            }
                        }
                        const bool exact_only = (ins.w & (1 << 17)) != 0;       // rm_tape.h: no child may be masked
                        cull_on = !exact_only && __reduce_and_sync(0xffffffffu, (sm > -CUDART_INF_F && rw < CUDART_INF_F) ? 1u : 0u) != 0u;
                        mask_base = -0x40000000;                               // first child: compute a mask
                        if (cull_on) RM_STAT(1, 1);
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
                        for (int k = 0; k < (CULL ? K : 1); ++k) hint[k] = a0[k];    // largest term at this point
                        cull_on = false;
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
                if constexpr (RECORD == 1) rec->y[pc][k] = d[k];
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
