// Tape instruction encoding shared by the kernels; mirrors ray_marching_b200/tape.py.
//
// An instruction is two 32-bit words:
//   word0: op[0:4] frame[4] fold[5:7] root[7] npost[8:10] post kinds[10:16] ordinal[16:32]
//   word1: parameter offset in float4 units [0:16], union node id [16:31], culling flag [31]
// The code block travels in the kernel-parameter space (constant bank): with a warp-uniform
// program counter every decode step (load, field extraction, compare, branch) is issued on the
// uniform datapath (LDCU / USHF / UISETP / BRA.U) and costs no vector registers, no BSSY/BSYNC
// reconvergence pairs and no shared-memory bandwidth.  Parameters (which change under an
// optimiser and live on the device) are staged into shared memory once per block.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace rm {

enum : int {
    OP_SPHERE = 1, OP_BOX = 2, OP_PLANE = 3, OP_LINE = 4, OP_DISK = 5, OP_TORUS = 6,
    OP_OPEN_MIN = 8, OP_OPEN_SMOOTH = 9, OP_CLOSE_MIN = 10, OP_CLOSE_SMOOTH = 11,
};
enum : int { FOLD_NONE = 0, FOLD_MIN = 1, FOLD_SMOOTH = 2 };
enum : int { POST_ROUND = 1, POST_ONION = 2 };

constexpr int MAX_DEPTH = 8;
constexpr int HEAD_FRAME = 1 << 4;   // leaf: 3 float4 rows of the world->local frame precede the primitive record
constexpr int HEAD_ROOT = 1 << 7;    // OPEN/CLOSE of the outermost union: no save/restore of the parent accumulator

__host__ __device__ __forceinline__ int head_op(int h) { return h & 15; }
__host__ __device__ __forceinline__ int head_fold(int h) { return (h >> 5) & 3; }
__host__ __device__ __forceinline__ int head_npost(int h) { return (h >> 8) & 3; }
__host__ __device__ __forceinline__ int head_post(int h, int i) { return (h >> (10 + 2 * i)) & 3; }
__host__ __device__ __forceinline__ int head_ordinal(int h) { return (int)((unsigned)h >> 16); }
__host__ __device__ __forceinline__ int word_par(int w) { return w & 0xFFFF; }
__host__ __device__ __forceinline__ int word_node(int w) { return (w >> 16) & 0x7FFF; }
// word1 bit 31: leaf -> a bounding-sphere record (world centre, radius / s_min) sits right before its
// first record; OPEN/CLOSE_SMOOTH -> this is the union whose children carry those records
__host__ __device__ __forceinline__ bool word_cull(int w) { return w < 0; }

// float4 records occupied by the primitive parameters of a leaf op
__host__ __device__ __forceinline__ int prim_f4(int op) {
    return op == OP_LINE ? 3 : (op == OP_PLANE ? 0 : 1);
}

// Code block passed BY VALUE as a __grid_constant__ kernel parameter.  Two capacities keep the
// launch-parameter copy small for the registry scenes (1.5 KB) while covering 512-primitive tapes.
// Each instruction is expanded on the host (expand_code) to 16 bytes:
//   x, y : word0, word1 as above (the backward kernels decode these)
//   z    : byte offset of the instruction's first parameter record in the staged block
//   w    : post-op record byte offset [0:20] | tail selector [20:24] | dispatch id [24:32]
// so that the forward interpreter needs one add for the record address, one switch on the
// dispatch id (leaf kind x has-frame, or union op) and one switch on the tail selector
// (fold kind x {no post-op, one rounding, one onion, generic chain}).
constexpr int CODE_CAP_SMALL = 96;
constexpr int CODE_CAP_LARGE = 1024;
template <int CAP>
struct TapeCode {
    // x, y: instruction range the forward interpreter walks; z: how it treats the root (0: walk the
    // whole tape; 1 / 2: the outermost hard / smooth union has no post-ops, so its OPEN and CLOSE are
    // not dispatched -- the accumulators start as that union's and the result is read from them);
    // w: byte offset of the root smooth union's record.  A registry scene is 4 leaves + this pair.
    int4 hdr;
    int4 ins[CAP + 1];     // + one pad entry: the interpreter prefetches ins[pc + 1] unconditionally
};

enum : int { DISP_CULL = 32 };   // dispatch flag: bounding-sphere culling applies (see word_cull)
enum : int { DISP_FRAME = 8, DISP_OPEN_MIN = 16, DISP_OPEN_SMOOTH = 17, DISP_CLOSE_MIN = 18, DISP_CLOSE_SMOOTH = 19 };
enum : int { TAIL_POST_NONE = 0, TAIL_POST_ROUND = 1, TAIL_POST_ONION = 2, TAIL_POST_CHAIN = 3 };  // tail = fold * 4 + post case

__host__ __device__ __forceinline__ int ins_disp(int w) { return (int)((unsigned)w >> 24); }
__host__ __device__ __forceinline__ int ins_tail(int w) { return (w >> 20) & 15; }
__host__ __device__ __forceinline__ int ins_post_off(int w) { return w & 0xFFFFF; }

// One instruction: (word0, word1) -> expanded 16 bytes.
// `with_cull`: set the DISP_CULL dispatch flag (only the culling kernels look at it).
__host__ __device__ __forceinline__ int4 expand_one(int head, int w1, bool with_cull) {
    const int op = head_op(head), np = head_npost(head), fold = head_fold(head);
    const int off = word_par(w1);
    int disp, post_f4 = off;
    if (op < OP_OPEN_MIN) {
        disp = op | ((head & HEAD_FRAME) ? DISP_FRAME : 0);
        post_f4 = off + ((head & HEAD_FRAME) ? 3 : 0) + prim_f4(op);
    } else {
        disp = DISP_OPEN_MIN + (op - OP_OPEN_MIN);
    }
    int post_case = TAIL_POST_NONE;
    if (np == 1) post_case = head_post(head, 0) == POST_ONION ? TAIL_POST_ONION : TAIL_POST_ROUND;
    else if (np > 1) post_case = TAIL_POST_CHAIN;
    const int tail = (op == OP_OPEN_MIN || op == OP_OPEN_SMOOTH) ? 0 : fold * 4 + post_case;
    if (with_cull && word_cull(w1) && op != OP_OPEN_MIN && op != OP_CLOSE_MIN) disp |= DISP_CULL;
    return make_int4(head, w1, off * 16, ((np ? post_f4 * 16 : 0) & 0xFFFFF) | (tail << 20) | (disp << 24));
}

// Host: [n][2] words -> expanded instructions.
// Culling launches: the OPEN of the culled smooth union additionally carries, in the (otherwise unused)
// post-op offset field of its w word, the pc of its CLOSE [0:16] and a flag [16] saying that the union is a
// plain child of a HARD union (its CLOSE folds with min, no post-op, not the root): then the whole union may be
// skipped, exactly, whenever a lower bound of its value exceeds the parent's running minimum (rm_eval.cuh);
// flag [17] (cull_mode 2) says that this exact skip is ALL the launch may do: no child is masked.
// (The ordinal field of that OPEN holds the float4 offset of the union's sphere table: tape.py.)
// In every launch the z word of that union's CLOSE (no parameter record of its own) holds the pc of its OPEN.
inline void expand_instructions(const int32_t* words, int n, int4* out, int cull_mode) {
    const bool with_cull = cull_mode != 0;
    for (int i = 0; i < n; ++i) out[i] = expand_one(words[2 * i], words[2 * i + 1], with_cull);
    out[n] = make_int4(0, 0, 0, 0);
    int open_at[MAX_DEPTH + 1], sp = 0;
    for (int i = 0; i < n; ++i) {
        const int op = head_op(out[i].x);
        if (op == OP_OPEN_MIN || op == OP_OPEN_SMOOTH) { if (sp <= MAX_DEPTH) open_at[sp] = i; ++sp; }
        else if (op == OP_CLOSE_MIN || op == OP_CLOSE_SMOOTH) {
            --sp;
            if (sp < 0 || sp > MAX_DEPTH) continue;
            const int o = open_at[sp];
            if (op != OP_CLOSE_SMOOTH || !word_cull(out[i].y)) continue;
            const bool plain_min_child = !(out[i].x & HEAD_ROOT) && head_fold(out[i].x) == FOLD_MIN &&
                                         head_npost(out[i].x) == 0;
            if (with_cull) out[o].w |= (i & 0xFFFF) | (plain_min_child ? (1 << 16) : 0) | (cull_mode == 2 ? (1 << 17) : 0);
            out[i].z = o;                    // every launch: the CLOSE of the culled union knows its OPEN (reverse sweep)
        }
    }
}
template <int CAP>
inline void expand_code(const int32_t* words, int n, TapeCode<CAP>& code, int cull_mode = 0) {
    const bool with_cull = cull_mode != 0;
    expand_instructions(words, n, code.ins, cull_mode);
    code.hdr = make_int4(0, n, 0, 0);
    if (n >= 2) {
        const int open = words[0], close = words[2 * (n - 1)];
        const int oop = head_op(open), cop = head_op(close);
        const bool pair = (open & HEAD_ROOT) && (close & HEAD_ROOT) && head_npost(close) == 0 &&
                          ((oop == OP_OPEN_MIN && cop == OP_CLOSE_MIN) || (oop == OP_OPEN_SMOOTH && cop == OP_CLOSE_SMOOTH));
        const bool culled_root = with_cull && word_cull(words[1]);       // the culling set-up lives in OPEN
        if (pair && !culled_root)
            code.hdr = make_int4(1, n - 1, oop == OP_OPEN_MIN ? 1 : 2, word_par(words[1]) * 16);
    }
}

}  // namespace rm
