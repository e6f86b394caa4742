// Tape instruction encoding shared by the kernels; mirrors ray_marching_b200/tape.py.
//
// An instruction is two 32-bit words:
//   word0: op[0:4] frame[4] fold[5:7] root[7] npost[8:10] post kinds[10:16] ordinal[16:32]
//   word1: parameter offset in float4 units [0:16], union node id [16:32]
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
__host__ __device__ __forceinline__ int word_node(int w) { return (int)((unsigned)w >> 16); }

// float4 records occupied by the primitive parameters of a leaf op
__host__ __device__ __forceinline__ int prim_f4(int op) {
    return op == OP_LINE ? 3 : (op == OP_PLANE ? 0 : 1);
}

// Code block passed BY VALUE as a __grid_constant__ kernel parameter.  Two capacities keep the
// launch-parameter copy small for the registry scenes (768 B) while covering 512-primitive tapes.
constexpr int CODE_CAP_SMALL = 96;
constexpr int CODE_CAP_LARGE = 1024;
template <int CAP>
struct TapeCode {
    int2 ins[CAP];
};

}  // namespace rm
