// Tape instruction encoding shared by the kernels; mirrors ray_marching_b200/tape.py.
#pragma once
#include <stdint.h>

namespace rm {

enum : int {
    OP_SPHERE = 1, OP_BOX = 2, OP_PLANE = 3, OP_LINE = 4, OP_DISK = 5, OP_TORUS = 6,
    OP_OPEN_MIN = 8, OP_OPEN_SMOOTH = 9, OP_CLOSE_MIN = 10, OP_CLOSE_SMOOTH = 11,
};
enum : int { FRAME_KEEP = 0, FRAME_WORLD = 1, FRAME_NEW = 2 };
enum : int { FOLD_NONE = 0, FOLD_MIN = 1, FOLD_SMOOTH = 2 };
enum : int { POST_ROUND = 1, POST_ONION = 2 };

constexpr int MAX_DEPTH = 8;
constexpr int HEAD_ROOT = 1 << 16;   // OPEN/CLOSE of the outermost union: no save/restore of the parent accumulator

__host__ __device__ __forceinline__ int head_op(int h) { return h & 15; }
__host__ __device__ __forceinline__ int head_frame(int h) { return (h >> 4) & 3; }
__host__ __device__ __forceinline__ int head_fold(int h) { return (h >> 6) & 3; }
__host__ __device__ __forceinline__ int head_npost(int h) { return (h >> 8) & 3; }
__host__ __device__ __forceinline__ int head_post(int h, int i) { return (h >> (10 + 2 * i)) & 3; }

// float4 records occupied by the primitive parameters of a leaf op
__host__ __device__ __forceinline__ int prim_f4(int op) {
    return op == OP_LINE ? 3 : (op == OP_PLANE ? 0 : 1);
}

// Tape as seen by a kernel: pointers into shared memory after staging.
struct TapeView {
    const int4* code;
    const float4* par;
    int n_instr;
};

}  // namespace rm
