// Host-side helpers shared by the ABI translation units: thread-local error string, argument
// and CUDA status checks.  Nothing here throws across the C ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>

#include <string.h>

#include "../../include/rm_b200.h"
#include "rm_tape.h"

namespace rm {

char* error_buffer();                       // thread-local, 512 bytes (rm_host.cu)
int fail(int status, const char* fmt, ...);

// Validates the host-side code block: every record an instruction touches must lie inside the
// parameter block (the kernels index shared memory with these offsets), OPEN/CLOSE must balance.
inline int check_tape(const RmTape* t) {
    if (t == nullptr || t->code_host == nullptr || t->n_instr <= 0)
        return fail(RM_ERR_INVALID, "tape is null or empty");
    if (t->n_par < 0 || (t->n_par & 3) || (t->n_par > 0 && t->par == nullptr))
        return fail(RM_ERR_INVALID, "tape parameter block must be a multiple of 4 floats");
    if (t->n_instr > CODE_CAP_LARGE)
        return fail(RM_ERR_UNSUPPORTED, "tape has %d instructions (limit %d)", t->n_instr, CODE_CAP_LARGE);
    if (t->max_depth > MAX_DEPTH) return fail(RM_ERR_UNSUPPORTED, "union nesting deeper than %d", MAX_DEPTH);
    const int n_f4 = t->n_par / 4;
    int depth = 0;
    for (int i = 0; i < t->n_instr; ++i) {
        const int head = t->code_host[2 * i], w1 = t->code_host[2 * i + 1];
        const int op = head_op(head), off = word_par(w1);
        int need = head_npost(head) ? 1 : 0;
        if (op >= OP_SPHERE && op <= OP_TORUS) need += prim_f4(op) + ((head & HEAD_FRAME) ? 3 : 0);
        else if (op == OP_OPEN_MIN || op == OP_OPEN_SMOOTH) { need = op == OP_OPEN_SMOOTH ? 1 : 0; ++depth; }
        else if (op == OP_CLOSE_MIN || op == OP_CLOSE_SMOOTH) --depth;
        else return fail(RM_ERR_INVALID, "tape instruction %d has unknown opcode %d", i, op);
        if (depth < 0 || depth > MAX_DEPTH) return fail(RM_ERR_INVALID, "tape unions are unbalanced at instruction %d", i);
        if (word_cull(w1) && op >= OP_SPHERE && op <= OP_TORUS && off < 1)
            return fail(RM_ERR_INVALID, "tape instruction %d: bounding-sphere record missing", i);
        if (need > 0 && off + need > n_f4)
            return fail(RM_ERR_INVALID, "tape instruction %d reads parameters beyond the block", i);
        if ((op == OP_OPEN_MIN || op == OP_OPEN_SMOOTH || op == OP_CLOSE_MIN || op == OP_CLOSE_SMOOTH) &&
            word_node(w1) >= (t->n_nodes > 0 ? t->n_nodes : 1))
            return fail(RM_ERR_INVALID, "tape instruction %d names union node %d of %d", i, word_node(w1), t->n_nodes);
    }
    if (depth != 0) return fail(RM_ERR_INVALID, "tape unions are unbalanced");
    return RM_OK;
}
inline size_t tape_smem_bytes(const RmTape* t) { return (size_t)t->n_par * 4; }

// Runs BODY with `CAP` (constexpr) and `code` (TapeCode<CAP>, filled from the host block) in scope.
// FNV-1a over the instruction words: identifies the scene structure a specialised library was
// generated for (ray_marching_b200/specialise.py computes the same value).
inline uint32_t tape_code_hash(const RmTape* t) {
    uint32_t h = 2166136261u;
    const unsigned char* b = reinterpret_cast<const unsigned char*>(t->code_host);
    for (size_t i = 0; i < (size_t)t->n_instr * 8; ++i) { h ^= b[i]; h *= 16777619u; }
    return h;
}

#define RM_WITH_CODE(tape, ...) RM_WITH_CODE_EX(tape, false, __VA_ARGS__)
#define RM_WITH_CODE_EX(tape, with_cull, ...)                                            \
    do {                                                                                 \
        if ((tape)->n_instr <= rm::CODE_CAP_SMALL) {                                     \
            constexpr int CAP = rm::CODE_CAP_SMALL;                                      \
            rm::TapeCode<CAP> code;                                                      \
            rm::expand_code((tape)->code_host, (tape)->n_instr, code, with_cull);   \
            __VA_ARGS__                                                                  \
        } else {                                                                         \
            constexpr int CAP = rm::CODE_CAP_LARGE;                                      \
            rm::TapeCode<CAP> code;                                                      \
            rm::expand_code((tape)->code_host, (tape)->n_instr, code, with_cull);   \
            __VA_ARGS__                                                                  \
        }                                                                                \
    } while (0)

#define RM_CUDA(call)                                                                    \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess)                                                           \
            return rm::fail(RM_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); \
    } while (0)

// Raises the dynamic shared memory limit of `kernel` when the tape needs more than 48 KB.
template <typename KernelT>
inline int prepare_smem(KernelT kernel, size_t bytes) {
    if (bytes > 200 * 1024)
        return fail(RM_ERR_UNSUPPORTED, "tape needs %zu bytes of shared memory (limit 204800)", bytes);
    if (bytes > 48 * 1024)
        RM_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return RM_OK;
}

}  // namespace rm
