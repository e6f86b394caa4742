// Host-side helpers shared by the ABI translation units: thread-local error string, argument
// and CUDA status checks.  Nothing here throws across the C ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>

#include "../../include/rm_b200.h"

namespace rm {

char* error_buffer();                       // thread-local, 512 bytes (rm_host.cu)
int fail(int status, const char* fmt, ...);

inline int check_tape(const RmTape* t) {
    if (t == nullptr || t->code == nullptr || t->n_instr <= 0)
        return fail(RM_ERR_INVALID, "tape is null or empty");
    if (t->n_par < 0 || (t->n_par & 3) || (t->n_par > 0 && t->par == nullptr))
        return fail(RM_ERR_INVALID, "tape parameter block must be a multiple of 4 floats");
    if (t->max_depth > 8) return fail(RM_ERR_UNSUPPORTED, "union nesting deeper than 8");
    return RM_OK;
}
inline size_t tape_smem_bytes(const RmTape* t) {
    return (size_t)t->n_instr * 16 + (size_t)t->n_par * 4;
}

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
