// FP32 roofline probe: dependent-chain-free FFMA loop on every SM.  Measurement infrastructure
// for bench.py (MEASURED_PEAKS.json has no FP32 CUDA-core figure); not on the
// render path.
#include "rm_host.h"
#include "rm_math.cuh"

namespace rm {

__global__ void __launch_bounds__(256) ffma_probe_kernel(int iters, float seed, float* sink) {
    float a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = seed + (float)(threadIdx.x + i);
    const float m = 0.999f + seed * 1e-9f, c = seed * 1e-7f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r) {
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], m, c);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += a[i];
    if (s == 123.456f) sink[0] = s;           // never true; keeps the loop alive
}

// Issue-slot probe for the packed FP32 instructions of sm_100 (fma.rn.f32x2 -> FFMA2): the march loop is
// issue-bound (profiles/r01/m_render_c2_summary.txt: 85 % of issue slots, FMA pipe 53 %, XU 64 %), so the
// question is what a warp gains when the FMA-pipe half of its instruction stream is packed two rays wide.
//   kind 0: 128 scalar FFMA per iteration                      (= ffma_probe_kernel)
//   kind 1: 64 FFMA2 per iteration (same 128 FMAs)
//   kind 2: march-like scalar mix per iteration: 64 FFMA + 32 FMNMX + 16 MUFU.SQRT
//   kind 3: the same mix with the FMAs packed: 32 FFMA2 + 32 FMNMX + 16 MUFU.SQRT
template <int KIND>
__global__ void __launch_bounds__(256) issue_probe_kernel(int iters, float seed, float* sink) {
    float a[16], b[8];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = seed + (float)(threadIdx.x + i);
#pragma unroll
    for (int i = 0; i < 8; ++i) b[i] = seed * 2.f + (float)(threadIdx.x + i);
    const float m = 0.999f + seed * 1e-9f, c = seed * 1e-7f;
    const float2 m2 = make_float2(m, m), c2 = make_float2(c, c);
    for (int it = 0; it < iters; ++it) {
        if (KIND == 0) {
#pragma unroll
            for (int r = 0; r < 8; ++r)
#pragma unroll
                for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], m, c);
        } else if (KIND == 1) {
#pragma unroll
            for (int r = 0; r < 8; ++r)
#pragma unroll
                for (int i = 0; i < 16; i += 2) {
                    const float2 v = __ffma2_rn(make_float2(a[i], a[i + 1]), m2, c2);
                    a[i] = v.x; a[i + 1] = v.y;
                }
        } else {
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                if (KIND == 2) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], m, c);
                } else {
#pragma unroll
                    for (int i = 0; i < 16; i += 2) {
                        const float2 v = __ffma2_rn(make_float2(a[i], a[i + 1]), m2, c2);
                        a[i] = v.x; a[i + 1] = v.y;
                    }
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) b[i] = fmaxf(b[i], a[i]);
#pragma unroll
                for (int i = 0; i < 4; ++i) b[i] = Math<true>::sqrt(b[i] + a[8 + i]);
            }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += a[i];
#pragma unroll
    for (int i = 0; i < 8; ++i) s += b[i];
    if (s == 123.456f) sink[0] = s;           // never true; keeps the loop alive
}

// Test hook for rm_math.cuh: the branch-free correctly rounded sqrt / div / reciprocal next to the
// compiler's sqrt.rn / div.rn on the same operands.
__global__ void math_probe_kernel(const float* a, const float* b, int64_t n, float* out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = sqrt_rn_inline(a[i]);
    out[n + i] = __fsqrt_rn(a[i]);
    out[2 * n + i] = div_rn_inline(a[i], b[i]);
    out[3 * n + i] = __fdiv_rn(a[i], b[i]);
    out[4 * n + i] = rcp_or_zero(b[i]);
    out[5 * n + i] = (b[i] > 0.f && b[i] < CUDART_INF_F) ? __fdiv_rn(1.f, b[i]) : 0.f;
}

}  // namespace rm

extern "C" int rm_probe_math(const float* a, const float* b, int64_t n, float* out, void* stream) {
    if (n < 0 || (n > 0 && (!a || !b || !out))) return rm::fail(RM_ERR_INVALID, "rm_probe_math: bad arguments");
    if (n == 0) return RM_OK;
    rm::math_probe_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(a, b, n, out);
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}

// Launches blocks x 256 threads; every thread issues iters * 128 FFMA (= 256 flop).
extern "C" int rm_probe_ffma(int32_t blocks, int32_t iters, float* sink, void* stream) {
    if (blocks <= 0 || iters <= 0 || sink == nullptr)
        return rm::fail(RM_ERR_INVALID, "rm_probe_ffma: bad arguments");
    rm::ffma_probe_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(iters, 1.0f, sink);
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}

// Issue-slot probe (see issue_probe_kernel): blocks x 256 threads, `iters` iterations of the chosen mix.
extern "C" int rm_probe_issue(int32_t kind, int32_t blocks, int32_t iters, float* sink, void* stream) {
    if (blocks <= 0 || iters <= 0 || sink == nullptr || kind < 0 || kind > 3)
        return rm::fail(RM_ERR_INVALID, "rm_probe_issue: bad arguments");
    switch (kind) {
        case 0: rm::issue_probe_kernel<0><<<blocks, 256, 0, (cudaStream_t)stream>>>(iters, 1.0f, sink); break;
        case 1: rm::issue_probe_kernel<1><<<blocks, 256, 0, (cudaStream_t)stream>>>(iters, 1.0f, sink); break;
        case 2: rm::issue_probe_kernel<2><<<blocks, 256, 0, (cudaStream_t)stream>>>(iters, 1.0f, sink); break;
        default: rm::issue_probe_kernel<3><<<blocks, 256, 0, (cudaStream_t)stream>>>(iters, 1.0f, sink); break;
    }
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}
