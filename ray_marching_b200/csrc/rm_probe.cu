// FP32 roofline probe: dependent-chain-free FFMA loop on every SM.  Measurement infrastructure
// for bench.py (MEASURED_PEAKS.json has no FP32 CUDA-core figure); not on the
// render path.
#include "rm_host.h"

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

}  // namespace rm

// Launches blocks x 256 threads; every thread issues iters * 128 FFMA (= 256 flop).
extern "C" int rm_probe_ffma(int32_t blocks, int32_t iters, float* sink, void* stream) {
    if (blocks <= 0 || iters <= 0 || sink == nullptr)
        return rm::fail(RM_ERR_INVALID, "rm_probe_ffma: bad arguments");
    rm::ffma_probe_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(iters, 1.0f, sink);
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}
