// Reverse-mode derivative of one scene evaluation (the VJP every backward kernel is built from).
//
// For K sample points per thread and an upstream adjoint abar[k] on f(p_k) it produces
//   * abar[k] * grad_p f(p_k)                      (returned in registers), and
//   * sum_k abar[k] * d f(p_k) / d par             (added into per-block shared accumulators),
// following torch.autograd's sub-gradient conventions for the reference ops (SURVEY.md §8(a),
// "Autograd semantics" table): norm at 0 -> 0, max/min ties -> first index, clamp bounds
// inclusive, where(x>0) strict, abs at 0 -> 0, logsumexp -> softmax weights (+ d/dk).
//
// Pass 1 re-evaluates the tape forward and records, per union node, its raw value (and for a
// hard union the winning child ordinal).  Pass 2 walks the tape backwards: every instruction
// recomputes its own value from the sample point (leaves) or from the records (union closes),
// derives its weight in the parent from the parent's recorded value, and pushes the adjoint down.
// Per-leaf parameter partials of the K samples are summed across the warp with a transposing
// butterfly (31 shuffles for 32 values) and added to shared memory by one lane per value.
#pragma once
#include "rm_eval.cuh"

namespace rm {

__device__ __forceinline__ float sgn0(float x) { return x > 0.f ? 1.f : (x < 0.f ? -1.f : 0.f); }
__device__ __forceinline__ float inv_or_zero(float n) { return rcp_or_zero(n); }

// Value, gradient w.r.t. the local point (gq) and w.r.t. the primitive's parameter record (gp, in
// the record layout written by tape.py: up to 3 float4).  IEEE math throughout.
__device__ __forceinline__ float prim_grad(int op, const float4* P, float x, float y, float z,
                                           float (&gq)[3], float (&gp)[12]) {
#pragma unroll
    for (int i = 0; i < 12; ++i) gp[i] = 0.f;
    if (op == OP_SPHERE) {                                   // primitives.py:11-17
        const float n = sqrt_rn_inline(fmaf(x, x, fmaf(y, y, z * z)));
        const float inv = inv_or_zero(n);
        gq[0] = x * inv; gq[1] = y * inv; gq[2] = z * inv;
        gp[0] = -1.f;
        return n - P[0].x;
    }
    if (op == OP_BOX) {                                      // primitives.py:25-33
        const float4 h = P[0];
        const float ax = fabsf(x) - h.x, ay = fabsf(y) - h.y, az = fabsf(z) - h.z;
        const int arg = (ax >= ay && ax >= az) ? 0 : (ay >= az ? 1 : 2);   // first index on ties
        const float mx = arg == 0 ? ax : (arg == 1 ? ay : az);
        const float ox = ax > 0.f ? ax : 0.f, oy = ay > 0.f ? ay : 0.f, oz = az > 0.f ? az : 0.f;
        const float n = sqrt_rn_inline(fmaf(ox, ox, fmaf(oy, oy, oz * oz)));
        const float inv = inv_or_zero(n);
        const float in = mx < 0.f ? 1.f : 0.f;
        const float gax = ox * inv + (arg == 0 ? in : 0.f);
        const float gay = oy * inv + (arg == 1 ? in : 0.f);
        const float gaz = oz * inv + (arg == 2 ? in : 0.f);
        gq[0] = sgn0(x) * gax; gq[1] = sgn0(y) * gay; gq[2] = sgn0(z) * gaz;
        gp[0] = -gax; gp[1] = -gay; gp[2] = -gaz;
        return n + (mx < 0.f ? mx : 0.f);
    }
    if (op == OP_LINE) {                                     // primitives.py:51-61
        const float4 p0 = P[0], ab = P[1], c = P[2];
        const float apx = x - p0.x, apy = y - p0.y, apz = z - p0.z;
        const float hr = fmaf(apx, c.x, fmaf(apy, c.y, apz * c.z));
        const bool inside = hr >= 0.f && hr <= 1.f;          // clamp passes gradient on [0,1]
        const float h = clamp_nan(hr, 0.f, 1.f);
        const float wx = fmaf(h, ab.x, -apx), wy = fmaf(h, ab.y, -apy), wz = fmaf(h, ab.z, -apz);
        const float n = sqrt_rn_inline(fmaf(wx, wx, fmaf(wy, wy, wz * wz)));
        const float inv = inv_or_zero(n);
        const float gwx = wx * inv, gwy = wy * inv, gwz = wz * inv;
        const float t = inside ? (gwx * ab.x + gwy * ab.y + gwz * ab.z) : 0.f;   // d/dh
        const float gx = fmaf(t, c.x, -gwx), gy = fmaf(t, c.y, -gwy), gz = fmaf(t, c.z, -gwz);
        gq[0] = gx; gq[1] = gy; gq[2] = gz;
        gp[0] = -gx; gp[1] = -gy; gp[2] = -gz; gp[3] = -1.f;         // start, radius
        gp[4] = h * gwx; gp[5] = h * gwy; gp[6] = h * gwz;           // AB
        gp[8] = t * apx; gp[9] = t * apy; gp[10] = t * apz;          // AB / |AB|^2
        return n - p0.w;
    }
    if (op == OP_DISK) {                                     // primitives.py:69-82
        const float ryz = sqrt_rn_inline(fmaf(y, y, z * z));
        const float rim = ryz - P[0].x;
        const bool pos = rim > 0.f;
        const float m = pos ? rim : 0.f;
        const float d = sqrt_rn_inline(fmaf(x, x, m * m));
        const float inv = inv_or_zero(d);
        const float grim = pos ? m * inv : 0.f;
        const float iyz = inv_or_zero(ryz);
        gq[0] = x * inv; gq[1] = grim * y * iyz; gq[2] = grim * z * iyz;
        gp[0] = -grim;
        return d;
    }
    if (op == OP_TORUS) {                                    // primitives.py:91-102
        const float rxz = sqrt_rn_inline(fmaf(x, x, z * z));
        const float a = rxz - P[0].x;
        const float n = sqrt_rn_inline(fmaf(a, a, y * y));
        const float inv = inv_or_zero(n);
        const float u0 = a * inv, ixz = inv_or_zero(rxz);
        gq[0] = u0 * x * ixz; gq[1] = y * inv; gq[2] = u0 * z * ixz;
        gp[0] = -u0; gp[1] = -1.f;
        return n - P[0].y;
    }
    gq[0] = 1.f; gq[1] = 0.f; gq[2] = 0.f;                   // OP_PLANE, primitives.py:40-41
    return x;
}

// Value only, at the frame-local point: the interpreter's leaf bodies behind one dispatch (the on-the-fly
// sweep's liveness test; same math policy as the replay that produced the union values it is compared with).
template <bool FAST>
__device__ __forceinline__ float prim_value(int op, const float4* P, float x, float y, float z) {
    switch (op) {
        case OP_SPHERE: return prim_sphere<FAST>(P[0], x, y, z);
        case OP_BOX: return prim_box<FAST>(P[0], x, y, z);
        case OP_TORUS: return prim_torus<FAST>(P[0], x, y, z);
        case OP_LINE: return prim_line<FAST>(P[0], P[1], P[2], x, y, z);
        case OP_DISK: return prim_disk<FAST>(P[0], x, y, z);
        default: return x;                                   // OP_PLANE
    }
}

// Sum each of 32 per-lane values across the warp; lane L ends up holding the total of value L.
__device__ __forceinline__ float warp_transpose_sum(float (&v)[32], int lane) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
        const bool upper = (lane & o) != 0;
#pragma unroll
        for (int i = 0; i < o; ++i) {
            const float send = upper ? v[i] : v[i + o];
            const float keep = upper ? v[i + o] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
    }
    return v[0];
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Pass 1 of the scene VJP: forward evaluation (IEEE math; approximate with FASTM, see forward_pass in
// rm_backward.cu) that fills the node records -- the
// interpreter of rm_eval.cuh in RECORD mode.  Returns f(p_k).
template <int K, int CAP, bool FASTM = false>
__device__ __forceinline__ void eval_record(const TapeCode<CAP>& C, const int n_instr,
                                            const float4* __restrict__ spar, const float (&px)[K],
                                            const float (&py)[K], const float (&pz)[K],
                                            NodeRecords<K, CAP>& rec, float (&out)[K]) {
    eval_scene<K, FASTM, CAP, 0, true>(C, n_instr, spar, px, py, pz, out, nullptr, nullptr, &rec);
}

}  // namespace rm
