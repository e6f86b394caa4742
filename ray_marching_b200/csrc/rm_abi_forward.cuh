// extern "C" entry points of the forward path.  Included by rm_forward.cu.
#pragma once

namespace rm {

inline unsigned grid_for(int64_t n) { return (unsigned)((n + (int64_t)BLOCK * K - 1) / ((int64_t)BLOCK * K)); }


// rm_render: threads per block.  A small frame in 256-thread blocks leaves the SMs unevenly filled
// (c1: 300 blocks on 148 SMs = 2 or 3 blocks each, at most 24 of 64 warp slots); blocks of 128 or 64
// threads spread the same warps evenly.  Large grids keep 256 (measured best, profiles/r01/
// variants.md); so do tapes whose staged parameters are too big to replicate per small block.
inline unsigned render_threads(int64_t n, size_t smem_bytes) {
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
        sms = 148;
    const int64_t tiles = (n + 32 * K - 1) / (32 * K);
    unsigned threads = BLOCK;
    while (threads > 64 && smem_bytes <= 8 * 1024 && tiles / (threads / 32) < (int64_t)sms * 8) threads /= 2;
    return threads;
}

// rm_render: persistent grid -- at most (SMs x resident blocks per SM) blocks, each walking a
// contiguous range of warp tiles (render_kernel); never more blocks than tiles.
inline unsigned render_grid(int64_t n, unsigned threads = BLOCK) {
    static int sms = 0;
    if (sms == 0) {
        int dev = 0, v = 0;
        if (cudaGetDevice(&dev) == cudaSuccess &&
            cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0)
            sms = v;
        else
            sms = 148;
    }
#if RM_PERSIST
    const int64_t tiles = (n + 32 * K - 1) / (32 * K);
    const int64_t cap = (int64_t)sms * RM_MINB;
    (void)threads;
    return (unsigned)(tiles < cap ? tiles : cap);
#else
    (void)sms;
    const int64_t per_block = (int64_t)threads * K;
    return (unsigned)((n + per_block - 1) / per_block);
#endif
}

}  // namespace rm

using namespace rm;

#ifndef RM_SPEC_INC      // a tape-specialised library exports rm_render only (plus version / error)
extern "C" int rm_sdf_eval(const RmTape* tape, const float* points, int64_t n, float* out, void* stream) {
    if (int s = check_tape(tape)) return s;
    if (n < 0 || (n > 0 && (points == nullptr || out == nullptr)))
        return fail(RM_ERR_INVALID, "rm_sdf_eval: null buffer or negative count");
    if (n == 0) return RM_OK;
    const size_t sm = tape_smem_bytes(tape);
    RM_WITH_CODE(tape, {
        if (int s = prepare_smem(sdf_eval_kernel<CAP>, sm)) return s;
        sdf_eval_kernel<CAP><<<grid_for(n), BLOCK, sm, (cudaStream_t)stream>>>(
            code, tape->par, tape->n_instr, tape->n_par, points, n, out);
    });
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}

extern "C" int rm_camera_rays(const RmCamera* cam, const float* poses, int64_t ray_begin, int64_t n,
                              float* origin, float* dirs, void* stream) {
    CamParams c;
    if (int s = make_cam(cam, &c)) return s;
    const int64_t total = (int64_t)c.n_views * c.width * c.height;
    if (poses == nullptr || ray_begin < 0 || n < 0 || ray_begin + n > total)
        return fail(RM_ERR_INVALID, "rm_camera_rays: ray range outside the camera batch");
    if (n == 0) return RM_OK;
    camera_kernel<<<grid_for(n), BLOCK, 0, (cudaStream_t)stream>>>(c, poses, ray_begin, n, origin, dirs);
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}

extern "C" int rm_camera_update(float* poses, const float* inputs, int32_t n_views, int32_t n_frames,
                                float translation_sensitivity, float rotation_sensitivity,
                                int32_t* frame_counter, void* stream) {
    if (n_views < 0 || n_frames < 1 || (n_views > 0 && (poses == nullptr || inputs == nullptr)))
        return fail(RM_ERR_INVALID, "rm_camera_update: null buffer, negative view count or no frames");
    if (frame_counter == nullptr && n_frames != 1)
        return fail(RM_ERR_INVALID, "rm_camera_update: a multi-frame path needs the device frame counter");
    if (n_views == 0) return RM_OK;
    camera_update_kernel<<<1, 128, 0, (cudaStream_t)stream>>>(poses, inputs, n_views, n_frames,
                                                              translation_sensitivity, rotation_sensitivity,
                                                              frame_counter);
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}

extern "C" int rm_march(const RmTape* tape, const float* origin, const float* dirs, int64_t n,
                        int32_t steps, int32_t fast_math, float* pos_out, void* stream) {
    if (int s = check_tape(tape)) return s;
    if (n < 0 || steps < 0 || (n > 0 && (!origin || !dirs || !pos_out)))
        return fail(RM_ERR_INVALID, "rm_march: null buffer or negative count");
    if (n == 0) return RM_OK;
    const size_t sm = tape_smem_bytes(tape);
    RM_WITH_CODE_EX(tape, fast_math >= 2, {
        if (fast_math >= 2) {
            if (int s = prepare_smem(march_kernel<true, CAP, true>, sm)) return s;
            march_kernel<true, CAP, true><<<grid_for(n), BLOCK, sm, (cudaStream_t)stream>>>(
                code, tape->par, tape->n_instr, tape->n_par, origin, dirs, n, steps, pos_out);
        } else if (fast_math) {
            if (int s = prepare_smem(march_kernel<true, CAP, false>, sm)) return s;
            march_kernel<true, CAP, false><<<grid_for(n), BLOCK, sm, (cudaStream_t)stream>>>(
                code, tape->par, tape->n_instr, tape->n_par, origin, dirs, n, steps, pos_out);
        } else {
            if (int s = prepare_smem(march_kernel<false, CAP, false>, sm)) return s;
            march_kernel<false, CAP, false><<<grid_for(n), BLOCK, sm, (cudaStream_t)stream>>>(
                code, tape->par, tape->n_instr, tape->n_par, origin, dirs, n, steps, pos_out);
        }
    });
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}

extern "C" int rm_normals(const RmTape* tape, const float* points, int64_t n, float eps,
                          float* normals, float* lap, void* stream) {
    if (int s = check_tape(tape)) return s;
    if (n < 0 || !(eps > 0.f) || (n > 0 && points == nullptr))
        return fail(RM_ERR_INVALID, "rm_normals: null buffer, negative count or eps <= 0");
    if (n == 0) return RM_OK;
    Stencil st;
    make_stencil(eps, &st);
    const size_t sm = tape_smem_bytes(tape);
    RM_WITH_CODE(tape, {
        if (int s = prepare_smem(normals_kernel<CAP>, sm)) return s;
        normals_kernel<CAP><<<grid_for(n), BLOCK, sm, (cudaStream_t)stream>>>(
            code, tape->par, tape->n_instr, tape->n_par, st, points, n, normals, lap);
    });
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}

extern "C" int rm_reduce_init(float* reduce, void* stream) {
    if (reduce == nullptr) return fail(RM_ERR_INVALID, "rm_reduce_init: null scratch");
    reduce_init_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(reduce);
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}

#endif  // RM_SPEC_INC

extern "C" int rm_render(const RmTape* tape, const RmCamera* cam, const float* poses,
                         const RmRenderArgs* a, const RmRenderOut* out, void* stream) {
    if (int s = check_tape(tape)) return s;
    RenderParams P;
    if (int s = make_cam(cam, &P.cam)) return s;
    if (a == nullptr || out == nullptr || poses == nullptr)
        return fail(RM_ERR_INVALID, "rm_render: null argument block");
    const int64_t total = (int64_t)P.cam.n_views * P.cam.width * P.cam.height;
    if (const char* why = check_shard(a->ray_begin, a->n_rays, a->strip_rays, a->strip_stride, total))
        return fail(RM_ERR_INVALID, "rm_render: %s", why);
    if (a->steps < 0 || a->bounces < 0 || !(a->normals_eps > 0.f))
        return fail(RM_ERR_INVALID, "rm_render: negative steps/bounces or eps <= 0");
    if (a->display_mode > 7) return fail(RM_ERR_INVALID, "rm_render: display_mode must be -1..7");
    const unsigned mask = a->mode_mask | (a->display_mode >= 0 ? (1u << a->display_mode) : 0u);
    for (int m = 0; m < 8; ++m)
        if ((a->mode_mask >> m & 1u) && out->image[m] == nullptr)
            return fail(RM_ERR_INVALID, "rm_render: mode %d requested but image[%d] is null", m, m);
    if ((mask & (RM_MODE_TANGENT | RM_MODE_SPIN)) && a->cmap == nullptr)
        return fail(RM_ERR_INVALID, "rm_render: modes 6/7 need the colour LUT");
    if ((mask & (RM_MODE_DISTANCE | RM_MODE_PROXIMITY | RM_MODE_LAPLACIAN)) && out->reduce == nullptr)
        return fail(RM_ERR_INVALID, "rm_render: modes 1/2/5 need the reduction scratch");
    if (a->display_mode >= 0 && out->rgba == nullptr)
        return fail(RM_ERR_INVALID, "rm_render: display_mode set but rgba is null");
    if (a->n_rays == 0) return RM_OK;
    make_stencil(a->normals_eps, &P.st);
    P.ray_begin = a->ray_begin; P.n_rays = a->n_rays; P.steps = a->steps; P.bounces = a->bounces;
    P.strip_rays = a->strip_rays; P.strip_stride = a->strip_stride;
    P.hit_tau = a->hit_tau; P.lift_delta = a->lift_delta; P.mode_mask = a->mode_mask;
    P.display_mode = a->display_mode; P.degree = a->degree; P.early_exit = a->early_exit;
    P.cmap = a->cmap; P.out = *out;
    const size_t sm = tape_smem_bytes(tape);
#if RM_PERSIST
    const unsigned threads = BLOCK;
#else
    const unsigned threads = render_threads(a->n_rays, sm);
#endif
    unsigned grid = render_grid(a->n_rays, threads);
    {
        const int64_t tiles = (a->n_rays + 32 * K - 1) / (32 * K);
        if (tiles > 0x7fffffff) return fail(RM_ERR_INVALID, "rm_render: ray range too large for one launch");
        P.tile_base = (int)(tiles / grid); P.tile_rem = (int)(tiles % grid);     // persistent tiling only
    }
    // Culled kernel: 2-D warp tiles (render_kernel) when the range is whole rows of a width that 64-pixel patch
    // columns divide; any other shape keeps the linear tiles.
    P.patch_cols = 0; P.patch_rows = 0;
#if !RM_PERSIST && !defined(RM_SPEC_INC)
    {
        const int64_t Wd = P.cam.width;
        if (a->fast_math >= 2 && K == 4 && Wd % 64 == 0 && a->ray_begin % Wd == 0 && a->n_rays % Wd == 0 &&
            (a->strip_rays == 0 || a->strip_rays % Wd == 0)) {
            const int64_t rows = a->n_rays / Wd;
            const int64_t tiles = (Wd / 64) * ((rows + 15) / 16) * 8;
            if (tiles * 32 * K <= 0x7fffffff && rows <= 0x7fffffff) {
                P.patch_cols = (int)(Wd / 64);
                P.patch_rows = (int)rows;
                grid = (unsigned)((tiles + threads / 32 - 1) / (threads / 32));
            }
        }
    }
#endif
    // Approximate-math stencil samples (stencil_eval<.., SFAST>): only when no output of this frame needs correctly
    // rounded samples -- no Laplacian, no reported f(p) (sdf plane, proximity shader), no G-buffer plane.
    const bool any_gbuffer = out->origin || out->dirs || out->pos || out->sdf || out->normals || out->laplacian;
    const bool sfast = a->fast_math >= 1 && !any_gbuffer &&
                       !(mask & (RM_MODE_LAPLACIAN | RM_MODE_PROXIMITY)) && RM_SFAST;
#ifdef RM_SPEC_INC
    // specialised library: one code capacity, no culling variant; refuses any other tape
    if (tape->n_instr > CODE_CAP_SMALL || tape_code_hash(tape) != RM_SPEC_HASH)
        return fail(RM_ERR_INVALID, "rm_render: this library is specialised for another tape (hash %08x)",
                    (unsigned)RM_SPEC_HASH);
    {
        constexpr int CAP = CODE_CAP_SMALL;
        TapeCode<CAP> code;
        expand_code(tape->code_host, tape->n_instr, code, false);
        if (sfast) {
            if (int s = prepare_smem(render_kernel<true, CAP, false, true>, sm)) return s;
            render_kernel<true, CAP, false, true><<<grid, threads, sm, (cudaStream_t)stream>>>(
                code, tape->par, tape->n_instr, tape->n_par, poses, P);
        } else if (a->fast_math) {
            if (int s = prepare_smem(render_kernel<true, CAP, false>, sm)) return s;
            render_kernel<true, CAP, false><<<grid, threads, sm, (cudaStream_t)stream>>>(
                code, tape->par, tape->n_instr, tape->n_par, poses, P);
        } else {
            if (int s = prepare_smem(render_kernel<false, CAP, false>, sm)) return s;
            render_kernel<false, CAP, false><<<grid, threads, sm, (cudaStream_t)stream>>>(
                code, tape->par, tape->n_instr, tape->n_par, poses, P);
        }
    }
#else
    RM_WITH_CODE_EX(tape, (a->fast_math >= 3 ? 2 : (a->fast_math >= 2 ? 1 : 0)), {
        if (sfast && a->fast_math >= 2) {
            if (int s = prepare_smem(render_kernel<true, CAP, true, true>, sm)) return s;
            render_kernel<true, CAP, true, true><<<grid, threads, sm, (cudaStream_t)stream>>>(
                code, tape->par, tape->n_instr, tape->n_par, poses, P);
        } else if (sfast) {
            if (int s = prepare_smem(render_kernel<true, CAP, false, true>, sm)) return s;
            render_kernel<true, CAP, false, true><<<grid, threads, sm, (cudaStream_t)stream>>>(
                code, tape->par, tape->n_instr, tape->n_par, poses, P);
        } else if (a->fast_math >= 2) {
            if (int s = prepare_smem(render_kernel<true, CAP, true>, sm)) return s;
            render_kernel<true, CAP, true><<<grid, threads, sm, (cudaStream_t)stream>>>(
                code, tape->par, tape->n_instr, tape->n_par, poses, P);
        } else if (a->fast_math) {
            if (int s = prepare_smem(render_kernel<true, CAP, false>, sm)) return s;
            render_kernel<true, CAP, false><<<grid, threads, sm, (cudaStream_t)stream>>>(
                code, tape->par, tape->n_instr, tape->n_par, poses, P);
        } else {
            if (int s = prepare_smem(render_kernel<false, CAP, false>, sm)) return s;
            render_kernel<false, CAP, false><<<grid, threads, sm, (cudaStream_t)stream>>>(
                code, tape->par, tape->n_instr, tape->n_par, poses, P);
        }
    });
#endif  // RM_SPEC_INC
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}

#ifndef RM_SPEC_INC
extern "C" int rm_shade_finalize(const float* reduce, uint32_t mode_mask, int64_t n,
                                 float* const* image, float* rgba, int32_t display_mode, void* stream) {
    if (reduce == nullptr || n < 0) return fail(RM_ERR_INVALID, "rm_shade_finalize: null scratch");
    FinalizeParams F;
    for (int m = 0; m < 8; ++m) F.image[m] = image ? image[m] : nullptr;
    F.rgba = rgba; F.mode_mask = mode_mask; F.display_mode = display_mode; F.n = n;
    const bool any = (mode_mask & (RM_MODE_DISTANCE | RM_MODE_PROXIMITY | RM_MODE_LAPLACIAN)) ||
                     (rgba && (display_mode == 1 || display_mode == 2 || display_mode == 5));
    if (!any || n == 0) return RM_OK;
    int64_t blocks = ((n >> 2) + 1 + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    shade_finalize_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(reduce, F);
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}

extern "C" int rm_planes_to_f16(uint32_t mode_mask, int64_t n, const float* const* image,
                                uint16_t* const* image_f16, void* stream) {
    if (n < 0 || image == nullptr || image_f16 == nullptr)
        return fail(RM_ERR_INVALID, "rm_planes_to_f16: null plane table or negative count");
    HalfParams H;
    int64_t longest = 0;
    for (int m = 0; m < 8; ++m) {
        const bool on = (mode_mask >> m & 1u) && image_f16[m] != nullptr;
        if (on && image[m] == nullptr)
            return fail(RM_ERR_INVALID, "rm_planes_to_f16: mode %d requested but image[%d] is null", m, m);
        const int channels = (m == 4 || m == 6 || m == 7) ? 3 : 1;
        H.src[m] = image[m];
        H.dst[m] = reinterpret_cast<__half*>(image_f16[m]);
        H.count[m] = on ? n * channels : 0;
        if (H.count[m] > longest) longest = H.count[m];
    }
    if (longest == 0) return RM_OK;
    int64_t blocks = ((longest >> 2) + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    planes_to_half_kernel<<<dim3((unsigned)blocks, 8), 256, 0, (cudaStream_t)stream>>>(H);
    RM_CUDA(cudaGetLastError());
    return RM_OK;
}
#endif  // RM_SPEC_INC
