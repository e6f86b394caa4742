/*
 * rm_b200.h -- C ABI of the B200-native differentiable sphere-tracing path.
 *
 * The reference (kyle-rosa/ray_marching) has no FFI layer: its boundary is the Python
 * nn.Module surface.  Each entry point below names the reference
 * interface whose *body* it replaces.  Conventions:
 *   - plain C, no C++/torch types; every pointer is a DEVICE pointer unless the name ends
 *     in _host (read synchronously during the call); the caller allocates and owns every buffer;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - returns 0 on success, a negative RmStatus otherwise; never throws; the message of the
 *     last failure on the calling thread is available from rm_last_error();
 *   - no global mutable state: safe to call concurrently on different streams/buffers;
 *   - all floating point data is IEEE float32, tensors are contiguous, channels last.
 */
#ifndef RM_B200_H
#define RM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RM_B200_ABI_VERSION 7

typedef enum RmStatus {
    RM_OK = 0,
    RM_ERR_INVALID = -1,     /* bad argument (null pointer, negative size, tape too deep ...) */
    RM_ERR_CUDA = -2,        /* a CUDA runtime call failed; see rm_last_error() */
    RM_ERR_UNSUPPORTED = -3  /* e.g. tape larger than shared memory */
} RmStatus;

/* Lowered scene (ray_marching_b200/tape.py).  Replaces the recursive nn.Module tree walk of
 * scene/primitives.py:11-102 and scene/transformations.py:33-132. */
typedef struct RmTape {
    const int32_t* code_host; /* HOST [n_instr][2] instruction words (static per scene structure):
                                 word0 = op[0:4] frame[4] fold[5:7] root[7] npost[8:10] posts[10:16]
                                 ordinal[16:32]; word1 = parameter offset in float4 units [0:16],
                                 union node id [16:31], culling flag [31] (csrc/rm_tape.h).
                                 Copied into the launch parameters.                               */
    const float*   par;       /* DEVICE [n_par] parameter records, 16-byte aligned (live values)  */
    int32_t n_instr;          /* <= 1024 */
    int32_t n_par;            /* multiple of 4 */
    int32_t max_depth;        /* union nesting depth (<= 8) */
    int32_t n_nodes;          /* number of union nodes */
} RmTape;

/* Pinhole camera intrinsics: PinholeCamera.__init__, rendering/ray_marching.py:10-55. */
typedef struct RmCamera {
    int32_t n_views;       /* num_cameras */
    int32_t width;         /* px_width  */
    int32_t height;        /* px_height */
    float   focal_length;
    float   sensor_width;
    float   sensor_height;
} RmCamera;

/* Per-view pose record, 8 floats (32-byte aligned rows): camera quaternion q[4] scalar first
 * (w,x,y,z), translation t[3], 1 pad -- the (orientation, translation) pair RenderLoop.forward
 * receives (control.py:231-236).  The kernels expand it to the rotation polynomial of
 * quaternion.py:55-72 and the forward axis of QuaternionToSO3 (quaternion.py:114-124). */
#define RM_POSE_STRIDE 8

/* Shader mode bits; bit m = mode index m of Shader.forward (rendering/shader.py:204-209). */
enum {
    RM_MODE_LAMBERTIAN = 1 << 0, RM_MODE_DISTANCE = 1 << 1, RM_MODE_PROXIMITY = 1 << 2,
    RM_MODE_VIGNETTE = 1 << 3, RM_MODE_NORMAL = 1 << 4, RM_MODE_LAPLACIAN = 1 << 5,
    RM_MODE_TANGENT = 1 << 6, RM_MODE_SPIN = 1 << 7
};

/* Global-reduction scratch written by rm_render and consumed by rm_shade_finalize (8 floats):
 * [0] -min, [1] max of the distance-shader log, [2] -min, [3] max of the proximity-shader log,
 * [4] max |laplacian|, [5] NaN seen in distance log, [6] in proximity log, [7] in laplacian.
 * Minima are stored negated so that EVERY entry combines with MAX: a tile- or view-sharded job
 * needs one 8-float MAX all-reduce between the two calls. */
#define RM_REDUCE_FLOATS 8

/* Output planes of the fused render; any pointer may be NULL (plane not produced).
 * A "ray" is one pixel of one view; planes are indexed by the LOCAL ray index
 * r - ray_begin, r = (view*height + row)*width + col. */
typedef struct RmRenderOut {
    float* origin;      /* [n][3] ray start = pixel position   (PinholeCamera.forward item 0/2) */
    float* dirs;        /* [n][3] unit ray direction           (PinholeCamera.forward item 3)   */
    float* pos;         /* [n][3] marched position             (SDFMarcher.forward)             */
    float* sdf;         /* [n]    f(pos)                       (control.py:244)                 */
    float* normals;     /* [n][3]                              (SDFNormals.forward item 0)      */
    float* laplacian;   /* [n]                                 (SDFNormals.forward item 1)      */
    float* image[8];    /* per shader mode: [n] for modes 0,1,2,3,5, [n][3] for modes 4,6,7     */
    float* rgba;        /* [n][4] headless display buffer for `display_mode` (alpha = 1); the
                           reference's torchwindow contract, torchwindow/window.py:146-174       */
    float* reduce;      /* [RM_REDUCE_FLOATS] scratch, see above; required for modes 1,2,5      */
} RmRenderOut;

typedef struct RmRenderArgs {
    int64_t ray_begin;      /* first global ray of this shard */
    int64_t n_rays;         /* rays in this shard */
    int32_t steps;          /* marching_steps (reference default 32) */
    int32_t bounces;        /* reflected-ray segments after the first (0 = reference behaviour) */
    float   normals_eps;    /* stencil size (RenderLoop default 5e-2) */
    float   hit_tau;        /* |f| threshold for a reflection hit (config 3: 1e-3) */
    float   lift_delta;     /* lift along the normal after a hit (config 3: 2*eps) */
    uint32_t mode_mask;     /* which image[] planes to shade */
    int32_t display_mode;   /* mode index shaded into rgba, or -1 */
    int32_t degree;         /* colour-wheel multiplicity of modes 6,7 */
    int32_t early_exit;     /* 1: stop marching a ray once p stops changing (bit-identical) */
    int32_t fast_math;      /* 1: approximate sqrt/exp inside march steps (stencil stays IEEE);
                               2: additionally skip smooth-union children whose bounding sphere proves
                               their softmax term < 2^-24 of the largest one (warp-parallel test);
                               3: like 1, plus only the EXACT skip of 2 -- a whole smooth union under a hard
                               union is not evaluated where its bounding spheres prove that it loses the
                               min for every ray of the warp (same pixels as 1, bit for bit)           */
    const float* cmap;      /* [4096][3] cyclic colour LUT for modes 6,7 (may be NULL otherwise) */
    float   bwd_converge_tol; /* rm_render_bwd only: see rm_march_bwd (0 = reverse every step) */
    /* Interleaved shard ("row-interleaved tiles"): with strip_rays > 0 the shard's LOCAL
     * ray l is the global ray  ray_begin + (l / strip_rays) * strip_stride + l % strip_rays,  i.e. strips
     * of strip_rays consecutive rays taken every strip_stride rays (rank g of G: ray_begin = g * strip,
     * strip_stride = G * strip).  strip_rays must be a multiple of 4; 0 = one contiguous range. */
    int64_t strip_rays;
    int64_t strip_stride;
} RmRenderArgs;

int rm_abi_version(void);
const char* rm_last_error(void);

/* scene(points) -> sdf.  Replaces the nn.Module call of every scene node
 * (scene/primitives.py:11,25,40,51,69,91; scene/transformations.py:33,67,90,117,131). */
int rm_sdf_eval(const RmTape* tape, const float* points, int64_t n_points, float* sdf_out,
                void* stream);

/* PinholeCamera.forward (rendering/ray_marching.py:57-64): world-space ray origins + directions
 * for rays [ray_begin, ray_begin + n_rays). */
int rm_camera_rays(const RmCamera* cam, const float* poses, int64_t ray_begin, int64_t n_rays,
                   float* origin_out, float* dirs_out, void* stream);

/* Camera-state integration of EventAggregator.get_state (control.py:148-165), in place on the
 * device pose records [n_views][RM_POSE_STRIDE]:
 *     t <- t + Q.rotation(translation_input * translation_sensitivity, q)        (old q)
 *     q <- normalize(Q.multiply(q, Q.to_versor(orientation_input * rotation_sensitivity)))
 * inputs: device [n_frames][n_views][6] = (translation input x,y,z, orientation input yz,zx,xy) --
 * the two vectors get_state builds from the keyboard / mouse state.  frame_counter: device int32;
 * frame (*frame_counter mod n_frames) is applied and the counter incremented, so a scripted path
 * advances inside a captured CUDA graph with no host work; NULL is allowed when n_frames == 1. */
int rm_camera_update(float* poses, const float* inputs, int32_t n_views, int32_t n_frames,
                     float translation_sensitivity, float rotation_sensitivity,
                     int32_t* frame_counter, void* stream);

/* SDFMarcher.forward (rendering/ray_marching.py:72-84): p <- f(p) v + p, `steps` times. */
int rm_march(const RmTape* tape, const float* origin, const float* dirs, int64_t n_rays,
             int32_t steps, int32_t fast_math, float* pos_out, void* stream);

/* SDFNormals.forward (rendering/ray_marching.py:115-125). */
int rm_normals(const RmTape* tape, const float* points, int64_t n_points, float normals_eps,
               float* normals_out, float* laplacian_out, void* stream);

/* RenderLoop.forward (control.py:231-258) fused: camera -> march (+reflected segments) ->
 * f(p) -> stencil -> per-pixel shaders.  Normalising shaders (modes 1,2,5) leave their
 * un-normalised value in image[m] and min/max in out->reduce; call rm_shade_finalize next. */
int rm_render(const RmTape* tape, const RmCamera* cam, const float* poses,
              const RmRenderArgs* args, const RmRenderOut* out, void* stream);

/* Resets the reduction scratch (call before the rm_render calls that share it). */
int rm_reduce_init(float* reduce, void* stream);

/* Second phase of DistanceShader / ProximityShader / LaplacianShader (shader.py:34-38,51-55,
 * 81-89): applies the global min/max normalisation and the 1/2.33 gamma in place. */
int rm_shade_finalize(const float* reduce, uint32_t mode_mask, int64_t n_rays,
                      float* const* image, float* rgba, int32_t display_mode, void* stream);

/* Half-precision copies of the shaded image planes for transport (the reference's
 * own CUDA run dtype is float16, main.py:21; all arithmetic here stays float32).  For every mode m in
 * mode_mask with image_f16[m] != NULL converts image[m] ([n] or [n][3] float32, finalised) to IEEE
 * binary16 with round-to-nearest-even; one launch for all planes.  Shaded pixels lie in [0, 1], where
 * binary16 is exact to 2.4e-4 -- inside the 1e-3 pixel tolerance. */
int rm_planes_to_f16(uint32_t mode_mask, int64_t n_rays, const float* const* image,
                     uint16_t* const* image_f16, void* stream);

/* Backward of a scene evaluation (what torch.autograd does through the nn.Module tree of
 * scene/primitives.py + scene/transformations.py): given d loss / d sdf at n points, writes
 * d loss / d points [n][3] (may be NULL) and d loss / d par [tape->n_par] (overwritten), with the
 * reference's sub-gradient conventions.  Supports tapes with at most 16 union nodes. */
int rm_sdf_vjp(const RmTape* tape, const float* points, const float* grad_sdf, int64_t n_points,
               float* dpoints, float* dpar, void* stream);

/* Backward of SDFMarcher.forward (rendering/ray_marching.py:72-84), i.e. of the unrolled
 * `steps`-step graph the reference differentiates: given d loss / d pos_out, writes d loss / d origin,
 * d loss / d dirs ([n][3] each, may be NULL) and d loss / d par [tape->n_par] (overwritten).
 * traj_scratch: caller-provided, rm_bwd_scratch_floats(n_rays, steps) floats (the replayed
 * trajectory).  The forward replay uses IEEE math, i.e. it matches rm_march(..., fast_math = 0).
 * converge_tol = 0 reverses every step of the unrolled graph.  converge_tol > 0: the steps after a
 * ray has settled (every later step length <= converge_tol * (1 + |p|_inf)) are reversed in closed
 * form (geometric series of the fixed-point recurrence = implicit-surface adjoint truncated to the
 * same step count); the result differs from the step-by-step sum by O(converge_tol). */
int rm_march_bwd(const RmTape* tape, const float* origin, const float* dirs, const float* grad_pos,
                 int64_t n_rays, int32_t steps, float converge_tol, float* traj_scratch,
                 float* dorigin, float* ddirs, float* dpar, void* stream);

/* rm_march_bwd for callers that may not allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving
 * loops, harnesses that forbid it).  Same inputs, outputs, launch geometry and results as rm_march_bwd, but the
 * trajectory scratch is a file-scope device array inside the library (RM_MARCH_BWD_STATIC_MAX_FLOATS floats =
 * 32 MiB, held by every GPU that loads the library until the process exits); the call performs no cudaMalloc /
 * cudaFree and calls no allocating routine.  Limit: rm_bwd_scratch_floats(n_rays, steps) must not exceed
 * RM_MARCH_BWD_STATIC_MAX_FLOATS (e.g. 87 296 rays at 32 steps, any ray count at <= 18 steps); beyond it the call
 * returns RM_ERR_UNSUPPORTED with a message naming need and limit.  The first call on a device fetches the
 * array's device address (cudaGetSymbolAddress, once per device): make that call -- n_rays = 0 suffices --
 * outside graph capture.  All calls on one GPU share the array, which keeps the last call's contents (nothing
 * in it needs to start at zero; dpar is cleared by the call): calls that OVERLAP on one GPU (other streams or
 * host threads) race on it -- serialise them or use rm_march_bwd.  Exported by librm_b200.so only (not by the
 * tape-specialised libraries). */
#define RM_MARCH_BWD_STATIC_MAX_FLOATS 8388608
int rm_march_bwd_static(const RmTape* tape, const float* origin, const float* dirs, const float* grad_pos,
                        int64_t n_rays, int32_t steps, float converge_tol,
                        float* dorigin, float* ddirs, float* dpar, void* stream);

/* Upstream gradients of the G-buffer planes for rm_render_bwd, indexed like RmRenderOut (local ray
 * index); any pointer may be NULL (= zero gradient). */
typedef struct RmRenderGrads {
    const float* origin;     /* [n][3] d loss / d ray start        */
    const float* dirs;       /* [n][3] d loss / d ray direction    */
    const float* pos;        /* [n][3] d loss / d marched position */
    const float* sdf;        /* [n]    d loss / d f(pos)           */
    const float* normals;    /* [n][3] d loss / d normal           */
    const float* laplacian;  /* [n]    d loss / d laplacian        */
} RmRenderGrads;

/* Floats of trajectory scratch rm_march_bwd / rm_render_bwd need for `n_rays` rays and `steps`
 * steps (bounded: the kernels index it by thread slot, not by ray). */
int64_t rm_bwd_scratch_floats(int64_t n_rays, int32_t steps);

/* Backward of the G-buffer pipeline of RenderLoop.forward (control.py:239-245: camera -> march ->
 * f(p) -> SDFNormals), i.e. what torch.autograd does through the reference's unrolled graph, in one
 * kernel.  Uses args->ray_begin, n_rays, steps, normals_eps (bounces must be 0).  Writes
 * d loss / d par [tape->n_par] and d loss / d poses [n_views][RM_POSE_STRIDE] (dq[4], dt[3], 0),
 * both overwritten.  The forward that produced the planes must have used fast_math = 0. */
int rm_render_bwd(const RmTape* tape, const RmCamera* cam, const float* poses,
                  const RmRenderArgs* args, const RmRenderGrads* grads, float* traj_scratch,
                  float* dpar, float* dposes, void* stream);

/* Targets and weights of the fused inverse-rendering loss (rm_render_loss_bwd), indexed by the LOCAL
 * ray index like RmRenderOut.  Either target may be NULL (its term is skipped). */
typedef struct RmLossArgs {
    const float* target_lambertian; /* [n] target of LambertianShader (shader.py:16-20)            */
    const float* target_depth;      /* [n] target of the raw depth |pos - origin|                  */
    float weight_lambertian;        /* weight of sum_r (lambertian_r - target_r)^2, e.g. 1/n for a mean */
    float weight_depth;             /* weight of sum_r (depth_r - target_r)^2                       */
    float* loss_out;                /* DEVICE [2], overwritten: the two weighted sums over this shard */
    int64_t scratch_floats;         /* size of the scratch passed to rm_render_loss_bwd, in floats (checked) */
} RmLossArgs;

/* Floats of scratch rm_render_loss_bwd needs for this tape, ray count and step count: the replayed
 * trajectory (as rm_bwd_scratch_floats) plus one record per replayed evaluation and union node, which the
 * reverse pass reads back instead of re-evaluating the scene. */
int64_t rm_loss_scratch_floats(const RmTape* tape, int64_t n_rays, int32_t steps);

/* Forward AND backward of the inverse-rendering loss of SURVEY.md 8(d) (config 4),
 *     loss = w_l sum_r (clamp(-v.n, 0, 1)_r - lam*_r)^2 + w_d sum_r (|p_S - o|_r - depth*_r)^2,
 * over the pipeline of RenderLoop.forward (control.py:239-245) + LambertianShader (shader.py:16-20), in
 * ONE kernel: rays from the pixel index, march replay, stencil, shading, residuals and adjoint seeds in
 * registers, reverse march, camera-pose contraction (tapes of >= 24 instructions: two kernels -- the replay at
 * four rays per thread writes per-ray planes into `scratch`, the reverse kernel reads them).  What torch.autograd does through the reference's
 * unrolled graph plus its loss algebra; no G-buffer or gradient plane touches memory.
 * Writes loss->loss_out[2], d loss / d par [tape->n_par] and d loss / d poses [n_views][RM_POSE_STRIDE]
 * (overwritten; gradients of the SUM of both terms with unit upstream).  dpar == dposes == NULL evaluates
 * the loss only.  Uses args->ray_begin, n_rays, steps, normals_eps, bwd_converge_tol (bounces must be 0);
 * a sharded job sums loss_out / dpar / dposes over shards (one all-reduce).
 * `scratch` holds loss->scratch_floats >= rm_loss_scratch_floats(tape, n_rays, steps) floats (else RM_ERR_INVALID). */
int rm_render_loss_bwd(const RmTape* tape, const RmCamera* cam, const float* poses,
                       const RmRenderArgs* args, const RmLossArgs* loss, float* scratch,
                       float* dpar, float* dposes, void* stream);

/* Measurement aid for the FP32 roofline denominator (bench.py): blocks x 256 threads, each issuing
 * iters*128 independent FFMA (256 flop).  Not part of the render path. */
int rm_probe_ffma(int32_t blocks, int32_t iters, float* sink, void* stream);

/* Measurement aid: issue-slot behaviour of the packed FP32 instructions of sm_100 (fma.rn.f32x2).
 * kind 0 = 128 scalar FFMA per thread and iteration, 1 = the same FMAs as 64 FFMA2, 2 = a march-like
 * scalar mix (64 FFMA + 32 FMNMX + 16 MUFU), 3 = that mix with the FMAs packed.  Not on the render path. */
int rm_probe_issue(int32_t kind, int32_t blocks, int32_t iters, float* sink, void* stream);

/* Test hook for the kernels' branch-free correctly rounded arithmetic (csrc/rm_math.cuh): for n
 * operand pairs writes out[6][n] = { sqrt_inline(a), sqrt.rn(a), div_inline(a,b), div.rn(a,b),
 * rcp_or_zero(b), (0 < b < inf ? 1/b : 0) }.  Not part of the render path. */
int rm_probe_math(const float* a, const float* b, int64_t n, float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RM_B200_H */
