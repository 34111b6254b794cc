/*
 * vegs_b200.h — C ABI of libvegs_b200.so: the B200 (sm_100a) Gaussian-splatting hot path of VEGS.
 *
 * Plain C: pointers and sizes only, no torch / C++ types.  Every pointer is a DEVICE pointer owned
 * by the caller unless its name ends in `_host` (then it is a host pointer; pinned memory if the
 * copy is to be asynchronous).  Nullable pointers mean "not provided", exactly as the reference
 * binding does for shs / colors_precomp / scales / rotations / cov3D_precomp
 * (reference: submodules/diff_gaussian_rasterization/rasterize_points.cu:36-134 and
 * diff_gaussian_rasterization/__init__.py:221-231).
 *
 * Every function returns 0 on success and a non-zero VEGS_E* code on failure; nothing throws or
 * aborts across the ABI.  vegs_last_error() returns a thread-local message for the last failure.
 * The library never allocates device memory: all state and scratch buffers are passed in, sized by
 * the *_bytes() queries below (one exception, stated where it is declared: vegs_dist2_noalloc
 * uses an array built into the library image).  `stream` is a cudaStream_t passed as void*; the
 * CUDA device it belongs to must be the calling thread's current device.
 *
 * All paths cited below are relative to the reference tree (deepshwang/vegs).
 */
#ifndef VEGS_B200_H_INCLUDED
#define VEGS_B200_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VEGS_ABI_VERSION 4

enum {
    VEGS_OK = 0,
    VEGS_EINVAL = 1,   /* bad argument (null pointer, bad size, inconsistent options) */
    VEGS_ECUDA = 2,    /* a CUDA call or kernel failed; see vegs_last_error() */
    VEGS_ESIZE = 3     /* a caller-provided buffer is too small */
};

/* Tile size of the rasterizer: 16x16 pixels (reference cuda_rasterizer/config.h:15-17). */
#define VEGS_TILE 16

/*
 * Forward arguments.  Replaces the 19 positional arguments of `_C.rasterize_gaussians`
 * (rasterize_points.h:17-38, rasterize_points.cu:36-134) and the three growable byte tensors
 * geomBuffer / binningBuffer / imgBuffer (rasterize_points.cu:80-85).
 */
typedef struct VegsForwardArgs {
    /* sizes */
    int32_t P;            /* number of Gaussians */
    int32_t D;            /* active SH degree (0..3) */
    int32_t M;            /* SH coefficients stored per Gaussian (sh.size(1)); 0 if shs == NULL */
    int32_t W, H;         /* image size in pixels, each <= 65535 */
    /* inputs */
    const float* background;     /* [3] */
    const float* means3D;        /* [P,3] */
    const float* shs;            /* [P,M,3] or NULL */
    const float* colors_precomp; /* [P,3] or NULL */
    const float* opacities;      /* [P] */
    const float* scales;         /* [P,3] or NULL */
    const float* rotations;      /* [P,4] (r,x,y,z) or NULL */
    const float* cov3D_precomp;  /* [P,6] or NULL */
    const float* viewmatrix;     /* [16], transposed storage (scene/cameras.py:76) */
    const float* projmatrix;     /* [16], transposed storage, full view*proj */
    const float* campos;         /* [3] */
    float scale_modifier;
    float tan_fovx, tan_fovy;
    int32_t prefiltered;         /* accepted for interface parity and ignored: the reference traps the
                                  * kernel (printf + __trap, auxiliary.h:154-160) when it is set and a
                                  * point fails the near cull; here such points are simply culled */
    int32_t debug;               /* non-zero: synchronise and check after every stage */
    /* outputs (rasterize_points.cu:69-76); every element is written by the kernels */
    float* out_color;     /* [3,H,W] */
    float* out_depth;     /* [1,H,W] */
    float* out_cov_quat;  /* [4,H,W] planes (w,x,y,z) */
    float* out_cov_scale; /* [3,H,W] */
    float* out_alpha;     /* [1,H,W] */
    int32_t* radii;       /* [P] */
    /* state kept for the backward pass (caller keeps these alive, as ctx.save_for_backward does) */
    void* geom;           /* vegs_geom_bytes(P) */
    void* img;            /* vegs_img_bytes(W,H) */
    void* binning;        /* vegs_binning_bytes(R); only needed by vegs_forward_render */
    /* scratch, free to release when the forward returns */
    void* scratch;        /* vegs_scratch_bytes(P) */
    void* bin_scratch;    /* vegs_bin_scratch_bytes(R,W,H); only needed by vegs_forward_render */
    /* TWO host words (pinned memory), written asynchronously on `stream` by vegs_forward_prepare:
     * [0] = num_rendered, [1] = number of emission work items (informational).  The library never
     * reads them back: num_rendered travels through the argument of vegs_forward_render. */
    int32_t* num_rendered_host;
    /* Fused activations and fused concatenation (no reference counterpart; the reference's caller
     * applies them in separate passes over all Gaussians, scene/gaussian_model.py:100-120).
     * activations: VEGS_ACT_* bits - the matching input array holds the RAW parameter and the
     * kernel applies sigmoid (opacities), exp (scales), x / max(|x|, 1e-12) (rotations) itself.
     * shs_rest: when not NULL, `shs` is f_dc [P,1,3] and shs_rest is f_rest [P,M-1,3]; the kernel
     * reads them as the concatenated [P,M,3] row.  Both default to 0 / NULL = the reference's inputs. */
    uint32_t activations;
    const float* shs_rest;
} VegsForwardArgs;

#define VEGS_ACT_OPACITY_SIGMOID 1u
#define VEGS_ACT_SCALE_EXP       2u
#define VEGS_ACT_ROT_NORMALIZE   4u

/*
 * Backward arguments.  Replaces the 28 positional arguments of `_C.rasterize_gaussians_backward`
 * (rasterize_points.h:40-71, rasterize_points.cu:136-240).  dL_dout_cov_quat / dL_dout_cov_scale
 * are accepted and ignored: in the reference their contribution is overwritten before it is
 * returned (cuda_rasterizer/backward.cu:624-647 vs :322-325,:340-341).
 */
typedef struct VegsBackwardArgs {
    int32_t P, D, M, W, H;
    int32_t R;                    /* num_rendered of the matching forward */
    const float* background;
    const float* means3D;
    const float* shs;
    const float* colors_precomp;
    const float* scales;
    const float* rotations;
    const float* cov3D_precomp;
    const float* viewmatrix;
    const float* projmatrix;
    const float* campos;
    float scale_modifier;
    float tan_fovx, tan_fovy;
    int32_t debug;
    const int32_t* radii;         /* [P] from the forward */
    const float* out_alpha;       /* [1,H,W] from the forward */
    /* cotangents of the five image outputs; NULL means zero */
    const float* dL_dout_color;     /* [3,H,W] */
    const float* dL_dout_depth;     /* [1,H,W] */
    const float* dL_dout_cov_quat;  /* [4,H,W]  (ignored, see above) */
    const float* dL_dout_cov_scale; /* [3,H,W]  (ignored, see above) */
    const float* dL_dout_alpha;     /* [1,H,W] */
    /* state from the forward (img is not const: the backward leaves per-tile diagnostics in it) */
    const void* geom;
    void* img;
    const void* binning;
    /* scratch: packed per-Gaussian blend-stage accumulators, vegs_bwd_scratch_bytes(P) */
    void* bwd_scratch;
    /* outputs; every element is written (zeros for Gaussians with radii == 0); any may be NULL */
    float* dL_dmeans3D;       /* [P,3] */
    float* dL_dmeans2D;       /* [P,3]  (x,y scaled by 0.5*W, 0.5*H; z = 0) */
    float* dL_dsh;            /* [P,M,3] */
    float* dL_dcolors;        /* [P,3]  gradient w.r.t. the per-Gaussian RGB (colors_precomp) */
    float* dL_dopacity;       /* [P] */
    float* dL_dscales;        /* [P,3] */
    float* dL_drotations;     /* [P,4] */
    float* dL_dcov3D;         /* [P,6] */
    /* Fused gradient accumulation (no reference counterpart; the reference's caller adds the
     * returned gradient to .grad in a separate pass).  For every VEGS_ACC_* bit set, the matching
     * output pointer must hold valid data and receives old + gradient (rows of Gaussians with
     * radii == 0 are left untouched); for bits not set the output is overwritten as above.
     * dL_dmeans2D is always overwritten. */
    uint32_t accumulate_mask;
    /* Fused densification statistics (no reference counterpart; the reference's caller computes
     * them from the returned means2D gradient and radii, scene/gaussian_model.py:411-413 and
     * train.py:300-301).  When stat_grad_accum is not NULL, for every Gaussian with radii > 0:
     *   stat_grad_accum[i] += || dL_dmeans2D[i].xy ||,  stat_denom[i] += 1,
     *   stat_max_radii[i]   = max(stat_max_radii[i], radii[i])
     * All three [P] float, all three or none. */
    float* stat_grad_accum;
    float* stat_denom;
    float* stat_max_radii;
    /* Backward of the fused activations / concatenation of the matching forward (see
     * VegsForwardArgs).  With a VEGS_ACT_* bit set the matching input is the RAW parameter and the
     * returned gradient is with respect to it; `opacities` (raw, [P]) is only read for
     * VEGS_ACT_OPACITY_SIGMOID.  With shs_rest given, dL_dsh receives the f_dc part [P,1,3] and
     * dL_dsh_rest the f_rest part [P,M-1,3] (VEGS_ACC_SH applies to both). */
    uint32_t activations;
    const float* opacities;
    const float* shs_rest;
    float* dL_dsh_rest;
    /* Opt-in LIVE gradient of the covariance maps (no reference counterpart: the reference computes
     * such terms and then overwrites them, SURVEY.md Q1, so by default - 0 - they are ignored here
     * too).  With live_map_gradients != 0 the cotangents dL_dout_cov_scale / dL_dout_cov_quat reach
     * dL_dscales / dL_drotations with the blending weights w = alpha * T held constant:
     *   dL_dscales[j]    += sum_pix w_j(pix) * dL_dout_cov_scale[:, pix]          (exact for fixed weights)
     *   dL_drotations[j] += sum_pix w_j(pix) * P_pix dL_dout_cov_quat[:, pix]     (first order in w: the
     *       slerp mix(identity, q_j, w) ~ identity + w (q_j - identity); P_pix projects onto the
     *       tangent space of the normalised output quaternion, out_cov_quat of the forward, required)
     * No term flows through the weights themselves (they do not feed dL/dalpha). */
    int32_t live_map_gradients;
    const float* out_cov_quat;    /* [4,H,W] from the forward; only read when live_map_gradients != 0 */
    /* Deferred SH gradient (no reference counterpart).  The gradient of the colour with respect to
     * the SH coefficients of one view is an outer product, dL_dsh[k][c] = basis_k(view direction) *
     * dL_dRGB[c] (clamped channels zeroed).  When dL_drgb_deferred [P,3] is given, the kernel writes
     * that dL_dRGB (zeros for radii == 0) INSTEAD of the dL_dsh rows (dL_dsh / dL_dsh_rest may then
     * be NULL; VEGS_ACC_SH is ignored) and vegs_sh_grad_from_views rebuilds the sum over a step's
     * views from 12 bytes per view and Gaussian instead of 12 M bytes - what a view-parallel step
     * exchanges between GPUs.  dL_dmeans3D still contains the SH direction term. */
    float* dL_drgb_deferred;
} VegsBackwardArgs;

#define VEGS_ACC_MEANS3D   1u
#define VEGS_ACC_SH        2u
#define VEGS_ACC_COLORS    4u
#define VEGS_ACC_OPACITY   8u
#define VEGS_ACC_SCALES    16u
#define VEGS_ACC_ROTATIONS 32u
#define VEGS_ACC_COV3D     64u

/* Library / ABI identification. */
int vegs_abi_version(void);
const char* vegs_last_error(void);

/* Buffer sizes (bytes).  R = num_rendered. */
size_t vegs_geom_bytes(int32_t P);
size_t vegs_scratch_bytes(int32_t P);
size_t vegs_img_bytes(int32_t W, int32_t H);
size_t vegs_binning_bytes(int64_t R);   /* 68 B per instance: the sorted list + room for the per-unit hit lists the
                                         * forward pass leaves for the backward (worst case; ~10 % is written) */
size_t vegs_bin_scratch_bytes(int64_t R, int32_t W, int32_t H);
size_t vegs_bwd_scratch_bytes(int32_t P);

/*
 * Forward, stage 1: per-Gaussian preprocess, depth sort, tile-count scan.
 * Replaces FORWARD::preprocess + InclusiveSum + the num_rendered read-back
 * (cuda_rasterizer/rasterizer_impl.cu:253-286).  Enqueues an asynchronous copy of num_rendered
 * into args->num_rendered_host[0] (and of the emission work-item count into [1]); the caller
 * synchronises `stream`, reads [0], allocates binning / bin_scratch, and calls
 * vegs_forward_render.
 */
int vegs_forward_prepare(const VegsForwardArgs* args, void* stream);

/*
 * Forward, stage 2: tile-instance emission, tile sort, tile ranges, front-to-back compositing.
 * Replaces duplicateWithKeys + SortPairs + identifyTileRanges + FORWARD::render
 * (cuda_rasterizer/rasterizer_impl.cu:288-350).
 */
int vegs_forward_render(const VegsForwardArgs* args, int32_t num_rendered, void* stream);

/*
 * Backward: blend-stage gradients (each compositing unit sums an entry's gradient over its pixels in
 * registers and issues three 16-byte vector reductions per entry) then the per-Gaussian stage.
 * Replaces CudaRasterizer::Rasterizer::backward (cuda_rasterizer/rasterizer_impl.cu:356-481).
 */
int vegs_backward(const VegsBackwardArgs* args, void* stream);

/*
 * markVisible: present[i] = (view * p_i).z > 0.2.  Replaces `_C.mark_visible`
 * (rasterize_points.cu:242-261, cuda_rasterizer/rasterizer_impl.cu:54-66,141-153).
 */
int vegs_mark_visible(int32_t P, const float* means3D, const float* viewmatrix,
                      const float* projmatrix, uint8_t* present, void* stream);

/*
 * distCUDA2: out[i] = mean of the three smallest squared distances from point i to other points.
 * Replaces `simple_knn._C.distCUDA2` (submodules/simple-knn/spatial.cu:15-26,
 * simple_knn.cu:185-221).  `workspace` must hold vegs_dist2_workspace_bytes(P) bytes.
 */
size_t vegs_dist2_workspace_bytes(int32_t P);
int vegs_dist2(int32_t P, const float* points, float* out, void* workspace, size_t workspace_bytes,
               void* stream);

/*
 * distCUDA2 for callers that cannot allocate device memory.  It has no counterpart in the
 * reference (whose knn allocates and frees temporaries on every call, simple_knn.cu:187-220).
 *
 * Meant for hosts where cudaMalloc / cudaFree are not an option: an application that gives this
 * library a fixed memory budget and keeps all allocation to itself, or a latency-sensitive loop
 * that cannot tolerate the device-wide serialisation those calls cause.  The caller passes only
 * the input points and the output array; the scratch for the search structure is a
 * VEGS_DIST2_NOALLOC_SCRATCH_BYTES array built into the library image, which every GPU that loads
 * the library keeps resident until the process exits.  That is why the point count is limited:
 * more than VEGS_DIST2_NOALLOC_MAX_POINTS points are refused with VEGS_ESIZE (use vegs_dist2).
 *
 * The first call on each device looks up the array's device address (a one-time set-up with that
 * device current: the calling thread's current device must be the one `stream` belongs to, as for
 * every entry point here); make that first call before starting any stream capture.  No call allocates or
 * frees memory.  All calls on a device share the one array, and it keeps whatever the previous
 * call left in it; nothing in it needs to be zero on entry, but two calls whose kernels overlap on
 * the GPU (different streams, or different host threads) race on it and give wrong results — run
 * them one after another.
 */
#define VEGS_DIST2_NOALLOC_MAX_POINTS 131072
#define VEGS_DIST2_NOALLOC_SCRATCH_BYTES (6u << 20)
int vegs_dist2_noalloc(int32_t P, const float* points, float* out, void* stream);

/*
 * Optimiser step over the view-parallel step's buffers in one pass (no counterpart inside the
 * reference's extensions; it replaces the caller's torch.optim.Adam step, zero_grad and the folding
 * of the step's densification statistics: scene/gaussian_model.py:154-172,411-413,
 * train.py:300-301,319-320).  Up to VEGS_ADAM_MAX_SEGMENTS contiguous float segments per call:
 *   kind 0  Adam:  exp_avg += (1-b1)(g - exp_avg);  exp_avg_sq = b2 exp_avg_sq + (1-b2) g g;
 *                  param -= (lr / (1-b1^step)) * exp_avg / (sqrt(exp_avg_sq) / sqrt(1-b2^step) + eps)
 *                  (torch.optim.Adam without weight decay / amsgrad; `step` counts from 1)
 *   kind 1  param += grad      (running statistic += this step's reduced value)
 *   kind 2  param  = max(param, grad)
 * With zero_grad != 0 every `grad` element is set to zero after it was read.
 */
#define VEGS_ADAM_MAX_SEGMENTS 12
typedef struct VegsAdamSegment {
    float* param;
    float* grad;
    float* exp_avg;       /* kind 0 only */
    float* exp_avg_sq;    /* kind 0 only */
    int64_t count;        /* elements */
    float lr;             /* kind 0 only */
    int32_t kind;
} VegsAdamSegment;
typedef struct VegsAdamArgs {
    VegsAdamSegment segments[VEGS_ADAM_MAX_SEGMENTS];
    int32_t num_segments;
    int32_t step;         /* t of the bias corrections, >= 1 */
    int32_t zero_grad;
    double beta1, beta2, eps;   /* doubles, as Python holds them: 1 - beta is formed in double and
                                 * rounded to float once, which is what torch's kernels receive */
} VegsAdamArgs;
int vegs_adam_step(const VegsAdamArgs* args, void* stream);

/*
 * SH gradient of a step from the per-view colour gradients (no counterpart in the reference; see
 * VegsBackwardArgs.dL_drgb_deferred):
 *   dL_dsh[i][k][c] (+)= sum_v basis_k(normalize(means3D[i] - campos[v])) * dL_drgb[v][i][c],  k < (D+1)^2,
 * the remaining coefficients of the M stored ones get zero.  campos [V,3] and dL_drgb [V,P,3] are
 * device arrays.  With dL_dsh_rest != NULL the result is split as VEGS's f_dc [P,1,3] (dL_dsh) and
 * f_rest [P,M-1,3] (dL_dsh_rest).  accumulate != 0 adds to the destination.  M <= 16.
 */
int vegs_sh_grad_from_views(int32_t P, int32_t D, int32_t M, int32_t V, const float* means3D,
                            const float* campos, const float* dL_drgb, float* dL_dsh, float* dL_dsh_rest,
                            int32_t accumulate, void* stream);

/*
 * Normal-guidance loss on the covariance maps, value and both gradients in one pass (no
 * counterpart inside the reference's extensions; it replaces the Python loss of
 * loss/normal_guidance.py:3-23):
 *   loss = w_rot * mean|Rot(q)^T n_w| + w_scale * mean|scale * (Rot(q)^T n_w)|,  n_w = R_cam2world n_cam,
 * the rotation detached in the second term as the reference does.  cov_quat [4,H,W] (w,x,y,z planes),
 * cov_scale [3,H,W], normal_cam [3,H,W] are device arrays; R_cam2world_host points at 9 HOST floats
 * (row-major); loss is one device float; d_cov_quat [4,H,W] / d_cov_scale [3,H,W] receive
 * dloss/dmap and may be NULL.  The reference uses w_rot = 0.8, w_scale = 0.2.
 */
int vegs_normal_guidance(int32_t W, int32_t H, const float* cov_quat, const float* cov_scale,
                         const float* normal_cam, const float* R_cam2world_host, float w_rot, float w_scale,
                         float* loss, float* d_cov_quat, float* d_cov_scale, void* stream);

/*
 * Test / parity support: re-express the binning state in the reference's format
 * (cuda_rasterizer/rasterizer_impl.cu:70-111,116-138): point_list_keys[R] = (tile << 32) | depth
 * bits, point_list[R], ranges[T] (uint2), tiles_touched[P], depths[P], means2D[P,2].  Any output
 * pointer may be NULL.  Not on the hot path.
 */
int vegs_export_state(const VegsForwardArgs* args, int32_t num_rendered, uint64_t* point_list_keys,
                      uint32_t* point_list, uint32_t* ranges, uint32_t* tiles_touched,
                      float* depths, float* means2D, float* conic_opacity, float* rgb,
                      uint32_t* n_contrib, void* stream);

/*
 * Diagnostics: SM cycles / 1024 each tile's CTA spent in the forward / backward compositing
 * kernel of the last call that used `img` (device arrays of T = ceil(W/16)*ceil(H/16) words).
 * `dbg` receives 8 extra counters per tile that are only filled by a -DVEGS_TILE_DEBUG=1 build.
 */
int vegs_export_tile_clocks(const void* img, int32_t W, int32_t H, uint32_t* fwd, uint32_t* bwd,
                            uint32_t* dbg /* [8*T], debug builds only, may be NULL */, void* stream);

/* Number of kernels launched by this library in this process so far (bench bookkeeping). */
uint64_t vegs_kernel_launch_count(void);

/*
 * Measurement support (bench.py's roofline leg): while enabled, a CUDA event is recorded on the
 * launching stream after every kernel; vegs_profile_collect synchronises those events and writes
 * one text line per kernel name, "name total_ms launches\n", into buf, then forgets them.
 */
void vegs_profile_enable(int32_t on);
int vegs_profile_collect(char* buf, size_t buf_bytes);

/*
 * Test hook (no reference counterpart).  The instance list carries, next to each Gaussian id, an
 * 8-bit mask of the compositing units it reaches whenever every id fits 24 bits (P < 2^24); larger
 * scenes use plain ids and test the box per entry.  on = 0 forces the plain-id path for any P so
 * the tests can cover it at small sizes; on = 1 restores the default.  Process-wide; takes effect
 * at the next vegs_forward_prepare.
 */
void vegs_debug_unit_masks(int32_t on);

#ifdef __cplusplus
}
#endif
#endif /* VEGS_B200_H_INCLUDED */
