// Forward per-Gaussian stage and markVisible.
//
// Replaces FORWARD::preprocess / preprocessCUDA (reference cuda_rasterizer/forward.cu:156-257),
// computeCov3D (:119-153), computeCov2D (:75-114), computeColorFromSH (:21-72) and checkFrustum
// (cuda_rasterizer/rasterizer_impl.cu:54-66).
//
// One thread per Gaussian, 128-thread blocks, every warp works on its own 32 consecutive Gaussians.
// What differs from the reference's kernel:
//   * a warp's inputs arrive as asynchronous bulk copies (bulk.cuh, the 1-D form of TMA): the
//     contiguous blocks of 32 means / scales / rotations / opacities by four copies issued by one
//     lane, and the 192-byte SH row of every Gaussian that survives the near cull by one copy per
//     row into a 208-byte-stride shared row (conflict-free for 16-byte reads), all tracked by
//     mbarriers - no load instruction, no register staging and no address arithmetic per element,
//     instead of 48 strided scalar loads per thread.  SH rows that cannot be bulk-copied (row size
//     not a multiple of 16 bytes, unaligned base) take a warp-cooperative 16-byte-load path;
//   * results are written as packed 16-byte records (recA..recE, see common.cuh) so the
//     compositing kernels gather one aligned float4 per field group;
//   * the depth sort key and the packed tile rectangle are produced here, so no separate
//     duplicateWithKeys pass has to re-derive them;
//   * optionally (VegsForwardArgs.activations / shs_rest, no reference counterpart) the caller's
//     activation passes are applied here: sigmoid(opacity), exp(scale), normalize(rotation) and the
//     concatenation of f_dc / f_rest (scene/gaussian_model.py:100-120).
// Roofline: HBM.  Algorithmic bytes per Gaussian: 20 (mean + radius + key) for all P plus
// 12+16+4+12*M inputs and 80+8+1 outputs for the visible ones.
#include <cuda_fp16.h>
#include "common.cuh"
#include "gsmath.cuh"
#include "bulk.cuh"
#include "pwide.cuh"

namespace vegs {

namespace {

struct PreParams {
    int P, D, M, W, H, gx, gy;
    const float* means;
    const float* scales;
    const float* rots;
    const float* opac;
    const float* shs;          // [P,M,3], or f_dc [P,1,3] when shs_rest is given
    const float* shs_rest;     // [P,M-1,3] or NULL
    const float* cov3D_precomp;
    const float* colors_precomp;
    float mod, tan_fovx, tan_fovy, focal_x, focal_y;
    int* radii;
    float4 *recA, *recB, *recC, *recD, *recE;
    uint2* rect;
    uint2* pbox;
    uint8_t* clamped;
    uint32_t* keys;
    const float* view;
    const float* proj;
    const float* campos;
    uint32_t act;              // VEGS_ACT_* bits
    bool in_bulk_ok;           // means / scales / rotations / opacities 16-byte aligned
    bool shs_vec_ok;
    bool rots_vec_ok;
};

// SH -> RGB (+0.5, clamp at 0, clamp bits).  dc0 points at coefficient 0 of this Gaussian's row
// (layout [k][rgb]) and sh at the same row's origin for coefficients >= 1, i.e. coefficient k >= 1
// is sh[3k + ch]: the two differ only for the split f_dc / f_rest layout.
__device__ __forceinline__ float3 sh_to_rgb(const float* dc0, const float* sh, int deg, float dx, float dy, float dz,
                                            uint32_t& clampbits) {
    const float len = sqrtf(__fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy))));
    const float x = dx / len, y = dy / len, z = dz / len;
    float c[3] = {kSH_C0 * dc0[0], kSH_C0 * dc0[1], kSH_C0 * dc0[2]};
    if (deg > 0) {
        const float a = kSH_C1 * y, b = kSH_C1 * z, d = kSH_C1 * x;
#pragma unroll
        for (int ch = 0; ch < 3; ch++) {
            float v = c[ch];
            v = v - a * sh[3 + ch] + b * sh[6 + ch] - d * sh[9 + ch];
            c[ch] = v;
        }
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z;
            const float xy = x * y, yz = y * z, xz = x * z;
            const float w4 = kSH_C2[0] * xy, w5 = kSH_C2[1] * yz,
                        w6 = kSH_C2[2] * (2.0f * zz - xx - yy), w7 = kSH_C2[3] * xz,
                        w8 = kSH_C2[4] * (xx - yy);
#pragma unroll
            for (int ch = 0; ch < 3; ch++) {
                float v = c[ch];
                v = v + w4 * sh[12 + ch] + w5 * sh[15 + ch] + w6 * sh[18 + ch] + w7 * sh[21 + ch] +
                    w8 * sh[24 + ch];
                c[ch] = v;
            }
            if (deg > 2) {
                const float w9 = kSH_C3[0] * y * (3.0f * xx - yy), w10 = kSH_C3[1] * xy * z,
                            w11 = kSH_C3[2] * y * (4.0f * zz - xx - yy),
                            w12 = kSH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy),
                            w13 = kSH_C3[4] * x * (4.0f * zz - xx - yy),
                            w14 = kSH_C3[5] * z * (xx - yy), w15 = kSH_C3[6] * x * (xx - 3.0f * yy);
#pragma unroll
                for (int ch = 0; ch < 3; ch++) {
                    float v = c[ch];
                    v = v + w9 * sh[27 + ch] + w10 * sh[30 + ch] + w11 * sh[33 + ch] +
                        w12 * sh[36 + ch] + w13 * sh[39 + ch] + w14 * sh[42 + ch] +
                        w15 * sh[45 + ch];
                    c[ch] = v;
                }
            }
        }
    }
    c[0] += 0.5f; c[1] += 0.5f; c[2] += 0.5f;
    clampbits = (c[0] < 0.0f ? 1u : 0u) | (c[1] < 0.0f ? 2u : 0u) | (c[2] < 0.0f ? 4u : 0u);
    return make_float3(fmaxf(c[0], 0.0f), fmaxf(c[1], 0.0f), fmaxf(c[2], 0.0f));
}

// kM: SH coefficients per Gaussian when known at compile time (16), 0 = read p.M.
// kSh: how the SH rows reach shared memory (pwide.cuh).
template <int kM, int kSh>
__global__ void __launch_bounds__(kPwThreads) preprocess_fwd_kernel(const PreParams p) {
    const int M = kM ? kM : p.M;
    __shared__ WarpStage s_stage[kPwWarps];
    __shared__ __align__(8) uint64_t s_bar[kPwWarps][2];     // [0] inputs, [1] SH rows
    __shared__ Camera c_cam;
    load_camera(c_cam, p.view, p.proj, p.campos);

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int idx = blockIdx.x * kPwThreads + threadIdx.x;
    const int warp_base = idx - lane;
    const bool in_range = idx < p.P;
    const bool full = warp_base + 32 <= p.P;
    WarpStage& st = s_stage[warp];
    uint64_t* bar_in = &s_bar[warp][0];
    uint64_t* bar_sh = &s_bar[warp][1];
    const bool have_sr = p.scales != nullptr;
    const bool in_bulk = full && p.in_bulk_ok;
    const bool use_sh = p.colors_precomp == nullptr;
    // split layout: the two blocks are copied whole, which needs the full warp in range
    const bool sh_bulk = use_sh && M <= kMaxStageCoeffs && (kSh == kShBulkRows || (kSh == kShSplit && full && p.shs_vec_ok));

    if (lane == 0) {
        mbar_init(bar_in, 1);
        mbar_init(bar_sh, kSh == kShBulkRows ? 32 : 1);
        mbar_init_fence();
    }
    __syncwarp();
    if (lane == 0) {
        if (in_bulk) issue_input_copies(st.in, bar_in, p.means, p.scales, p.rots, p.opac, warp_base);
        if (kSh == kShSplit && sh_bulk) {
            const uint32_t rest_bytes = 32u * 12u * (uint32_t)(M - 1);
            mbar_arrive_expect_tx(bar_sh, 384u + rest_bytes);
            bulk_g2s(st.sh + 32 * 3 * (M - 1), p.shs + 3 * (size_t)warp_base, 384u, bar_sh);
            if (rest_bytes) bulk_g2s(st.sh, p.shs_rest + (size_t)warp_base * 3 * (M - 1), rest_bytes, bar_sh);
        }
    }

    bool visible = false;
    float mx = 0.f, my = 0.f, mz = 0.f;
    float depth = 0.f, pix_x = 0.f, pix_y = 0.f, opacity = 0.f;
    float3 conic = make_float3(0.f, 0.f, 0.f);
    int radius = 0;
    int x0 = 0, y0 = 0, x1 = 0, y1 = 0;
    float in_sx = 0.f, in_sy = 0.f, in_sz = 0.f;
    float4 in_q = make_float4(1.f, 0.f, 0.f, 0.f);

    LaneInputs li;
    if (in_bulk) {
        mbar_wait(bar_in, 0);
        li = read_inputs_smem(st.in, lane, have_sr, true);
    } else if (in_range) {
        li = read_inputs_global(p.means, p.scales, p.rots, p.opac, idx, p.rots_vec_ok);
    } else {
        li = LaneInputs{0.f, 0.f, 0.f, 0.f, 0.f, 0.f, make_float4(1.f, 0.f, 0.f, 0.f), 0.f};
    }
    bool near_ok = false;
    if (in_range) {
        mx = li.mx; my = li.my; mz = li.mz;
        depth = affine_row(c_cam.view, 2, 6, 10, 14, mx, my, mz);
        // Near cull, the only frustum test (auxiliary.h:154); NaN depth passes, as in the reference.
        near_ok = !(depth <= 0.2f);
    }
    // The SH row of every Gaussian in front of the near plane is requested now, before the
    // projection arithmetic, so the copy runs underneath it.
    if (kSh == kShBulkRows && use_sh) {
        if (near_ok) {
            const uint32_t row_bytes = 12u * (uint32_t)M;
            mbar_arrive_expect_tx(bar_sh, row_bytes);
            bulk_g2s(st.sh + lane * kShRowStride, p.shs + (size_t)idx * 3 * M, row_bytes, bar_sh);
        } else {
            mbar_arrive(bar_sh);
        }
    }
    if (near_ok) {
        in_sx = li.sx; in_sy = li.sy; in_sz = li.sz; in_q = li.q;
        float in_op = li.op;
        if (p.act & VEGS_ACT_SCALE_EXP) { in_sx = expf(in_sx); in_sy = expf(in_sy); in_sz = expf(in_sz); }
        if (p.act & VEGS_ACT_ROT_NORMALIZE) { float dn; in_q = act_normalize(in_q, dn); }
        if (p.act & VEGS_ACT_OPACITY_SIGMOID) in_op = act_sigmoid(in_op);
        const float hx = affine_row(c_cam.proj, 0, 4, 8, 12, mx, my, mz);
        const float hy = affine_row(c_cam.proj, 1, 5, 9, 13, mx, my, mz);
        const float hw = affine_row(c_cam.proj, 3, 7, 11, 15, mx, my, mz);
        const float p_w = __fdiv_rn(1.0f, __fadd_rn(hw, 0.0000001f));
        const float projx = __fmul_rn(hx, p_w);
        const float projy = __fmul_rn(hy, p_w);

        float c3[6];
        if (p.cov3D_precomp != nullptr) {
#pragma unroll
            for (int k = 0; k < 6; k++) c3[k] = p.cov3D_precomp[6 * (size_t)idx + k];
        } else {
            cov3d_from_scale_rot(in_sx, in_sy, in_sz, p.mod, in_q.x, in_q.y, in_q.z, in_q.w, c3);
        }
        Cov2DTerms tt;
        const float3 cov = cov2d_project(mx, my, mz, p.focal_x, p.focal_y, p.tan_fovx,
                                         p.tan_fovy, c3, c_cam.view, tt);
        const float det = __fmaf_rn(cov.x, cov.z, -__fmul_rn(cov.y, cov.y));
        if (det != 0.0f) {
            const float det_inv = __fdiv_rn(1.0f, det);
            conic = make_float3(__fmul_rn(cov.z, det_inv), __fmul_rn(-cov.y, det_inv),
                                __fmul_rn(cov.x, det_inv));
            const float mid = __fmul_rn(0.5f, __fadd_rn(cov.x, cov.z));
            const float sq = __fsqrt_rn(fmaxf(0.1f, __fmaf_rn(mid, mid, -det)));
            const float lambda1 = __fadd_rn(mid, sq);
            const float lambda2 = __fsub_rn(mid, sq);
            const float rad_f = ceilf(__fmul_rn(3.0f, __fsqrt_rn(fmaxf(lambda1, lambda2))));
            radius = (int)rad_f;  // float -> int conversion, as `radii[idx] = my_radius`
            pix_x = ndc2pix_ref(projx, p.W);
            pix_y = ndc2pix_ref(projy, p.H);
            tile_rect(pix_x, pix_y, radius, p.gx, p.gy, x0, y0, x1, y1);
            if ((x1 - x0) * (y1 - y0) != 0) {
                visible = true;
                opacity = in_op;
            }
        }
    }

    // ---- per-Gaussian integer outputs for every Gaussian
    if (in_range) {
        p.radii[idx] = visible ? radius : 0;
        p.keys[idx] = visible ? __float_as_uint(depth) : 0xFFFFFFFFu;
        p.rect[(size_t)kRectStride * idx] = visible ? make_uint2((uint32_t)x0 | ((uint32_t)x1 << 16),
                                           (uint32_t)y0 | ((uint32_t)y1 << 16))
                              : make_uint2(0u, 0u);
    }

    // ---- colour
    float3 rgb = make_float3(0.f, 0.f, 0.f);
    uint32_t clampbits = 0;
    if (!use_sh) {
        if (visible) {
            rgb.x = p.colors_precomp[3 * (size_t)idx + 0];
            rgb.y = p.colors_precomp[3 * (size_t)idx + 1];
            rgb.z = p.colors_precomp[3 * (size_t)idx + 2];
        }
    } else {
        const float ddx = mx - c_cam.campos[0];
        const float ddy = my - c_cam.campos[1];
        const float ddz = mz - c_cam.campos[2];
        if (M > kMaxStageCoeffs) {
            // no staging for more than 16 coefficients (never the case in VEGS)
            if (visible) {
                const float* row = p.shs + (size_t)idx * 3 * M;
                rgb = sh_to_rgb(row, row, p.D, ddx, ddy, ddz, clampbits);
            }
        } else if (kSh == kShBulkRows) {
            mbar_wait(bar_sh, 0);
            if (visible) {
                // 16-byte reads of the 208-byte-stride row; the values land in registers
                float shv[3 * kMaxStageCoeffs];
                const float4* row = reinterpret_cast<const float4*>(st.sh + lane * kShRowStride);
                const int nv = (3 * M) >> 2;
#pragma unroll
                for (int j = 0; j < 3 * kMaxStageCoeffs / 4; j++) {
                    if (j < nv) {
                        const float4 v = row[j];
                        shv[4 * j + 0] = v.x; shv[4 * j + 1] = v.y; shv[4 * j + 2] = v.z; shv[4 * j + 3] = v.w;
                    }
                }
                rgb = sh_to_rgb(shv, shv, p.D, ddx, ddy, ddz, clampbits);
            }
        } else if (kSh == kShSplit) {
            const int rs = 3 * (M - 1);            // rest row stride (odd for even M: conflict-free)
            float* dc = st.sh + 32 * rs + 3 * lane;
            float* rest = st.sh + lane * rs;
            if (sh_bulk) {
                mbar_wait(bar_sh, 0);
            } else {
                if (visible) {
                    const int need = 3 * (p.D + 1) * (p.D + 1) - 3;
                    for (int k = 0; k < 3; k++) dc[k] = p.shs[3 * (size_t)idx + k];
                    const float* src = p.shs_rest + (size_t)idx * rs;
                    for (int k = 0; k < need && k < rs; k++) rest[k] = src[k];
                }
                __syncwarp();
            }
            // coefficient k >= 1 of the cat'ed row is rest[3(k-1) + ch] = (rest - 3)[3k + ch]
            if (visible) rgb = sh_to_rgb(dc, rest - 3, p.D, ddx, ddy, ddz, clampbits);
        } else {
            const unsigned vis = __ballot_sync(0xffffffffu, visible);
            if (vis != 0u) {
                const int rowlen = 3 * M;
                stage_sh_ldg(st.sh, p.shs, rowlen, warp_base, p.P, vis, visible, idx, lane, p.shs_vec_ok,
                             3 * (p.D + 1) * (p.D + 1));
                __syncwarp();
                const float* row = st.sh + lane * (rowlen | 1);
                if (visible) rgb = sh_to_rgb(row, row, p.D, ddx, ddy, ddz, clampbits);
            }
        }
    }

    // ---- packed records for the Gaussians that will be rasterised
    if (visible) {
        float4 d4 = make_float4(0.f, 0.f, 0.f, 1.f);   // scale 0, identity rotation when absent
        float4 e4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (p.scales != nullptr) { d4.x = in_sx; d4.y = in_sy; d4.z = in_sz; }
        if (p.rots != nullptr) {
            d4.w = in_q.x;
            e4.x = in_q.y; e4.y = in_q.z; e4.z = in_q.w;
        }
        const float theta = acosf(d4.w);
        e4.w = theta;
        // Half extents (ex, ey) of the box that contains every pixel where this Gaussian can reach
        // alpha >= 1/255: that needs power >= -tau with tau = log(255*opacity), i.e. the ellipse
        // A dx^2 + 2 B dx dy + C dy^2 <= 2 tau, whose extents are sqrt(2 tau C / det), sqrt(2 tau A / det).
        // Margins (0.01 on tau, 1e-3 relative + 0.01 px on the extents, upward fp16 rounding) keep the
        // box conservative, so the compositing kernels may skip the pixels outside it.  Encodings:
        // (+inf, +inf) = "cannot bound, test the whole tile" (non-finite inputs, for which the
        // reference's fminf(0.99, NaN) = 0.99 makes every pixel contribute); (-1, -1) = no pixel.
        float ex = __int_as_float(0x7f800000), ey = ex;
        {
            const float tau = logf(255.0f * opacity) + 0.01f;
            const bool fin = isfinite(pix_x) && isfinite(pix_y) && isfinite(conic.x) &&
                             isfinite(conic.y) && isfinite(conic.z) && isfinite(tau);
            if (fin) {
                const float cdet = conic.x * conic.z - conic.y * conic.y;
                if (tau < 0.0f) {
                    ex = ey = -1.0f;
                } else if (cdet > 0.0f) {
                    // The reference evaluates the power in float; its rounding error scales with the
                    // cancelling terms A dx^2, C dy^2, 2 B dx dy, which reach 1e5 - 1e6 at the far end
                    // of a long thin splat.  Widen tau by 2e-6 (~32 float epsilons) of their size at
                    // the corner of the first-guess box, twice over because the box grows with tau.
                    float s2 = 2.0f * tau / cdet;
                    float tx = sqrtf(s2 * conic.z), ty = sqrtf(s2 * conic.x);
                    const float mag = conic.x * tx * tx + conic.z * ty * ty + 2.0f * fabsf(conic.y) * tx * ty;
                    s2 = 2.0f * (tau + 4e-6f * mag) / cdet;
                    tx = sqrtf(s2 * conic.z) * 1.002f + 0.01f;
                    ty = sqrtf(s2 * conic.x) * 1.002f + 0.01f;
                    if (isfinite(tx) && isfinite(ty)) { ex = tx; ey = ty; }
                }
            }
        }
        const __half2 eh = __floats2half2_rn(ex, ey);   // 1.002 above covers the 2^-11 rounding
        p.recA[(size_t)kGeomRecStride * idx] = make_float4(pix_x, pix_y, __uint_as_float(*reinterpret_cast<const uint32_t*>(&eh)), depth);
        {
            // The same box in whole pixels, clipped to the image, for the emission kernel's unit
            // masks: float operations of unit_box() (composite_common.cuh) with the image as the block.
            const float2 e = __half22float2(eh);
            const int bx0 = (int)ceilf(fmaxf(pix_x - e.x, 0.0f));
            const int bx1 = (int)floorf(fminf(pix_x + e.x, (float)(p.W - 1)));
            const int by0 = (int)ceilf(fmaxf(pix_y - e.y, 0.0f));
            const int by1 = (int)floorf(fminf(pix_y + e.y, (float)(p.H - 1)));
            p.pbox[(size_t)kRectStride * idx] = make_uint2((uint32_t)min(bx0, p.W) | ((uint32_t)(max(bx1, -1) + 1) << 16),
                                     (uint32_t)min(by0, p.H) | ((uint32_t)(max(by1, -1) + 1) << 16));
        }
        p.recB[(size_t)kGeomRecStride * idx] = make_float4(conic.x, conic.y, conic.z, opacity);
        p.recC[(size_t)kGeomRecStride * idx] = make_float4(rgb.x, rgb.y, rgb.z, __fdiv_rn(1.0f, sinf(theta)));
        p.recC[(size_t)kGeomRecStride * idx + 1] = make_float4(0.f, 0.f, 0.f, 0.f);   // pad: whole sectors are written
        p.recD[(size_t)kScratchRecStride * idx] = d4;
        p.recE[(size_t)kScratchRecStride * idx] = e4;
        p.clamped[idx] = (uint8_t)clampbits;
    }
}

__global__ void __launch_bounds__(256) mark_visible_kernel(int P, const float* __restrict__ means,
                                                           const float* __restrict__ view,
                                                           uint8_t* __restrict__ present) {
    __shared__ Camera c_cam;
    load_camera(c_cam, view, nullptr, nullptr);
    const int idx = blockIdx.x * 256 + threadIdx.x;
    if (idx >= P) return;
    const float x = means[3 * (size_t)idx + 0];
    const float y = means[3 * (size_t)idx + 1];
    const float z = means[3 * (size_t)idx + 2];
    const float depth = affine_row(c_cam.view, 2, 6, 10, 14, x, y, z);
    present[idx] = (depth <= 0.2f) ? 0 : 1;
}

}  // namespace

int launch_preprocess_fwd(const VegsForwardArgs& a, const GeomView& g, const ScratchView& s,
                          cudaStream_t stream) {
    PreParams p;
    p.view = a.viewmatrix; p.proj = a.projmatrix; p.campos = a.campos;
    p.P = a.P; p.D = a.D; p.M = a.M; p.W = a.W; p.H = a.H;
    p.gx = (a.W + kTile - 1) / kTile;
    p.gy = (a.H + kTile - 1) / kTile;
    p.means = a.means3D; p.scales = a.scales; p.rots = a.rotations; p.opac = a.opacities;
    p.shs = a.shs; p.shs_rest = a.shs_rest; p.cov3D_precomp = a.cov3D_precomp; p.colors_precomp = a.colors_precomp;
    p.mod = a.scale_modifier; p.tan_fovx = a.tan_fovx; p.tan_fovy = a.tan_fovy;
    p.focal_y = a.H / (2.0f * a.tan_fovy);
    p.focal_x = a.W / (2.0f * a.tan_fovx);
    p.radii = a.radii;
    p.recA = g.recA; p.recB = g.recB; p.recC = g.recC; p.recD = s.recD; p.recE = s.recE;
    p.rect = s.rect; p.pbox = s.pbox; p.clamped = g.clamped; p.keys = s.keys0;
    p.act = a.activations;
    auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };
    p.in_bulk_ok = al16(a.means3D) && al16(a.opacities) && al16(a.scales) && al16(a.rotations);
    p.shs_vec_ok = a.shs != nullptr && al16(a.shs) && al16(a.shs_rest);
    p.rots_vec_ok = a.rotations != nullptr && al16(a.rotations);
    if (a.shs_rest != nullptr && (a.shs == nullptr || a.M < 1 || a.M > kMaxStageCoeffs)) {
        set_error("shs_rest needs shs (= f_dc) and 1 <= M <= %d", kMaxStageCoeffs);
        return VEGS_EINVAL;
    }
    // how the SH rows travel (pwide.cuh)
    int mode = kShLdg;
    if (a.shs_rest != nullptr) mode = kShSplit;
    else if (a.shs != nullptr && a.M <= kMaxStageCoeffs && (12 * a.M) % 16 == 0 && p.shs_vec_ok && !sh_bulk_disabled())
        mode = kShBulkRows;
    const int blocks = (a.P + kPwThreads - 1) / kPwThreads;
#define VEGS_PRE_LAUNCH(KM, MODE) preprocess_fwd_kernel<KM, MODE><<<blocks, kPwThreads, 0, stream>>>(p)
    if (a.M == 16) {
        if (mode == kShBulkRows) VEGS_PRE_LAUNCH(16, kShBulkRows);
        else if (mode == kShSplit) VEGS_PRE_LAUNCH(16, kShSplit);
        else VEGS_PRE_LAUNCH(16, kShLdg);
    } else {
        if (mode == kShBulkRows) VEGS_PRE_LAUNCH(0, kShBulkRows);
        else if (mode == kShSplit) VEGS_PRE_LAUNCH(0, kShSplit);
        else VEGS_PRE_LAUNCH(0, kShLdg);
    }
#undef VEGS_PRE_LAUNCH
    VEGS_LAUNCH_CHECK("preprocess_fwd", stream, a.debug);
    return VEGS_OK;
}

int launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present,
                        cudaStream_t stream) {
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, means3D, view, present);
    VEGS_LAUNCH_CHECK("mark_visible", stream, false);
    return VEGS_OK;
}

}  // namespace vegs
