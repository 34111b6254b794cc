// Forward per-Gaussian stage and markVisible.
//
// Replaces FORWARD::preprocess / preprocessCUDA (reference cuda_rasterizer/forward.cu:156-257),
// computeCov3D (:119-153), computeCov2D (:75-114), computeColorFromSH (:21-72) and checkFrustum
// (cuda_rasterizer/rasterizer_impl.cu:54-66).
//
// One thread per Gaussian, 128-thread blocks.  What differs from the reference's kernel:
//   * the 192-byte SH rows are fetched warp-cooperatively with coalesced 16-byte loads (only rows of
//     Gaussians that survived all culls) into padded shared memory, instead of 48 strided scalar
//     loads per thread;
//   * results are written as packed 16-byte records (recA..recE, see common.cuh) so the
//     compositing kernels gather one aligned float4 per field group;
//   * the depth sort key and the packed tile rectangle are produced here, so no separate
//     duplicateWithKeys pass has to re-derive them.
// Roofline: HBM.  Algorithmic bytes per Gaussian: 20 (mean + radius + key) for all P plus
// 12+16+4+12*M inputs and 80+8+1 outputs for the visible ones.
#include <cuda_fp16.h>
#include "common.cuh"
#include "gsmath.cuh"

namespace vegs {

namespace {

constexpr int kPreThreads = 128;
constexpr int kPreWarps = kPreThreads / 32;
constexpr int kMaxStageCoeffs = 16;                       // SH degree <= 3
constexpr int kShStrideMax = 3 * kMaxStageCoeffs + 1;     // 49 floats, odd -> conflict-free rows

struct PreParams {
    int P, D, M, W, H, gx, gy;
    const float* means;
    const float* scales;
    const float* rots;
    const float* opac;
    const float* shs;
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
    bool shs_vec_ok;
    bool rots_vec_ok;
};

__device__ __forceinline__ float4 ldg_stream4(const float4* p) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p));
    return v;
}

// SH -> RGB (+0.5, clamp at 0, clamp bits).  sh points at this Gaussian's row, layout [k][rgb].
__device__ __forceinline__ float3 sh_to_rgb(const float* sh, int deg, float dx, float dy, float dz,
                                            uint32_t& clampbits) {
    const float len = sqrtf(__fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy))));
    const float x = dx / len, y = dy / len, z = dz / len;
    float c[3] = {kSH_C0 * sh[0], kSH_C0 * sh[1], kSH_C0 * sh[2]};
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

// kM: SH coefficients per Gaussian when known at compile time (1, 4, 9, 16), 0 = read p.M.  The
// staging loop divides element indices by the row length; with a constant that is a multiply.
template <int kM>
__global__ void __launch_bounds__(kPreThreads) preprocess_fwd_kernel(const PreParams p) {
    const int M = kM ? kM : p.M;
    __shared__ float s_sh[kPreWarps][32 * kShStrideMax];
    __shared__ Camera c_cam;
    load_camera(c_cam, p.view, p.proj, p.campos);

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int idx = blockIdx.x * kPreThreads + threadIdx.x;
    const int warp_base = idx - lane;
    const bool in_range = idx < p.P;

    bool visible = false;
    float mx = 0.f, my = 0.f, mz = 0.f;
    float depth = 0.f, pix_x = 0.f, pix_y = 0.f, opacity = 0.f;
    float3 conic = make_float3(0.f, 0.f, 0.f);
    int radius = 0;
    int x0 = 0, y0 = 0, x1 = 0, y1 = 0;

    // Every per-Gaussian input is requested up front (nine in ten Gaussians survive the culls):
    // the loads overlap instead of queueing behind the depth test.
    float in_sx = 0.f, in_sy = 0.f, in_sz = 0.f, in_op = 0.f;
    float4 in_q = make_float4(1.f, 0.f, 0.f, 0.f);
    if (in_range) {
        mx = p.means[3 * (size_t)idx + 0];
        my = p.means[3 * (size_t)idx + 1];
        mz = p.means[3 * (size_t)idx + 2];
        if (p.scales != nullptr) {
            in_sx = p.scales[3 * (size_t)idx + 0];
            in_sy = p.scales[3 * (size_t)idx + 1];
            in_sz = p.scales[3 * (size_t)idx + 2];
        }
        if (p.rots != nullptr) in_q = load_quat(p.rots, idx, p.rots_vec_ok);
        in_op = p.opac[idx];
        depth = affine_row(c_cam.view, 2, 6, 10, 14, mx, my, mz);
        // Near cull, the only frustum test (auxiliary.h:154); NaN depth passes, as in the reference.
        if (!(depth <= 0.2f)) {
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
    if (p.colors_precomp != nullptr) {
        if (visible) {
            rgb.x = p.colors_precomp[3 * (size_t)idx + 0];
            rgb.y = p.colors_precomp[3 * (size_t)idx + 1];
            rgb.z = p.colors_precomp[3 * (size_t)idx + 2];
        }
    } else {
        const int rowlen = 3 * M;
        const unsigned vis = __ballot_sync(0xffffffffu, visible);
        if (vis != 0u) {
            const float ddx = mx - c_cam.campos[0];
            const float ddy = my - c_cam.campos[1];
            const float ddz = mz - c_cam.campos[2];
            if (M <= kMaxStageCoeffs) {
                const int stride = rowlen | 1;
                float* srow = s_sh[warp];
                const bool full = (warp_base + 32 <= p.P);
                if (full && p.shs_vec_ok) {
                    const float4* src =
                        reinterpret_cast<const float4*>(p.shs + (size_t)warp_base * rowlen);
                    const int nvec = 8 * rowlen;  // 32 rows * rowlen floats / 4
                    for (int f = lane; f < nvec; f += 32) {
                        const int e0 = 4 * f;
                        const int r0 = e0 / rowlen;
                        const int r1 = (e0 + 3) / rowlen;
                        if (((vis >> r0) | (vis >> r1)) & 1u) {
                            const float4 v = ldg_stream4(src + f);
                            const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                            for (int k = 0; k < 4; k++) {
                                const int e = e0 + k;
                                const int r = e / rowlen;
                                srow[r * stride + (e - r * rowlen)] = vv[k];
                            }
                        }
                    }
                } else if (visible) {
                    const float* src = p.shs + (size_t)idx * rowlen;
                    const int need = 3 * (p.D + 1) * (p.D + 1);
                    for (int k = 0; k < need && k < rowlen; k++) srow[lane * stride + k] = src[k];
                }
                __syncwarp();
                if (visible) rgb = sh_to_rgb(srow + lane * stride, p.D, ddx, ddy, ddz, clampbits);
            } else if (visible) {
                rgb = sh_to_rgb(p.shs + (size_t)idx * rowlen, p.D, ddx, ddy, ddz, clampbits);
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
    p.shs = a.shs; p.cov3D_precomp = a.cov3D_precomp; p.colors_precomp = a.colors_precomp;
    p.mod = a.scale_modifier; p.tan_fovx = a.tan_fovx; p.tan_fovy = a.tan_fovy;
    p.focal_y = a.H / (2.0f * a.tan_fovy);
    p.focal_x = a.W / (2.0f * a.tan_fovx);
    p.radii = a.radii;
    p.recA = g.recA; p.recB = g.recB; p.recC = g.recC; p.recD = s.recD; p.recE = s.recE;
    p.rect = s.rect; p.pbox = s.pbox; p.clamped = g.clamped; p.keys = s.keys0;
    p.shs_vec_ok = a.shs != nullptr && (reinterpret_cast<uintptr_t>(a.shs) & 15u) == 0;
    p.rots_vec_ok = a.rotations != nullptr && (reinterpret_cast<uintptr_t>(a.rotations) & 15u) == 0;
    const int blocks = (a.P + kPreThreads - 1) / kPreThreads;
    switch (a.M) {
        case 16: preprocess_fwd_kernel<16><<<blocks, kPreThreads, 0, stream>>>(p); break;
        case 9:  preprocess_fwd_kernel<9><<<blocks, kPreThreads, 0, stream>>>(p); break;
        case 4:  preprocess_fwd_kernel<4><<<blocks, kPreThreads, 0, stream>>>(p); break;
        case 1:  preprocess_fwd_kernel<1><<<blocks, kPreThreads, 0, stream>>>(p); break;
        default: preprocess_fwd_kernel<0><<<blocks, kPreThreads, 0, stream>>>(p); break;
    }
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
