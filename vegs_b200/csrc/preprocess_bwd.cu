// Per-Gaussian backward stage, one fused P-wide kernel.
//
// Replaces BACKWARD::preprocess = computeCov2DCUDA + preprocessCUDA + computeColorFromSH +
// computeCov3D (reference cuda_rasterizer/backward.cu:20-139, :144-274, :278-343, :348-414) and
// the ten torch::zeros fills of RasterizeGaussiansBackwardCUDA (rasterize_points.cu:181-191).
//
// What differs from the reference:
//   * one kernel instead of two, and no zero-fill passes: every element of every returned gradient
//     tensor is written exactly once here (zeros for Gaussians with radii == 0);
//   * every warp works on its own 32 consecutive Gaussians and moves their data with asynchronous
//     bulk copies (bulk.cuh, the 1-D form of TMA) tracked by mbarriers: the contiguous input blocks
//     (means, scales, rotations, radii, the packed 48-byte blend-stage accumulators) by a handful
//     of copies issued by one lane, the 192-byte SH row of every visible Gaussian by one copy per
//     row; on the way out the warp's gradient blocks are staged in shared memory and leave as bulk
//     stores - or, for fused gradient accumulation, as bulk REDUCTIONS (cp.reduce.async.bulk
//     .add.f32) that do the read-modify-write in L2, so the old .grad never travels to the SM;
//   * the SH backward runs in a streaming per-coefficient form (sh_bwd.cuh), four coefficients =
//     three 16-byte shared-memory accesses at a time;
//   * the 3D covariance is recomputed from scale/rotation instead of being stored by the forward
//     and re-read (saves 48 B per Gaussian of traffic);
//   * optionally the Jacobians of the caller's activations (sigmoid / exp / normalize, split
//     f_dc / f_rest) are applied here (VegsBackwardArgs.activations, no reference counterpart).
// The gradient formulas are the reference's; dL/dscale and dL/drot do not receive the quaternion /
// scale-map cotangents unless the opt-in live-map gradient is requested and the
// quaternion gradient has no normalisation Jacobian of its own (backward.cu:340-341).
// Roofline: HBM.  Algorithmic bytes per Gaussian: reads 4 (radius) for all, +12+28+48+1+12*M for
// visible; writes 12+12+4+12+16+12*M (+12 colours, +24 cov3D when requested) for all.
#include "common.cuh"
#include "gsmath.cuh"
#include "bulk.cuh"
#include "pwide.cuh"
#include "sh_bwd.cuh"

namespace vegs {

namespace {

struct PbParams {
    int P, D, M, W, H;
    const float* means;
    const int* radii;
    const float* shs;            // [P,M,3], or f_dc when shs_rest is given
    const float* shs_rest;
    const uint8_t* clamped;
    const float* scales;
    const float* rots;
    const float* opac;           // raw opacities, only read when VEGS_ACT_OPACITY_SIGMOID
    const float* cov3D_precomp;
    float mod, tan_fovx, tan_fovy, focal_x, focal_y;
    const float* view;
    const float* proj;
    const float* campos;
    const float4* acc;
    const float4* map_acc;       // [2P] live quaternion / scale-map gradients (opt-in) or NULL
    float* dL_dmeans3D;
    float* dL_dmeans2D;
    float* dL_dsh;
    float* dL_dsh_rest;
    float* dL_drgb;      // deferred SH gradient: the clamp-masked dL/dRGB [P,3] instead of dL/dSH rows, or NULL
    float* dL_dcolors;
    float* dL_dopacity;
    float* dL_dscales;
    float* dL_drots;
    float* dL_dcov3D;
    bool shs_vec_ok, dsh_vec_ok, rots_vec_ok;
    bool in_bulk_ok;             // means, scales, rotations, radii, acc 16-byte aligned
    bool out_bulk_ok;            // the dense outputs that are present are 16-byte aligned
    uint32_t accum;      // VEGS_ACC_* bits: those outputs receive old + gradient
    uint32_t act;        // VEGS_ACT_* bits
    float *stat_grad, *stat_denom, *stat_maxr;   // fused densification statistics, or NULL
};

// Dense output staging of one warp (overlays the input blocks once they are in registers).
constexpr int kOutMeans3D = 0, kOutMeans2D = 96, kOutScales = 192, kOutRots = 288, kOutOpac = 416, kOutRgb = 448;   // 544 floats in all

struct __align__(16) BwdWarpStage {
    float sh[kShWarpFloats];              // SH rows in, dL/dSH rows out
    float io[736];                        // inputs: means 96 | scales 96 | rots 128 | opac 32 | acc 384; then outputs
    int radii[32];
};
constexpr int kIoAcc = kInWarpFloats;    // 352: the 32 x 12 accumulator floats follow the four input blocks

// One group of four SH coefficients of one lane's row: load, backward, store.
//   kShBulkRows  row = the lane's 52-float-stride row, three 16-byte accesses
//   kShSplit     dc = coefficient 0, row[3k + ch] = coefficient k >= 1 (row = rest - 3), scalar accesses
//   kShLdg       row[3k + ch], dc = row, scalar accesses
template <int kSh, int G>
__device__ __forceinline__ void sh_bwd_row_group(ShBwdState& s, float* row, float* dc, int M, bool zero_only) {
    if (4 * G >= M) return;
    float g[12];
    if (kSh == kShBulkRows) {
        float4* r4 = reinterpret_cast<float4*>(row) + 3 * G;
        if (!zero_only) {
            const float4 a = r4[0], b = r4[1], c = r4[2];
            g[0] = a.x; g[1] = a.y; g[2] = a.z; g[3] = a.w; g[4] = b.x; g[5] = b.y; g[6] = b.z; g[7] = b.w;
            g[8] = c.x; g[9] = c.y; g[10] = c.z; g[11] = c.w;
            sh_bwd_group<G>(s, g);
        } else {
#pragma unroll
            for (int i = 0; i < 12; i++) g[i] = 0.f;
        }
        r4[0] = make_float4(g[0], g[1], g[2], g[3]);
        r4[1] = make_float4(g[4], g[5], g[6], g[7]);
        r4[2] = make_float4(g[8], g[9], g[10], g[11]);
    } else {
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int k = 4 * G + j;
            float* ptr = (k == 0) ? dc : row + 3 * k;
#pragma unroll
            for (int ch = 0; ch < 3; ch++) g[3 * j + ch] = (!zero_only && k < M) ? ptr[ch] : 0.f;
        }
        if (!zero_only) sh_bwd_group<G>(s, g);
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int k = 4 * G + j;
            if (k < M) {
                float* ptr = (k == 0) ? dc : row + 3 * k;
#pragma unroll
                for (int ch = 0; ch < 3; ch++) ptr[ch] = g[3 * j + ch];
            }
        }
    }
}

// kM: SH coefficients per Gaussian when known at compile time (16), 0 = read p.M.
// kSh: how the SH rows travel (pwide.cuh).
template <int kM, int kSh>
__global__ void __launch_bounds__(kPwThreads, 5) preprocess_bwd_kernel(const PbParams p) {
    const int M = kM ? kM : p.M;
    __shared__ BwdWarpStage s_stage[kPwWarps];
    __shared__ __align__(8) uint64_t s_bar[kPwWarps][2];     // [0] inputs, [1] SH rows
    __shared__ Camera cam;
    load_camera(cam, p.view, p.proj, p.campos);

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int idx = blockIdx.x * kPwThreads + threadIdx.x;
    const int warp_base = idx - lane;
    const bool in_range = idx < p.P;
    const bool full = warp_base + 32 <= p.P;
    BwdWarpStage& st = s_stage[warp];
    uint64_t* bar_in = &s_bar[warp][0];
    uint64_t* bar_sh = &s_bar[warp][1];
    const bool have_sr = p.cov3D_precomp == nullptr;
    const bool in_bulk = full && p.in_bulk_ok;
    const bool out_bulk = full && p.out_bulk_ok;
    const bool defer_sh = p.dL_drgb != nullptr;    // only dL/dRGB leaves the kernel; dL/dSH is rebuilt later (sh_views.cu)
    const bool use_sh = p.shs != nullptr && (p.dL_dsh != nullptr || defer_sh);
    const bool staged_sh = use_sh && M <= kMaxStageCoeffs;
    const bool sh_block_bulk = staged_sh && kSh == kShSplit && full && p.shs_vec_ok;   // whole-warp blocks in
    const bool act_op = (p.act & VEGS_ACT_OPACITY_SIGMOID) != 0u;

    if (lane == 0) {
        mbar_init(bar_in, 1);
        mbar_init(bar_sh, kSh == kShBulkRows ? 32 : 1);
        mbar_init_fence();
    }
    __syncwarp();
    if (lane == 0) {
        if (in_bulk) {
            issue_input_copies(st.io, bar_in, p.means, have_sr ? p.scales : nullptr, have_sr ? p.rots : nullptr,
                               act_op ? p.opac : nullptr, warp_base, 128u + 1536u);
            bulk_g2s(st.radii, p.radii + warp_base, 128u, bar_in);
            bulk_g2s(st.io + kIoAcc, p.acc + 3 * (size_t)warp_base, 1536u, bar_in);
        }
        if (sh_block_bulk) {
            const uint32_t rest_bytes = 32u * 12u * (uint32_t)(M - 1);
            mbar_arrive_expect_tx(bar_sh, 384u + rest_bytes);
            bulk_g2s(st.sh + 32 * 3 * (M - 1), p.shs + 3 * (size_t)warp_base, 384u, bar_sh);
            if (rest_bytes) bulk_g2s(st.sh, p.shs_rest + (size_t)warp_base * 3 * (M - 1), rest_bytes, bar_sh);
        }
    }

    // ---- inputs
    LaneInputs li = LaneInputs{0.f, 0.f, 0.f, 0.f, 0.f, 0.f, make_float4(1.f, 0.f, 0.f, 0.f), 0.f};
    float4 in_a0 = make_float4(0.f, 0.f, 0.f, 0.f), in_a1 = in_a0, in_a2 = in_a0;
    int in_radius = 0;
    if (in_bulk) {
        mbar_wait(bar_in, 0);
        li = read_inputs_smem(st.io, lane, have_sr, act_op);
        in_radius = st.radii[lane];
        const float4* a4 = reinterpret_cast<const float4*>(st.io + kIoAcc) + 3 * lane;
        in_a0 = a4[0]; in_a1 = a4[1]; in_a2 = a4[2];
    } else if (in_range) {
        li = read_inputs_global(p.means, have_sr ? p.scales : nullptr, p.rots, act_op ? p.opac : nullptr, idx,
                                p.rots_vec_ok);
        in_radius = p.radii[idx];
        in_a0 = p.acc[3 * (size_t)idx + 0];
        in_a1 = p.acc[3 * (size_t)idx + 1];
        in_a2 = p.acc[3 * (size_t)idx + 2];
    }
    const bool visible = in_range && in_radius > 0;
    // the SH row of every visible Gaussian is requested now and arrives underneath the arithmetic
    if (kSh == kShBulkRows && staged_sh) {
        if (visible) {
            const uint32_t row_bytes = 12u * (uint32_t)M;
            mbar_arrive_expect_tx(bar_sh, row_bytes);
            bulk_g2s(st.sh + lane * kShRowStride, p.shs + (size_t)idx * 3 * M, row_bytes, bar_sh);
        } else {
            mbar_arrive(bar_sh);
        }
    }

    // activated scale / rotation (what the forward rasterised) and what their Jacobians need
    float sx = li.sx, sy = li.sy, sz = li.sz;
    float4 q = li.q;
    float q_denom = 1.0f;
    if (have_sr) {
        if (p.act & VEGS_ACT_SCALE_EXP) { sx = expf(sx); sy = expf(sy); sz = expf(sz); }
        if (p.act & VEGS_ACT_ROT_NORMALIZE) q = act_normalize(q, q_denom);
    }

    float dmean[3] = {0.f, 0.f, 0.f};
    float dm2d[2] = {0.f, 0.f};
    float dcol[3] = {0.f, 0.f, 0.f};
    float dopac = 0.f;
    float dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float dscale[3] = {0.f, 0.f, 0.f};
    float drot[4] = {0.f, 0.f, 0.f, 0.f};
    float mx = 0.f, my = 0.f, mz = 0.f;

    if (visible) {
        mx = li.mx; my = li.my; mz = li.mz;
        const float4 a0 = in_a0;  // dmean2D.xy, dconic.xx, dconic.xy
        const float4 a1 = in_a1;  // dconic.yy, dopacity, dcol.r, dcol.g
        const float4 a2 = in_a2;  // dcol.b, ddepth
        dm2d[0] = a0.x; dm2d[1] = a0.y;
        const float dcx = a0.z, dcy = a0.w, dcz = a1.x;
        dopac = a1.y;
        dcol[0] = a1.z; dcol[1] = a1.w; dcol[2] = a2.x;
        const float ddepth = a2.y;

        // ---- 3D covariance (recomputed) from the activated scale / rotation
        float c3[6];
        if (p.cov3D_precomp != nullptr) {
#pragma unroll
            for (int k = 0; k < 6; k++) c3[k] = p.cov3D_precomp[6 * (size_t)idx + k];
        } else {
            cov3d_from_scale_rot(sx, sy, sz, p.mod, q.x, q.y, q.z, q.w, c3);
        }

        // ---- conic -> cov2D -> cov3D and mean (computeCov2DCUDA, backward.cu:144-274)
        Cov2DTerms t;
        const float3 cov = cov2d_project(mx, my, mz, p.focal_x, p.focal_y, p.tan_fovx, p.tan_fovy,
                                         c3, cam.view, t);
        const float limx = 1.3f * p.tan_fovx, limy = 1.3f * p.tan_fovy;
        const float x_grad_mul = (t.txtz < -limx || t.txtz > limx) ? 0.f : 1.f;
        const float y_grad_mul = (t.tytz < -limy || t.tytz > limy) ? 0.f : 1.f;
        const float a = cov.x, b = cov.y, c = cov.z;
        const float denom = a * c - b * b;
        float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        if (denom2inv != 0.f) {
            dL_da = denom2inv * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcz);
            dL_dc = denom2inv * (-a * a * dcz + 2 * a * b * dcy + (denom - a * c) * dcx);
            dL_db = denom2inv * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
            dcov[0] = (t.T00 * t.T00 * dL_da + t.T00 * t.T10 * dL_db + t.T10 * t.T10 * dL_dc);
            dcov[3] = (t.T01 * t.T01 * dL_da + t.T01 * t.T11 * dL_db + t.T11 * t.T11 * dL_dc);
            dcov[5] = (t.T02 * t.T02 * dL_da + t.T02 * t.T12 * dL_db + t.T12 * t.T12 * dL_dc);
            dcov[1] = 2 * t.T00 * t.T01 * dL_da + (t.T00 * t.T11 + t.T01 * t.T10) * dL_db +
                      2 * t.T10 * t.T11 * dL_dc;
            dcov[2] = 2 * t.T00 * t.T02 * dL_da + (t.T00 * t.T12 + t.T02 * t.T10) * dL_db +
                      2 * t.T10 * t.T12 * dL_dc;
            dcov[4] = 2 * t.T02 * t.T01 * dL_da + (t.T01 * t.T12 + t.T02 * t.T11) * dL_db +
                      2 * t.T11 * t.T12 * dL_dc;
        }
        // Vrk columns: V0 = (c0,c1,c2), V1 = (c1,c3,c4), V2 = (c2,c4,c5)
        const float T0V0 = t.T00 * c3[0] + t.T01 * c3[1] + t.T02 * c3[2];
        const float T0V1 = t.T00 * c3[1] + t.T01 * c3[3] + t.T02 * c3[4];
        const float T0V2 = t.T00 * c3[2] + t.T01 * c3[4] + t.T02 * c3[5];
        const float T1V0 = t.T10 * c3[0] + t.T11 * c3[1] + t.T12 * c3[2];
        const float T1V1 = t.T10 * c3[1] + t.T11 * c3[3] + t.T12 * c3[4];
        const float T1V2 = t.T10 * c3[2] + t.T11 * c3[4] + t.T12 * c3[5];
        const float dL_dT00 = 2 * T0V0 * dL_da + T1V0 * dL_db;
        const float dL_dT01 = 2 * T0V1 * dL_da + T1V1 * dL_db;
        const float dL_dT02 = 2 * T0V2 * dL_da + T1V2 * dL_db;
        const float dL_dT10 = 2 * T1V0 * dL_dc + T0V0 * dL_db;
        const float dL_dT11 = 2 * T1V1 * dL_dc + T0V1 * dL_db;
        const float dL_dT12 = 2 * T1V2 * dL_dc + T0V2 * dL_db;
        const float* v = cam.view;
        // W[0] = (v0,v4,v8), W[1] = (v1,v5,v9), W[2] = (v2,v6,v10)
        const float dL_dJ00 = v[0] * dL_dT00 + v[4] * dL_dT01 + v[8] * dL_dT02;
        const float dL_dJ02 = v[2] * dL_dT00 + v[6] * dL_dT01 + v[10] * dL_dT02;
        const float dL_dJ11 = v[1] * dL_dT10 + v[5] * dL_dT11 + v[9] * dL_dT12;
        const float dL_dJ12 = v[2] * dL_dT10 + v[6] * dL_dT11 + v[10] * dL_dT12;
        const float tz = 1.f / t.tz;
        const float tz2 = tz * tz;
        const float tz3 = tz2 * tz;
        const float h_x = p.focal_x, h_y = p.focal_y;
        const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
        const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
        const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 +
                             (2 * h_x * t.tx) * tz3 * dL_dJ02 + (2 * h_y * t.ty) * tz3 * dL_dJ12;
        // transformVec4x3Transpose
        dmean[0] = v[0] * dL_dtx + v[1] * dL_dty + v[2] * dL_dtz;
        dmean[1] = v[4] * dL_dtx + v[5] * dL_dty + v[6] * dL_dtz;
        dmean[2] = v[8] * dL_dtx + v[9] * dL_dty + v[10] * dL_dtz;

        // ---- screen-space mean -> 3D mean (backward.cu:376-391)
        const float* pr = cam.proj;
        const float m_hw = pr[3] * mx + pr[7] * my + pr[11] * mz + pr[15];
        const float m_w = 1.0f / (m_hw + 0.0000001f);
        const float mul1 = (pr[0] * mx + pr[4] * my + pr[8] * mz + pr[12]) * m_w * m_w;
        const float mul2 = (pr[1] * mx + pr[5] * my + pr[9] * mz + pr[13]) * m_w * m_w;
        dmean[0] += (pr[0] * m_w - pr[3] * mul1) * dm2d[0] + (pr[1] * m_w - pr[3] * mul2) * dm2d[1];
        dmean[1] += (pr[4] * m_w - pr[7] * mul1) * dm2d[0] + (pr[5] * m_w - pr[7] * mul2) * dm2d[1];
        dmean[2] += (pr[8] * m_w - pr[11] * mul1) * dm2d[0] + (pr[9] * m_w - pr[11] * mul2) * dm2d[1];

        // ---- depth -> 3D mean (backward.cu:393-405)
        const float mul3 = v[2] * mx + v[6] * my + v[10] * mz + v[14];
        dmean[0] += (v[2] - v[3] * mul3) * ddepth;
        dmean[1] += (v[6] - v[7] * mul3) * ddepth;
        dmean[2] += (v[10] - v[11] * mul3) * ddepth;

        // ---- cov3D -> scale, rotation (backward.cu:278-343)
        if (p.scales != nullptr) {
            const float r = q.x, x = q.y, y = q.z, z = q.w;
            // R[col][row]
            const float R[3][3] = {
                {1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
            const float s[3] = {p.mod * sx, p.mod * sy, p.mod * sz};
            float M[3][3];
#pragma unroll
            for (int cc = 0; cc < 3; cc++)
#pragma unroll
                for (int rr = 0; rr < 3; rr++) M[cc][rr] = s[rr] * R[cc][rr];
            const float dS[3][3] = {{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]},
                                    {0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
                                    {0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}};
            // dL_dM = (2 M) * dSigma : dM[c][r] = sum_k 2 M[k][r] dS[c][k]
            float dM[3][3];
#pragma unroll
            for (int cc = 0; cc < 3; cc++)
#pragma unroll
                for (int rr = 0; rr < 3; rr++)
                    dM[cc][rr] = 2.0f * M[0][rr] * dS[cc][0] + 2.0f * M[1][rr] * dS[cc][1] +
                                 2.0f * M[2][rr] * dS[cc][2];
            // dL_dscale_i = sum_r R[r][i] * dM[r][i]
#pragma unroll
            for (int i = 0; i < 3; i++)
                dscale[i] = R[0][i] * dM[0][i] + R[1][i] * dM[1][i] + R[2][i] * dM[2][i];
            // dMt[c][r] = s_c * dM[r][c]
            float dMt[3][3];
#pragma unroll
            for (int cc = 0; cc < 3; cc++)
#pragma unroll
                for (int rr = 0; rr < 3; rr++) dMt[cc][rr] = s[cc] * dM[rr][cc];
            drot[0] = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) +
                      2 * x * (dMt[1][2] - dMt[2][1]);
            drot[1] = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) +
                      2 * r * (dMt[1][2] - dMt[2][1]) - 4 * x * (dMt[2][2] + dMt[1][1]);
            drot[2] = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) +
                      2 * z * (dMt[1][2] + dMt[2][1]) - 4 * y * (dMt[2][2] + dMt[0][0]);
            drot[3] = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) +
                      2 * y * (dMt[1][2] + dMt[2][1]) - 4 * z * (dMt[1][1] + dMt[0][0]);
        }
    }

    // ---- SH backward: dL/dSH rows built in the staging rows, dL/dmean through the view direction
    const bool acc_sh = (p.accum & VEGS_ACC_SH) != 0u;
    bool dsh_done = !use_sh || defer_sh;
    float dRGB[3] = {0.f, 0.f, 0.f};
    if (use_sh) {
        if (visible) {
            const uint32_t cb = p.clamped[idx];
            dRGB[0] = (cb & 1u) ? 0.f : dcol[0];
            dRGB[1] = (cb & 2u) ? 0.f : dcol[1];
            dRGB[2] = (cb & 4u) ? 0.f : dcol[2];
        }
        const float dox = mx - cam.campos[0], doy = my - cam.campos[1], doz = mz - cam.campos[2];
        if (staged_sh) {
            const int rowlen = 3 * M;
            float *row, *dc;
            if (kSh == kShBulkRows) {
                row = st.sh + lane * kShRowStride; dc = row;
                mbar_wait(bar_sh, 0);
            } else if (kSh == kShSplit) {
                const int rs = 3 * (M - 1);
                dc = st.sh + 32 * rs + 3 * lane;
                row = st.sh + lane * rs - 3;
                if (sh_block_bulk) {
                    mbar_wait(bar_sh, 0);
                } else {
                    if (visible) {
                        for (int k = 0; k < 3; k++) dc[k] = p.shs[3 * (size_t)idx + k];
                        const float* src = p.shs_rest + (size_t)idx * rs;
                        for (int k = 0; k < rs; k++) row[3 + k] = src[k];
                    }
                    __syncwarp();
                }
            } else {
                row = st.sh + lane * (rowlen | 1); dc = row;
                const unsigned vis = __ballot_sync(0xffffffffu, visible);
                if (vis != 0u) stage_sh_ldg(st.sh, p.shs, rowlen, warp_base, p.P, vis, visible, idx, lane, p.shs_vec_ok, rowlen);
                __syncwarp();
            }
            ShBwdState sb = sh_bwd_begin(p.D, visible ? dox : 1.f, doy, doz, dRGB);
            const bool zero_only = !visible;       // rows of culled Gaussians: zeros (ignored when accumulating)
            sh_bwd_row_group<kSh, 0>(sb, row, dc, M, zero_only);
            sh_bwd_row_group<kSh, 1>(sb, row, dc, M, zero_only);
            sh_bwd_row_group<kSh, 2>(sb, row, dc, M, zero_only);
            sh_bwd_row_group<kSh, 3>(sb, row, dc, M, zero_only);
            if (visible) {
                const float3 dm = sh_bwd_end(sb, dox, doy, doz);
                dmean[0] += dm.x; dmean[1] += dm.y; dmean[2] += dm.z;
            }
            // ---- dL/dSH rows to global memory (not when the SH gradient is deferred)
            if (defer_sh) {
            } else if (kSh == kShBulkRows && p.dsh_vec_ok) {
                // one bulk store (or L2 reduction, when accumulating) per row, issued by the row's lane
                fence_smem_to_async_proxy();
                if (in_range && (visible || !acc_sh)) {
                    float* dst = p.dL_dsh + (size_t)idx * rowlen;
                    if (acc_sh) bulk_s2g_add_f32(dst, row, 12u * (uint32_t)M);
                    else bulk_s2g(dst, row, 12u * (uint32_t)M);
                }
                dsh_done = true;
            } else if (kSh == kShSplit && full && p.dsh_vec_ok) {
                // the warp's two dense blocks leave as they came
                fence_smem_to_async_proxy();
                __syncwarp();
                if (lane == 0) {
                    const int rs = 3 * (M - 1);
                    float* d0 = p.dL_dsh + 3 * (size_t)warp_base;
                    float* d1 = p.dL_dsh_rest + (size_t)warp_base * rs;
                    if (acc_sh) {
                        bulk_s2g_add_f32(d0, st.sh + 32 * rs, 384u);
                        if (rs) bulk_s2g_add_f32(d1, st.sh, 32u * 4u * (uint32_t)rs);
                    } else {
                        bulk_s2g(d0, st.sh + 32 * rs, 384u);
                        if (rs) bulk_s2g(d1, st.sh, 32u * 4u * (uint32_t)rs);
                    }
                }
                dsh_done = true;
            } else if (kSh == kShLdg && full && p.dsh_vec_ok) {
                __syncwarp();
                const unsigned vis = __ballot_sync(0xffffffffu, visible);
                const int stride = rowlen | 1;
                const int nvec = 8 * rowlen;
                float4* dst = reinterpret_cast<float4*>(p.dL_dsh + (size_t)warp_base * rowlen);
                for (int f = lane; f < nvec; f += 32) {
                    float vv[4];
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const int e = 4 * f + k;
                        const int r = e / rowlen;
                        vv[k] = st.sh[r * stride + (e - r * rowlen)];
                    }
                    if (acc_sh) {
                        // rows of culled Gaussians add nothing: leave chunks that only touch them alone
                        const int r0 = (4 * f) / rowlen, r1 = (4 * f + 3) / rowlen;
                        if ((((vis >> r0) | (vis >> r1)) & 1u) == 0u) continue;
                        const float4 o = dst[f];
                        vv[0] += o.x; vv[1] += o.y; vv[2] += o.z; vv[3] += o.w;
                    }
                    dst[f] = make_float4(vv[0], vv[1], vv[2], vv[3]);
                }
                dsh_done = true;
            }
            if (!dsh_done && in_range && (!acc_sh || visible)) {
                // partial warp / unaligned output: the lane writes its own row
                if (kSh == kShSplit) {
                    float* d0 = p.dL_dsh + 3 * (size_t)idx;
                    float* d1 = p.dL_dsh_rest + (size_t)idx * 3 * (M - 1);
                    for (int k = 0; k < 3; k++) d0[k] = acc_sh ? d0[k] + dc[k] : dc[k];
                    for (int k = 0; k < 3 * (M - 1); k++) d1[k] = acc_sh ? d1[k] + row[3 + k] : row[3 + k];
                } else {
                    float* dst = p.dL_dsh + (size_t)idx * rowlen;
                    for (int k = 0; k < rowlen; k++) dst[k] = acc_sh ? dst[k] + row[k] : row[k];
                }
            }
        } else if (in_range && !defer_sh) {
            // M > 16: no staging; work in a local copy of the first 16 coefficients
            // (accumulation and deferral are refused for M > 16 by the launcher)
            float* dst = p.dL_dsh + (size_t)idx * 3 * M;
            if (visible) {
                const float* src = p.shs + (size_t)idx * 3 * M;
                float loc[48];
                for (int k = 0; k < 48; k++) loc[k] = src[k];
                ShBwdState sb = sh_bwd_begin(p.D, dox, doy, doz, dRGB);
                sh_bwd_group<0>(sb, loc); sh_bwd_group<1>(sb, loc + 12);
                sh_bwd_group<2>(sb, loc + 24); sh_bwd_group<3>(sb, loc + 36);
                const float3 dm = sh_bwd_end(sb, dox, doy, doz);
                dmean[0] += dm.x; dmean[1] += dm.y; dmean[2] += dm.z;
                for (int k = 0; k < 48; k++) dst[k] = loc[k];
                for (int k = 48; k < 3 * M; k++) dst[k] = 0.0f;
            } else {
                for (int k = 0; k < 3 * M; k++) dst[k] = 0.0f;
            }
        }
    }

    // ---- opt-in live gradient of the covariance maps (composite_bwd.cu), with respect to the
    // activated scale / rotation like the terms above
    if (p.map_acc != nullptr && visible && have_sr) {
        const float4 ms = p.map_acc[2 * (size_t)idx], mq = p.map_acc[2 * (size_t)idx + 1];
        dscale[0] += ms.x; dscale[1] += ms.y; dscale[2] += ms.z;
        drot[0] += mq.x; drot[1] += mq.y; drot[2] += mq.z; drot[3] += mq.w;
    }

    // ---- Jacobians of the fused activations (scene/gaussian_model.py:100-120 applied by the caller otherwise)
    float dm2_norm = 0.f;
    if (visible) {
        dm2_norm = sqrtf(dm2d[0] * dm2d[0] + dm2d[1] * dm2d[1]);
        if (act_op) { const float o = act_sigmoid(li.op); dopac *= o * (1.0f - o); }
        if (have_sr) {
            if (p.act & VEGS_ACT_SCALE_EXP) { dscale[0] *= sx; dscale[1] *= sy; dscale[2] *= sz; }
            if (p.act & VEGS_ACT_ROT_NORMALIZE) {
                // n = q_raw / max(|q_raw|, eps):  dL/dq_raw = (g - n <n, g>) / max(|q_raw|, eps)
                const float ng = q.x * drot[0] + q.y * drot[1] + q.z * drot[2] + q.w * drot[3];
                const float inv = 1.0f / q_denom;
                drot[0] = (drot[0] - q.x * ng) * inv; drot[1] = (drot[1] - q.y * ng) * inv;
                drot[2] = (drot[2] - q.z * ng) * inv; drot[3] = (drot[3] - q.w * ng) * inv;
            }
        }
    }

    // ---- dense outputs: overwritten, or added to (visible rows only) where accumulation is requested
    if (out_bulk) {
        // the warp's blocks are staged in shared memory (over the input blocks, which every lane has
        // in registers by now) and leave as one bulk store / reduction each; culled rows hold zeros,
        // which is what an overwrite must write and what an addition may add
        __syncwarp();
        float* o = st.io;
#pragma unroll
        for (int k = 0; k < 3; k++) {
            o[kOutMeans3D + 3 * lane + k] = dmean[k];
            o[kOutMeans2D + 3 * lane + k] = (k < 2) ? dm2d[k] : 0.0f;
            o[kOutScales + 3 * lane + k] = dscale[k];
        }
        *reinterpret_cast<float4*>(&o[kOutRots + 4 * lane]) = make_float4(drot[0], drot[1], drot[2], drot[3]);
        o[kOutOpac + lane] = dopac;
        if (defer_sh) {
#pragma unroll
            for (int k = 0; k < 3; k++) o[kOutRgb + 3 * lane + k] = dRGB[k];
        }
        fence_smem_to_async_proxy();
        __syncwarp();
        if (lane == 0) {
            auto put_block = [&](float* dst, int off, uint32_t bytes, uint32_t bit) {
                if (dst == nullptr) return;
                if (p.accum & bit) bulk_s2g_add_f32(dst, o + off, bytes);
                else bulk_s2g(dst, o + off, bytes);
            };
            put_block(p.dL_dmeans3D ? p.dL_dmeans3D + 3 * (size_t)warp_base : nullptr, kOutMeans3D, 384u, VEGS_ACC_MEANS3D);
            put_block(p.dL_dmeans2D ? p.dL_dmeans2D + 3 * (size_t)warp_base : nullptr, kOutMeans2D, 384u, 0u);
            put_block(p.dL_dopacity ? p.dL_dopacity + (size_t)warp_base : nullptr, kOutOpac, 128u, VEGS_ACC_OPACITY);
            put_block(p.dL_dscales ? p.dL_dscales + 3 * (size_t)warp_base : nullptr, kOutScales, 384u, VEGS_ACC_SCALES);
            put_block(p.dL_drots ? p.dL_drots + 4 * (size_t)warp_base : nullptr, kOutRots, 512u, VEGS_ACC_ROTATIONS);
            put_block(defer_sh ? p.dL_drgb + 3 * (size_t)warp_base : nullptr, kOutRgb, 384u, 0u);
        }
    }
    if (in_range) {
        const size_t i3 = 3 * (size_t)idx;
        auto put = [&](float* dst, int n, const float* v, uint32_t bit) {
            if (dst == nullptr) return;
            const bool acc = (p.accum & bit) != 0u;
            if (acc && !visible) return;
            for (int k = 0; k < n; k++) dst[k] = acc ? dst[k] + v[k] : v[k];
        };
        if (!out_bulk) {
            const float dm2[3] = {dm2d[0], dm2d[1], 0.0f};
            put(p.dL_dmeans3D ? p.dL_dmeans3D + i3 : nullptr, 3, dmean, VEGS_ACC_MEANS3D);
            put(p.dL_dmeans2D ? p.dL_dmeans2D + i3 : nullptr, 3, dm2, 0u);
            put(p.dL_dopacity ? p.dL_dopacity + idx : nullptr, 1, &dopac, VEGS_ACC_OPACITY);
            put(p.dL_dscales ? p.dL_dscales + i3 : nullptr, 3, dscale, VEGS_ACC_SCALES);
            put(p.dL_drots ? p.dL_drots + 4 * (size_t)idx : nullptr, 4, drot, VEGS_ACC_ROTATIONS);
            put(defer_sh ? p.dL_drgb + i3 : nullptr, 3, dRGB, 0u);
        }
        // the two rarely used outputs (colors_precomp / cov3D_precomp modes) are written directly
        put(p.dL_dcolors ? p.dL_dcolors + i3 : nullptr, 3, dcol, VEGS_ACC_COLORS);
        put(p.dL_dcov3D ? p.dL_dcov3D + 6 * (size_t)idx : nullptr, 6, dcov, VEGS_ACC_COV3D);
        // densification statistics of the reference's training loop (scene/gaussian_model.py:411-413,
        // train.py:300-301), for the Gaussians this view rendered
        if (p.stat_grad != nullptr && visible) {
            p.stat_grad[idx] += dm2_norm;
            p.stat_denom[idx] += 1.0f;
            p.stat_maxr[idx] = fmaxf(p.stat_maxr[idx], (float)in_radius);
        }
    }
    // the shared-memory sources of this thread's bulk stores must stay intact until they were read
    bulk_commit();
    bulk_wait_read_all();
}

}  // namespace

int launch_preprocess_bwd(const VegsBackwardArgs& a, const GeomView& g, const float4* acc,
                          const float4* map_acc, cudaStream_t stream) {
    PbParams p;
    p.P = a.P; p.D = a.D; p.M = a.M; p.W = a.W; p.H = a.H;
    p.means = a.means3D; p.radii = a.radii; p.shs = a.shs; p.shs_rest = a.shs_rest; p.clamped = g.clamped;
    p.scales = a.scales; p.rots = a.rotations; p.opac = a.opacities; p.cov3D_precomp = a.cov3D_precomp;
    p.mod = a.scale_modifier; p.tan_fovx = a.tan_fovx; p.tan_fovy = a.tan_fovy;
    p.focal_y = a.H / (2.0f * a.tan_fovy);
    p.focal_x = a.W / (2.0f * a.tan_fovx);
    p.view = a.viewmatrix; p.proj = a.projmatrix; p.campos = a.campos;
    p.acc = acc;
    p.map_acc = map_acc;
    p.dL_dmeans3D = a.dL_dmeans3D; p.dL_dmeans2D = a.dL_dmeans2D; p.dL_dsh = a.dL_dsh; p.dL_dsh_rest = a.dL_dsh_rest;
    p.dL_drgb = a.dL_drgb_deferred;
    p.dL_dcolors = a.dL_dcolors; p.dL_dopacity = a.dL_dopacity; p.dL_dscales = a.dL_dscales;
    p.dL_drots = a.dL_drotations; p.dL_dcov3D = a.dL_dcov3D;
    auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };   // NULL counts as aligned
    p.shs_vec_ok = a.shs != nullptr && al16(a.shs) && al16(a.shs_rest);
    p.dsh_vec_ok = a.dL_dsh != nullptr && al16(a.dL_dsh) && al16(a.dL_dsh_rest);
    p.rots_vec_ok = a.rotations != nullptr && al16(a.rotations);
    p.act = a.activations;
    p.accum = a.accumulate_mask;
    const bool bulk_off = sh_bulk_disabled();
    p.in_bulk_ok = !bulk_off && al16(a.means3D) && al16(a.radii) && al16(acc) &&
                   (a.cov3D_precomp != nullptr || (al16(a.scales) && al16(a.rotations))) &&
                   (!(a.activations & VEGS_ACT_OPACITY_SIGMOID) || al16(a.opacities));
    p.out_bulk_ok = !bulk_off && al16(a.dL_dmeans3D) && al16(a.dL_dmeans2D) && al16(a.dL_dopacity) &&
                    al16(a.dL_dscales) && al16(a.dL_drotations) && al16(a.dL_drgb_deferred);
    p.stat_grad = a.stat_grad_accum; p.stat_denom = a.stat_denom; p.stat_maxr = a.stat_max_radii;
    if ((p.stat_grad != nullptr) != (p.stat_denom != nullptr) || (p.stat_grad != nullptr) != (p.stat_maxr != nullptr)) {
        set_error("vegs_backward: stat_grad_accum, stat_denom and stat_max_radii go together");
        return VEGS_EINVAL;
    }
    if ((p.accum & VEGS_ACC_SH) && a.M > kMaxStageCoeffs) {
        set_error("vegs_backward: VEGS_ACC_SH needs M <= %d", kMaxStageCoeffs);
        return VEGS_EINVAL;
    }
    if ((a.activations & VEGS_ACT_OPACITY_SIGMOID) && a.opacities == nullptr) {
        set_error("vegs_backward: VEGS_ACT_OPACITY_SIGMOID needs the raw opacities");
        return VEGS_EINVAL;
    }
    if (a.dL_drgb_deferred != nullptr && (a.shs == nullptr || a.M > kMaxStageCoeffs)) {
        set_error("vegs_backward: dL_drgb_deferred needs shs and M <= %d", kMaxStageCoeffs);
        return VEGS_EINVAL;
    }
    if (a.shs_rest != nullptr && (a.shs == nullptr || (a.dL_dsh_rest == nullptr && a.dL_drgb_deferred == nullptr) || a.M < 1 ||
                                  a.M > kMaxStageCoeffs)) {
        set_error("vegs_backward: shs_rest needs shs (= f_dc), dL_dsh_rest (or the deferred form) and 1 <= M <= %d", kMaxStageCoeffs);
        return VEGS_EINVAL;
    }
    int mode = kShLdg;
    if (a.shs_rest != nullptr) mode = kShSplit;
    else if (a.shs != nullptr && (a.dL_dsh != nullptr || a.dL_drgb_deferred != nullptr) && a.M <= kMaxStageCoeffs && (12 * a.M) % 16 == 0 &&
             p.shs_vec_ok && !bulk_off)
        mode = kShBulkRows;
    const int blocks = (a.P + kPwThreads - 1) / kPwThreads;
#define VEGS_PB_LAUNCH(KM, MODE) preprocess_bwd_kernel<KM, MODE><<<blocks, kPwThreads, 0, stream>>>(p)
    if (a.M == 16) {
        if (mode == kShBulkRows) VEGS_PB_LAUNCH(16, kShBulkRows);
        else if (mode == kShSplit) VEGS_PB_LAUNCH(16, kShSplit);
        else VEGS_PB_LAUNCH(16, kShLdg);
    } else {
        if (mode == kShBulkRows) VEGS_PB_LAUNCH(0, kShBulkRows);
        else if (mode == kShSplit) VEGS_PB_LAUNCH(0, kShSplit);
        else VEGS_PB_LAUNCH(0, kShLdg);
    }
#undef VEGS_PB_LAUNCH
    VEGS_LAUNCH_CHECK("preprocess_bwd", stream, a.debug);
    return VEGS_OK;
}

}  // namespace vegs
