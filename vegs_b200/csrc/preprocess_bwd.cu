// Per-Gaussian backward stage, one fused P-wide kernel.
//
// Replaces BACKWARD::preprocess = computeCov2DCUDA + preprocessCUDA + computeColorFromSH +
// computeCov3D (reference cuda_rasterizer/backward.cu:20-139, :144-274, :278-343, :348-414) and
// the ten torch::zeros fills of RasterizeGaussiansBackwardCUDA (rasterize_points.cu:181-191).
//
// What differs from the reference:
//   * one kernel instead of two, and no zero-fill passes: every element of every returned gradient
//     tensor is written exactly once here (zeros for Gaussians with radii == 0);
//   * the blend-stage gradients arrive as one packed 48-byte record per Gaussian (three 16-byte
//     loads) instead of five separate arrays;
//   * SH coefficients are read and dL/dSH written warp-cooperatively with coalesced 16-byte
//     accesses through padded shared memory (the reference does 48 strided scalar loads and 48
//     strided scalar stores per thread);
//   * the 3D covariance is recomputed from scale/rotation instead of being stored by the forward
//     and re-read (saves 48 B per Gaussian of traffic).
// The gradient formulas are the reference's; dL/dscale and dL/drot do not receive the quaternion /
// scale-map cotangents and the quaternion gradient has no normalisation Jacobian
// (backward.cu:340-341).
// Roofline: HBM.  Algorithmic bytes per Gaussian: reads 4 (radius) for all, +12+28+48+1+12*M for
// visible; writes 12+12+4+12+16+12*M (+12 colours, +24 cov3D when requested) for all.
#include "common.cuh"
#include "gsmath.cuh"

namespace vegs {

namespace {

constexpr int kPbThreads = 128;
constexpr int kPbWarps = kPbThreads / 32;
constexpr int kMaxStageCoeffs = 16;
constexpr int kShStrideMax = 3 * kMaxStageCoeffs + 1;

struct PbParams {
    int P, D, M, W, H;
    const float* means;
    const int* radii;
    const float* shs;
    const uint8_t* clamped;
    const float* scales;
    const float* rots;
    const float* cov3D_precomp;
    float mod, tan_fovx, tan_fovy, focal_x, focal_y;
    const float* view;
    const float* proj;
    const float* campos;
    const float4* acc;
    float* dL_dmeans3D;
    float* dL_dmeans2D;
    float* dL_dsh;
    float* dL_dcolors;
    float* dL_dopacity;
    float* dL_dscales;
    float* dL_drots;
    float* dL_dcov3D;
    bool shs_vec_ok, dsh_vec_ok, rots_vec_ok;
    uint32_t accum;      // VEGS_ACC_* bits: those outputs receive old + gradient
    float *stat_grad, *stat_denom, *stat_maxr;   // fused densification statistics, or NULL
};

__device__ __forceinline__ float4 ldg_stream4(const float4* p) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p));
    return v;
}

// SH backward for one Gaussian.  `row` holds its coefficients [k][rgb] on entry (first
// (deg+1)^2 valid) and receives dL/dSH (all M coefficients; unused ones zero) on exit.
// Returns dL/d(mean) through the view direction.  Reference: backward.cu:20-139.
__device__ __forceinline__ float3 sh_backward(float* row, int deg, int M, float dox, float doy,
                                              float doz, const float dRGB[3]) {
    const float len = sqrtf(dox * dox + doy * doy + doz * doz);
    const float x = dox / len, y = doy / len, z = doz / len;
    float ddx[3] = {0.f, 0.f, 0.f}, ddy[3] = {0.f, 0.f, 0.f}, ddz[3] = {0.f, 0.f, 0.f};

    auto put = [&](int k, float w) {
#pragma unroll
        for (int ch = 0; ch < 3; ch++) row[3 * k + ch] = w * dRGB[ch];
    };
    put(0, kSH_C0);
    int used = 1;
    if (deg > 0) {
        float sh1[3], sh2[3], sh3[3];
#pragma unroll
        for (int ch = 0; ch < 3; ch++) { sh1[ch] = row[3 + ch]; sh2[ch] = row[6 + ch]; sh3[ch] = row[9 + ch]; }
        put(1, -kSH_C1 * y);
        put(2, kSH_C1 * z);
        put(3, -kSH_C1 * x);
#pragma unroll
        for (int ch = 0; ch < 3; ch++) {
            ddx[ch] = -kSH_C1 * sh3[ch];
            ddy[ch] = -kSH_C1 * sh1[ch];
            ddz[ch] = kSH_C1 * sh2[ch];
        }
        used = 4;
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z;
            const float xy = x * y, yz = y * z, xz = x * z;
            float s4[3], s5[3], s6[3], s7[3], s8[3];
#pragma unroll
            for (int ch = 0; ch < 3; ch++) {
                s4[ch] = row[12 + ch]; s5[ch] = row[15 + ch]; s6[ch] = row[18 + ch];
                s7[ch] = row[21 + ch]; s8[ch] = row[24 + ch];
            }
            put(4, kSH_C2[0] * xy);
            put(5, kSH_C2[1] * yz);
            put(6, kSH_C2[2] * (2.f * zz - xx - yy));
            put(7, kSH_C2[3] * xz);
            put(8, kSH_C2[4] * (xx - yy));
#pragma unroll
            for (int ch = 0; ch < 3; ch++) {
                ddx[ch] += kSH_C2[0] * y * s4[ch] + kSH_C2[2] * 2.f * -x * s6[ch] +
                           kSH_C2[3] * z * s7[ch] + kSH_C2[4] * 2.f * x * s8[ch];
                ddy[ch] += kSH_C2[0] * x * s4[ch] + kSH_C2[1] * z * s5[ch] +
                           kSH_C2[2] * 2.f * -y * s6[ch] + kSH_C2[4] * 2.f * -y * s8[ch];
                ddz[ch] += kSH_C2[1] * y * s5[ch] + kSH_C2[2] * 2.f * 2.f * z * s6[ch] +
                           kSH_C2[3] * x * s7[ch];
            }
            used = 9;
            if (deg > 2) {
                float s9[3], s10[3], s11[3], s12[3], s13[3], s14[3], s15[3];
#pragma unroll
                for (int ch = 0; ch < 3; ch++) {
                    s9[ch] = row[27 + ch]; s10[ch] = row[30 + ch]; s11[ch] = row[33 + ch];
                    s12[ch] = row[36 + ch]; s13[ch] = row[39 + ch]; s14[ch] = row[42 + ch];
                    s15[ch] = row[45 + ch];
                }
                put(9, kSH_C3[0] * y * (3.f * xx - yy));
                put(10, kSH_C3[1] * xy * z);
                put(11, kSH_C3[2] * y * (4.f * zz - xx - yy));
                put(12, kSH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
                put(13, kSH_C3[4] * x * (4.f * zz - xx - yy));
                put(14, kSH_C3[5] * z * (xx - yy));
                put(15, kSH_C3[6] * x * (xx - 3.f * yy));
#pragma unroll
                for (int ch = 0; ch < 3; ch++) {
                    ddx[ch] += kSH_C3[0] * s9[ch] * 3.f * 2.f * xy + kSH_C3[1] * s10[ch] * yz +
                               kSH_C3[2] * s11[ch] * -2.f * xy + kSH_C3[3] * s12[ch] * -3.f * 2.f * xz +
                               kSH_C3[4] * s13[ch] * (-3.f * xx + 4.f * zz - yy) +
                               kSH_C3[5] * s14[ch] * 2.f * xz + kSH_C3[6] * s15[ch] * 3.f * (xx - yy);
                    ddy[ch] += kSH_C3[0] * s9[ch] * 3.f * (xx - yy) + kSH_C3[1] * s10[ch] * xz +
                               kSH_C3[2] * s11[ch] * (-3.f * yy + 4.f * zz - xx) +
                               kSH_C3[3] * s12[ch] * -3.f * 2.f * yz + kSH_C3[4] * s13[ch] * -2.f * xy +
                               kSH_C3[5] * s14[ch] * -2.f * yz + kSH_C3[6] * s15[ch] * -3.f * 2.f * xy;
                    ddz[ch] += kSH_C3[1] * s10[ch] * xy + kSH_C3[2] * s11[ch] * 4.f * 2.f * yz +
                               kSH_C3[3] * s12[ch] * 3.f * (2.f * zz - xx - yy) +
                               kSH_C3[4] * s13[ch] * 4.f * 2.f * xz + kSH_C3[5] * s14[ch] * (xx - yy);
                }
                used = 16;
            }
        }
    }
    for (int k = 3 * used; k < 3 * M; k++) row[k] = 0.0f;

    const float dLx = ddx[0] * dRGB[0] + ddx[1] * dRGB[1] + ddx[2] * dRGB[2];
    const float dLy = ddy[0] * dRGB[0] + ddy[1] * dRGB[1] + ddy[2] * dRGB[2];
    const float dLz = ddz[0] * dRGB[0] + ddz[1] * dRGB[1] + ddz[2] * dRGB[2];
    // dnormvdv (auxiliary.h:107-117)
    const float sum2 = dox * dox + doy * doy + doz * doz;
    const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    float3 r;
    r.x = ((+sum2 - dox * dox) * dLx - doy * dox * dLy - doz * dox * dLz) * invsum32;
    r.y = (-dox * doy * dLx + (sum2 - doy * doy) * dLy - doz * doy * dLz) * invsum32;
    r.z = (-dox * doz * dLx - doy * doz * dLy + (sum2 - doz * doz) * dLz) * invsum32;
    return r;
}

// kM: SH coefficients per Gaussian when known at compile time (1, 4, 9, 16), 0 = read p.M.  The
// staging loops divide element indices by the row length; with a constant that is a multiply.
template <int kM>
__global__ void __launch_bounds__(kPbThreads) preprocess_bwd_kernel(const PbParams p) {
    const int M = kM ? kM : p.M;
    __shared__ float s_sh[kPbWarps][32 * kShStrideMax];
    __shared__ Camera cam;
    load_camera(cam, p.view, p.proj, p.campos);

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int idx = blockIdx.x * kPbThreads + threadIdx.x;
    const int warp_base = idx - lane;
    const bool in_range = idx < p.P;
    // Every per-Gaussian input is requested before the visibility test (nine in ten Gaussians are
    // visible): the loads overlap instead of queueing behind the radius.
    float in_m[3] = {0.f, 0.f, 0.f}, in_s[3] = {0.f, 0.f, 0.f};
    float4 in_q = make_float4(1.f, 0.f, 0.f, 0.f);
    float4 in_a0 = make_float4(0.f, 0.f, 0.f, 0.f), in_a1 = in_a0, in_a2 = in_a0;
    int in_radius = 0;
    if (in_range) {
        in_radius = p.radii[idx];
#pragma unroll
        for (int k = 0; k < 3; k++) in_m[k] = p.means[3 * (size_t)idx + k];
        in_a0 = p.acc[3 * (size_t)idx + 0];
        in_a1 = p.acc[3 * (size_t)idx + 1];
        in_a2 = p.acc[3 * (size_t)idx + 2];
        if (p.cov3D_precomp == nullptr) {
#pragma unroll
            for (int k = 0; k < 3; k++) in_s[k] = p.scales[3 * (size_t)idx + k];
            in_q = load_quat(p.rots, idx, p.rots_vec_ok);
        }
    }
    const bool visible = in_range && in_radius > 0;

    float dmean[3] = {0.f, 0.f, 0.f};
    float dm2d[2] = {0.f, 0.f};
    float dcol[3] = {0.f, 0.f, 0.f};
    float dopac = 0.f;
    float dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float dscale[3] = {0.f, 0.f, 0.f};
    float drot[4] = {0.f, 0.f, 0.f, 0.f};
    float mx = 0.f, my = 0.f, mz = 0.f;

    if (visible) {
        mx = in_m[0]; my = in_m[1]; mz = in_m[2];
        const float4 a0 = in_a0;  // dmean2D.xy, dconic.xx, dconic.xy
        const float4 a1 = in_a1;  // dconic.yy, dopacity, dcol.r, dcol.g
        const float4 a2 = in_a2;  // dcol.b, ddepth
        dm2d[0] = a0.x; dm2d[1] = a0.y;
        const float dcx = a0.z, dcy = a0.w, dcz = a1.x;
        dopac = a1.y;
        dcol[0] = a1.z; dcol[1] = a1.w; dcol[2] = a2.x;
        const float ddepth = a2.y;

        // ---- 3D covariance (recomputed) and scale/rotation
        float c3[6];
        float sx = 0.f, sy = 0.f, sz = 0.f;
        float4 q = make_float4(1.f, 0.f, 0.f, 0.f);
        if (p.cov3D_precomp != nullptr) {
#pragma unroll
            for (int k = 0; k < 6; k++) c3[k] = p.cov3D_precomp[6 * (size_t)idx + k];
        } else {
            sx = in_s[0]; sy = in_s[1]; sz = in_s[2];
            q = in_q;
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

    // ---- SH backward (row staged in shared memory, coalesced both ways)
    if (p.shs != nullptr && p.dL_dsh != nullptr) {
        const int rowlen = 3 * M;
        const unsigned vis = __ballot_sync(0xffffffffu, visible);
        float dRGB[3] = {0.f, 0.f, 0.f};
        if (visible) {
            const uint32_t cb = p.clamped[idx];
            dRGB[0] = (cb & 1u) ? 0.f : dcol[0];
            dRGB[1] = (cb & 2u) ? 0.f : dcol[1];
            dRGB[2] = (cb & 4u) ? 0.f : dcol[2];
        }
        const float dox = mx - cam.campos[0], doy = my - cam.campos[1], doz = mz - cam.campos[2];
        if (M <= kMaxStageCoeffs) {
            const int stride = rowlen | 1;
            float* srow = s_sh[warp];
            float* myrow = srow + lane * stride;
            const bool full = (warp_base + 32 <= p.P);
            const int nvec = 8 * rowlen;
            if (vis != 0u) {
                if (full && p.shs_vec_ok) {
                    const float4* src =
                        reinterpret_cast<const float4*>(p.shs + (size_t)warp_base * rowlen);
                    for (int f = lane; f < nvec; f += 32) {
                        const int e0 = 4 * f;
                        const int r0 = e0 / rowlen;
                        const int r1 = (e0 + 3) / rowlen;
                        if (((vis >> r0) | (vis >> r1)) & 1u) {
                            const float4 v4 = ldg_stream4(src + f);
                            const float vv[4] = {v4.x, v4.y, v4.z, v4.w};
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
                    for (int k = 0; k < rowlen; k++) myrow[k] = src[k];
                }
            }
            __syncwarp();
            if (visible) {
                const float3 dm = sh_backward(myrow, p.D, M, dox, doy, doz, dRGB);
                dmean[0] += dm.x; dmean[1] += dm.y; dmean[2] += dm.z;
            } else {
                for (int k = 0; k < rowlen; k++) myrow[k] = 0.0f;
            }
            __syncwarp();
            const bool acc_sh = (p.accum & VEGS_ACC_SH) != 0u;
            if (full && p.dsh_vec_ok) {
                float4* dst = reinterpret_cast<float4*>(p.dL_dsh + (size_t)warp_base * rowlen);
                for (int f = lane; f < nvec; f += 32) {
                    float vv[4];
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const int e = 4 * f + k;
                        const int r = e / rowlen;
                        vv[k] = srow[r * stride + (e - r * rowlen)];
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
            } else if (in_range && (!acc_sh || visible)) {
                float* dst = p.dL_dsh + (size_t)idx * rowlen;
                for (int k = 0; k < rowlen; k++) dst[k] = acc_sh ? dst[k] + myrow[k] : myrow[k];
            }
        } else if (in_range) {
            // M > 16: no staging; work in the output row directly
            // (accumulation is refused for M > 16 by the launcher)
            float* dst = p.dL_dsh + (size_t)idx * rowlen;
            if (visible) {
                const float* src = p.shs + (size_t)idx * rowlen;
                for (int k = 0; k < 48; k++) dst[k] = src[k];
                const float3 dm = sh_backward(dst, p.D, M, dox, doy, doz, dRGB);
                dmean[0] += dm.x; dmean[1] += dm.y; dmean[2] += dm.z;
            } else {
                for (int k = 0; k < rowlen; k++) dst[k] = 0.0f;
            }
        }
    }

    // ---- dense outputs: overwritten, or added to (visible rows only) where accumulation is requested
    if (in_range) {
        const size_t i3 = 3 * (size_t)idx;
        auto put = [&](float* dst, int n, const float* v, uint32_t bit) {
            if (dst == nullptr) return;
            const bool acc = (p.accum & bit) != 0u;
            if (acc && !visible) return;
            for (int k = 0; k < n; k++) dst[k] = acc ? dst[k] + v[k] : v[k];
        };
        const float dm2[3] = {dm2d[0], dm2d[1], 0.0f};
        put(p.dL_dmeans3D ? p.dL_dmeans3D + i3 : nullptr, 3, dmean, VEGS_ACC_MEANS3D);
        put(p.dL_dmeans2D ? p.dL_dmeans2D + i3 : nullptr, 3, dm2, 0u);
        put(p.dL_dcolors ? p.dL_dcolors + i3 : nullptr, 3, dcol, VEGS_ACC_COLORS);
        put(p.dL_dopacity ? p.dL_dopacity + idx : nullptr, 1, &dopac, VEGS_ACC_OPACITY);
        put(p.dL_dscales ? p.dL_dscales + i3 : nullptr, 3, dscale, VEGS_ACC_SCALES);
        put(p.dL_drots ? p.dL_drots + 4 * (size_t)idx : nullptr, 4, drot, VEGS_ACC_ROTATIONS);
        put(p.dL_dcov3D ? p.dL_dcov3D + 6 * (size_t)idx : nullptr, 6, dcov, VEGS_ACC_COV3D);
        // densification statistics of the reference's training loop (scene/gaussian_model.py:411-413,
        // train.py:300-301), for the Gaussians this view rendered
        if (p.stat_grad != nullptr && visible) {
            p.stat_grad[idx] += sqrtf(dm2d[0] * dm2d[0] + dm2d[1] * dm2d[1]);
            p.stat_denom[idx] += 1.0f;
            p.stat_maxr[idx] = fmaxf(p.stat_maxr[idx], (float)in_radius);
        }
    }
}

}  // namespace

int launch_preprocess_bwd(const VegsBackwardArgs& a, const GeomView& g, const float4* acc,
                          cudaStream_t stream) {
    PbParams p;
    p.P = a.P; p.D = a.D; p.M = a.M; p.W = a.W; p.H = a.H;
    p.means = a.means3D; p.radii = a.radii; p.shs = a.shs; p.clamped = g.clamped;
    p.scales = a.scales; p.rots = a.rotations; p.cov3D_precomp = a.cov3D_precomp;
    p.mod = a.scale_modifier; p.tan_fovx = a.tan_fovx; p.tan_fovy = a.tan_fovy;
    p.focal_y = a.H / (2.0f * a.tan_fovy);
    p.focal_x = a.W / (2.0f * a.tan_fovx);
    p.view = a.viewmatrix; p.proj = a.projmatrix; p.campos = a.campos;
    p.acc = acc;
    p.dL_dmeans3D = a.dL_dmeans3D; p.dL_dmeans2D = a.dL_dmeans2D; p.dL_dsh = a.dL_dsh;
    p.dL_dcolors = a.dL_dcolors; p.dL_dopacity = a.dL_dopacity; p.dL_dscales = a.dL_dscales;
    p.dL_drots = a.dL_drotations; p.dL_dcov3D = a.dL_dcov3D;
    p.shs_vec_ok = a.shs != nullptr && (reinterpret_cast<uintptr_t>(a.shs) & 15u) == 0;
    p.dsh_vec_ok = a.dL_dsh != nullptr && (reinterpret_cast<uintptr_t>(a.dL_dsh) & 15u) == 0;
    p.rots_vec_ok = a.rotations != nullptr && (reinterpret_cast<uintptr_t>(a.rotations) & 15u) == 0;
    p.accum = a.accumulate_mask;
    p.stat_grad = a.stat_grad_accum; p.stat_denom = a.stat_denom; p.stat_maxr = a.stat_max_radii;
    if ((p.stat_grad != nullptr) != (p.stat_denom != nullptr) || (p.stat_grad != nullptr) != (p.stat_maxr != nullptr)) {
        set_error("vegs_backward: stat_grad_accum, stat_denom and stat_max_radii go together");
        return VEGS_EINVAL;
    }
    if ((p.accum & VEGS_ACC_SH) && a.M > kMaxStageCoeffs) {
        set_error("vegs_backward: VEGS_ACC_SH needs M <= %d", kMaxStageCoeffs);
        return VEGS_EINVAL;
    }
    const int blocks = (a.P + kPbThreads - 1) / kPbThreads;
    switch (a.M) {
        case 16: preprocess_bwd_kernel<16><<<blocks, kPbThreads, 0, stream>>>(p); break;
        case 9:  preprocess_bwd_kernel<9><<<blocks, kPbThreads, 0, stream>>>(p); break;
        case 4:  preprocess_bwd_kernel<4><<<blocks, kPbThreads, 0, stream>>>(p); break;
        case 1:  preprocess_bwd_kernel<1><<<blocks, kPbThreads, 0, stream>>>(p); break;
        default: preprocess_bwd_kernel<0><<<blocks, kPbThreads, 0, stream>>>(p); break;
    }
    VEGS_LAUNCH_CHECK("preprocess_bwd", stream, a.debug);
    return VEGS_OK;
}

}  // namespace vegs
