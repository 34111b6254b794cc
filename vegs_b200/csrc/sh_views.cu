// Rebuilds the SH gradient of a step from the per-view colour gradients.
//
// No counterpart in the reference (it is single-GPU and writes dL/dSH per view).  The colour of a
// Gaussian is linear in its SH coefficients, colour_c = 0.5 + sum_k basis_k(dir) * sh[k][c]
// (reference cuda_rasterizer/forward.cu:21-72), so the gradient one view contributes is the outer
// product basis(dir_view) x dL/dRGB_view (clamped channels zeroed), and the gradient of a step of
// V views is
//     dL/dsh[k][c] = sum_v basis_k(normalize(mean - campos_v)) * dL/dRGB_v[c] .
// The backward kernel can therefore hand out 12 bytes per view and Gaussian
// (VegsBackwardArgs.dL_drgb_deferred) instead of accumulating 12 M bytes, and a view-parallel step
// exchanges those 3 words per view between the GPUs (an all-gather of V x P x 3 floats) instead of
// all-reducing the 3 M words of the SH gradient: 24 instead of 48 words per Gaussian on the wire
// for an 8-view step at M = 16, and no read-modify-write of the SH gradient per view.
// One thread per Gaussian, a warp per 32 consecutive Gaussians; the rows leave through shared
// memory as bulk stores like the backward kernel's (bulk.cuh).
// Roofline: HBM; 12 V + 12 bytes read and 12 M written per Gaussian.
#include "common.cuh"
#include "gsmath.cuh"
#include "bulk.cuh"
#include "pwide.cuh"
#include "sh_bwd.cuh"

namespace vegs {

namespace {

struct ShViewsParams {
    int P, D, M, V;
    const float* means;
    const float* campos;     // [V,3] device
    const float* drgb;       // [V,P,3]
    float* dsh;              // [P,M,3], or the f_dc part [P,1,3] when dsh_rest is given
    float* dsh_rest;         // [P,M-1,3] or NULL
    int accumulate;          // add to the destination instead of overwriting it
    bool vec_ok;             // destinations 16-byte aligned
};

template <int K>
__device__ __forceinline__ void add_coeff(float acc[48], int used, float x, float y, float z, const float g[3]) {
    if (K < used) {
        float w, dx, dy, dz;
        sh_basis<K>(x, y, z, w, dx, dy, dz);
        acc[3 * K + 0] = __fmaf_rn(w, g[0], acc[3 * K + 0]);
        acc[3 * K + 1] = __fmaf_rn(w, g[1], acc[3 * K + 1]);
        acc[3 * K + 2] = __fmaf_rn(w, g[2], acc[3 * K + 2]);
    }
}

__global__ void __launch_bounds__(kPwThreads) sh_grad_from_views_kernel(const ShViewsParams p) {
    __shared__ __align__(16) float s_rows[kPwWarps][kShWarpFloats];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int idx = blockIdx.x * kPwThreads + threadIdx.x;
    const int warp_base = idx - lane;
    const bool in_range = idx < p.P;
    const bool full = warp_base + 32 <= p.P;
    const int M = p.M;
    const int used = (p.D + 1) * (p.D + 1);
    float acc[48];
#pragma unroll
    for (int k = 0; k < 48; k++) acc[k] = 0.f;
    if (in_range) {
        const float mx = p.means[3 * (size_t)idx], my = p.means[3 * (size_t)idx + 1], mz = p.means[3 * (size_t)idx + 2];
        // four views at a time: their twelve gradient words are requested before any of them is
        // used, so that the loop waits for memory once per four views instead of once per view
        for (int v0 = 0; v0 < p.V; v0 += 4) {
            float g[4][3];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const bool on = v0 + u < p.V;
                const float* g3 = p.drgb + ((size_t)(on ? v0 + u : v0) * p.P + idx) * 3;
                g[u][0] = on ? __ldg(g3) : 0.f;
                g[u][1] = on ? __ldg(g3 + 1) : 0.f;
                g[u][2] = on ? __ldg(g3 + 2) : 0.f;
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                if (g[u][0] == 0.f && g[u][1] == 0.f && g[u][2] == 0.f) continue;     // not rendered by this view
                const int v = v0 + u;
                const float dox = mx - p.campos[3 * v], doy = my - p.campos[3 * v + 1], doz = mz - p.campos[3 * v + 2];
                const float len = sqrtf(dox * dox + doy * doy + doz * doz);   // as sh_bwd_begin
                const float x = dox / len, y = doy / len, z = doz / len;
                add_coeff<0>(acc, used, x, y, z, g[u]); add_coeff<1>(acc, used, x, y, z, g[u]); add_coeff<2>(acc, used, x, y, z, g[u]);
                add_coeff<3>(acc, used, x, y, z, g[u]); add_coeff<4>(acc, used, x, y, z, g[u]); add_coeff<5>(acc, used, x, y, z, g[u]);
                add_coeff<6>(acc, used, x, y, z, g[u]); add_coeff<7>(acc, used, x, y, z, g[u]); add_coeff<8>(acc, used, x, y, z, g[u]);
                add_coeff<9>(acc, used, x, y, z, g[u]); add_coeff<10>(acc, used, x, y, z, g[u]); add_coeff<11>(acc, used, x, y, z, g[u]);
                add_coeff<12>(acc, used, x, y, z, g[u]); add_coeff<13>(acc, used, x, y, z, g[u]); add_coeff<14>(acc, used, x, y, z, g[u]);
                add_coeff<15>(acc, used, x, y, z, g[u]);
            }
        }
    }
    const int rowlen = 3 * M;
    if (p.dsh_rest == nullptr && M == 16 && p.vec_ok) {
        // 192-byte rows: 16-byte shared stores into 208-byte-stride rows, one bulk store / reduction per row
        float4* r4 = reinterpret_cast<float4*>(s_rows[warp] + lane * kShRowStride);
#pragma unroll
        for (int j = 0; j < 12; j++) r4[j] = make_float4(acc[4 * j], acc[4 * j + 1], acc[4 * j + 2], acc[4 * j + 3]);
        fence_smem_to_async_proxy();
        if (in_range) {
            float* dst = p.dsh + (size_t)idx * 48;
            if (p.accumulate) bulk_s2g_add_f32(dst, r4, 192u);
            else bulk_s2g(dst, r4, 192u);
        }
    } else if (p.dsh_rest != nullptr && full && p.vec_ok) {
        // split layout: the warp's dense f_dc block (384 B) and f_rest block (32 x 12 (M-1) B) as two bulk stores
        const int rs = 3 * (M - 1);
        float* rest = s_rows[warp] + lane * rs;
        float* dc = s_rows[warp] + 32 * rs + 3 * lane;
        // every index into acc[] is a compile-time constant (guards instead of run-time bounds), so
        // the 48 sums stay in registers
#pragma unroll
        for (int k = 0; k < 3; k++) dc[k] = acc[k];
#pragma unroll
        for (int k = 0; k < 45; k++)
            if (k < rs) rest[k] = acc[3 + k];
        fence_smem_to_async_proxy();
        __syncwarp();
        if (lane == 0) {
            float* d0 = p.dsh + 3 * (size_t)warp_base;
            float* d1 = p.dsh_rest + (size_t)warp_base * rs;
            if (p.accumulate) {
                bulk_s2g_add_f32(d0, s_rows[warp] + 32 * rs, 384u);
                if (rs) bulk_s2g_add_f32(d1, s_rows[warp], 32u * 4u * (uint32_t)rs);
            } else {
                bulk_s2g(d0, s_rows[warp] + 32 * rs, 384u);
                if (rs) bulk_s2g(d1, s_rows[warp], 32u * 4u * (uint32_t)rs);
            }
        }
    } else if (in_range) {
        float* d0 = p.dsh + (size_t)idx * (p.dsh_rest ? 3 : rowlen);
        const int n0 = p.dsh_rest ? 3 : rowlen;
#pragma unroll
        for (int k = 0; k < 48; k++)
            if (k < n0) d0[k] = p.accumulate ? d0[k] + acc[k] : acc[k];
        if (p.dsh_rest) {
            float* d1 = p.dsh_rest + (size_t)idx * 3 * (M - 1);
#pragma unroll
            for (int k = 0; k < 45; k++)
                if (k < 3 * (M - 1)) d1[k] = p.accumulate ? d1[k] + acc[3 + k] : acc[3 + k];
        }
    }
    bulk_commit();
    bulk_wait_read_all();
}

}  // namespace

}  // namespace vegs

extern "C" int vegs_sh_grad_from_views(int32_t P, int32_t D, int32_t M, int32_t V, const float* means3D,
                                       const float* campos, const float* dL_drgb, float* dL_dsh, float* dL_dsh_rest,
                                       int32_t accumulate, void* stream_) {
    using namespace vegs;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    if (P < 0 || V < 0 || D < 0 || D > 3 || M < (D + 1) * (D + 1) || M > kMaxStageCoeffs ||
        (P > 0 && (!means3D || !dL_dsh || (V > 0 && (!campos || !dL_drgb))))) {
        set_error("vegs_sh_grad_from_views: bad argument (0 <= D <= 3, (D+1)^2 <= M <= %d)", kMaxStageCoeffs);
        return VEGS_EINVAL;
    }
    if (P == 0) return VEGS_OK;
    ShViewsParams p;
    p.P = P; p.D = D; p.M = M; p.V = V;
    p.means = means3D; p.campos = campos; p.drgb = dL_drgb; p.dsh = dL_dsh; p.dsh_rest = dL_dsh_rest;
    p.accumulate = accumulate;
    auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };
    p.vec_ok = al16(dL_dsh) && al16(dL_dsh_rest);
    sh_grad_from_views_kernel<<<(P + kPwThreads - 1) / kPwThreads, kPwThreads, 0, stream>>>(p);
    VEGS_LAUNCH_CHECK("sh_grad_from_views", stream, false);
    return VEGS_OK;
}
