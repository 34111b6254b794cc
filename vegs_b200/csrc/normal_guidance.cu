// Normal-guidance loss on the rasterizer's covariance maps, value and gradients in one pass.
//
// No counterpart inside the reference's extensions: it replaces the reference's Python loss
// (loss/normal_guidance.py:3-23, with utils/graphics_utils.py:204-246 quaternion_to_matrix and
// :362-368 cam_normal_to_world_normal), which runs a dozen element-wise / permute / repeat torch
// kernels over [Npix,3,3] temporaries per iteration.  Per pixel:
//     n_w   = R_cam2world * n_cam                       the predicted normal in world space
//     Rot   = matrix of the (un-normalised) map quaternion, two_s = 2 / |q|^2
//     t[b]  = sum_a Rot[a][b] * n_w[a]                  (b = 0, 1, 2: the three covariance axes)
//     loss  = w_rot * mean |t[b]|  +  w_scale * mean |scale[b] * t[b]|        (means over 3 Npix)
// with, as in the reference, the rotation DETACHED in the second term: the quaternion map receives
// the gradient of the first term only, the scale map that of the second.
// One thread per pixel; the loss is reduced per block and added to one device word.
// Roofline: HBM, 10 floats read and 7 written per pixel.
#include "common.cuh"

namespace vegs {

namespace {

struct NgParams {
    int npix;
    const float* quat;      // [4, Npix] planes (w, x, y, z)
    const float* scale;     // [3, Npix]
    const float* normal;    // [3, Npix] camera-space normals
    float R[9];             // cam -> world, row-major
    float c_rot, c_scale;   // w_rot / (3 Npix), w_scale / (3 Npix)
    float* loss;            // [1], zeroed before the launch
    float* d_quat;          // [4, Npix] or NULL
    float* d_scale;         // [3, Npix] or NULL
};

__global__ void __launch_bounds__(256) normal_guidance_kernel(const NgParams p) {
    __shared__ float s_part[8];
    const int pix = blockIdx.x * 256 + threadIdx.x;
    float local = 0.f;
    if (pix < p.npix) {
        const float r = p.quat[pix], i = p.quat[p.npix + pix], j = p.quat[2 * p.npix + pix], k = p.quat[3 * p.npix + pix];
        const float nc0 = p.normal[pix], nc1 = p.normal[p.npix + pix], nc2 = p.normal[2 * p.npix + pix];
        const float n[3] = {p.R[0] * nc0 + p.R[1] * nc1 + p.R[2] * nc2, p.R[3] * nc0 + p.R[4] * nc1 + p.R[5] * nc2,
                            p.R[6] * nc0 + p.R[7] * nc1 + p.R[8] * nc2};
        const float s = r * r + i * i + j * j + k * k;
        const float ts = 2.0f / s;
        // Rot = I + ts * B(q), B bilinear in q (utils/graphics_utils.py:233-246)
        const float B[3][3] = {{-(j * j + k * k), i * j - k * r, i * k + j * r},
                               {i * j + k * r, -(i * i + k * k), j * k - i * r},
                               {i * k - j * r, j * k + i * r, -(i * i + j * j)}};
        float t[3], g[3];     // t[b], and dL/dt[b] of the first term
        const float sc[3] = {p.scale[pix], p.scale[p.npix + pix], p.scale[2 * p.npix + pix]};
#pragma unroll
        for (int b = 0; b < 3; b++) {
            t[b] = n[b] + ts * (B[0][b] * n[0] + B[1][b] * n[1] + B[2][b] * n[2]);
            local += p.c_rot * fabsf(t[b]) + p.c_scale * fabsf(sc[b] * t[b]);
            g[b] = p.c_rot * (t[b] > 0.f ? 1.f : (t[b] < 0.f ? -1.f : 0.f));
        }
        if (p.d_scale != nullptr) {
#pragma unroll
            for (int b = 0; b < 3; b++) {
                const float v = sc[b] * t[b];
                p.d_scale[b * p.npix + pix] = p.c_scale * (v > 0.f ? 1.f : (v < 0.f ? -1.f : 0.f)) * t[b];
            }
        }
        if (p.d_quat != nullptr) {
            // G[a][b] = dL/dRot[a][b] = g[b] * n[a];  dL/dq_m = ts * <G, dB/dq_m> - (2 q_m / s) * ts * <G, B>
            float GB = 0.f;
#pragma unroll
            for (int a = 0; a < 3; a++)
#pragma unroll
                for (int b = 0; b < 3; b++) GB += g[b] * n[a] * B[a][b];
            auto G = [&](int a, int b) { return g[b] * n[a]; };
            // dB/dr, dB/di, dB/dj, dB/dk contracted with G
            const float dr = -k * G(0, 1) + j * G(0, 2) + k * G(1, 0) - i * G(1, 2) - j * G(2, 0) + i * G(2, 1);
            const float di = j * G(0, 1) + k * G(0, 2) + j * G(1, 0) - 2.f * i * G(1, 1) - r * G(1, 2) + k * G(2, 0) +
                             r * G(2, 1) - 2.f * i * G(2, 2);
            const float dj = -2.f * j * G(0, 0) + i * G(0, 1) + r * G(0, 2) + i * G(1, 0) + k * G(1, 2) - r * G(2, 0) +
                             k * G(2, 1) - 2.f * j * G(2, 2);
            const float dk = -2.f * k * G(0, 0) - r * G(0, 1) + i * G(0, 2) + r * G(1, 0) - 2.f * k * G(1, 1) + j * G(1, 2) +
                             i * G(2, 0) + j * G(2, 1);
            const float corr = 2.0f / s * ts * GB;
            p.d_quat[pix] = ts * dr - corr * r;
            p.d_quat[p.npix + pix] = ts * di - corr * i;
            p.d_quat[2 * p.npix + pix] = ts * dj - corr * j;
            p.d_quat[3 * p.npix + pix] = ts * dk - corr * k;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = local;
    __syncthreads();
    if (threadIdx.x == 0) {
        float tot = 0.f;
#pragma unroll
        for (int w = 0; w < 8; w++) tot += s_part[w];
        atomicAdd(p.loss, tot);
    }
}

}  // namespace

}  // namespace vegs

extern "C" int vegs_normal_guidance(int32_t W, int32_t H, const float* cov_quat, const float* cov_scale,
                                    const float* normal_cam, const float* R_cam2world_host, float w_rot, float w_scale,
                                    float* loss, float* d_cov_quat, float* d_cov_scale, void* stream_) {
    using namespace vegs;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    if (W <= 0 || H <= 0 || !cov_quat || !cov_scale || !normal_cam || !R_cam2world_host || !loss) {
        set_error("vegs_normal_guidance: bad argument");
        return VEGS_EINVAL;
    }
    NgParams p;
    p.npix = W * H;
    p.quat = cov_quat; p.scale = cov_scale; p.normal = normal_cam;
    for (int k = 0; k < 9; k++) p.R[k] = R_cam2world_host[k];
    p.c_rot = w_rot / (3.0f * (float)p.npix);
    p.c_scale = w_scale / (3.0f * (float)p.npix);
    p.loss = loss; p.d_quat = d_cov_quat; p.d_scale = d_cov_scale;
    VEGS_CUDA_OK(cudaMemsetAsync(loss, 0, sizeof(float), stream));
    normal_guidance_kernel<<<(p.npix + 255) / 256, 256, 0, stream>>>(p);
    VEGS_LAUNCH_CHECK("normal_guidance", stream, false);
    return VEGS_OK;
}
