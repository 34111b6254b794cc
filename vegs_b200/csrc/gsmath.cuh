// Per-Gaussian projection arithmetic, stated once for the forward and backward P-wide kernels.
//
// The values produced here feed integer results that must equal the reference's bit for bit
// (radii, tile rectangles, depth sort keys), so the float operation ORDER is pinned with explicit
// round-to-nearest intrinsics (__fmaf_rn / __fmul_rn / __fadd_rn are never re-associated or
// contracted by nvcc).  The order is the one nvcc 12.9 emits for the reference's expressions
// (SASS of the sm_100a build of cuda_rasterizer/forward.cu):
//     a*b + c*d + e*f        ->  fma(e, f, fma(a, b, c*d))
//     a*b + c*d + e*f + g    ->  fma(e, f, fma(a, b, c*d)) + g
//     a*b - c*c              ->  fma(a, b, -(c*c))
// with shared sub-products (x*z, r*x, r*z, y*y, z*z in the rotation matrix) kept as plain
// products.  oracle/vegs_oracle.c states the same sequence with fmaf().
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "common.cuh"

namespace vegs {

// Camera matrices are read from global memory into shared memory once per block.
__device__ __forceinline__ void load_camera(Camera& cam, const float* view, const float* proj,
                                            const float* campos) {
    const int t = threadIdx.x;
    if (t < 16) {
        cam.view[t] = view[t];
        if (proj != nullptr) cam.proj[t] = proj[t];
    }
    if (t < 3 && campos != nullptr) cam.campos[t] = campos[t];
    __syncthreads();
}

__device__ __forceinline__ float4 load_quat(const float* rots, int idx, bool vec_ok) {
    if (vec_ok) return reinterpret_cast<const float4*>(rots)[idx];
    return make_float4(rots[4 * (size_t)idx + 0], rots[4 * (size_t)idx + 1],
                       rots[4 * (size_t)idx + 2], rots[4 * (size_t)idx + 3]);
}


// m[a]*x + m[b]*y + m[c]*z + m[d]   (reference: transformPoint4x3/4x4, auxiliary.h:58-77)
__device__ __forceinline__ float affine_row(const float* __restrict__ m, int a, int b, int c,
                                            int d, float x, float y, float z) {
    return __fadd_rn(__fmaf_rn(m[c], z, __fmaf_rn(m[a], x, __fmul_rn(m[b], y))), m[d]);
}

// a0*b0 + a1*b1 + a2*b2  in the reference's contraction order
__device__ __forceinline__ float dot3_ref(float a0, float b0, float a1, float b1, float a2,
                                          float b2) {
    return __fmaf_rn(a2, b2, __fmaf_rn(a0, b0, __fmul_rn(a1, b1)));
}

// 3D covariance from scale and (un-normalised) quaternion (r,x,y,z).
// Reference: computeCov3D, cuda_rasterizer/forward.cu:119-153.
__device__ __forceinline__ void cov3d_from_scale_rot(float sx, float sy, float sz, float mod,
                                                     float r, float x, float y, float z,
                                                     float* __restrict__ cov) {
    const float s0 = __fmul_rn(mod, sx), s1 = __fmul_rn(mod, sy), s2 = __fmul_rn(mod, sz);
    const float xz = __fmul_rn(x, z), rx = __fmul_rn(r, x), rz = __fmul_rn(r, z);
    const float yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
    // rotation matrix entries R[col][row] (glm column-major constructor)
    const float R00 = __fsub_rn(1.0f, __fmul_rn(2.0f, __fadd_rn(yy, zz)));
    const float R01 = __fmul_rn(2.0f, __fmaf_rn(x, y, -rz));
    const float R02 = __fmul_rn(2.0f, __fmaf_rn(r, y, xz));
    const float R10 = __fmul_rn(2.0f, __fmaf_rn(x, y, rz));
    const float R11 = __fsub_rn(1.0f, __fmul_rn(2.0f, __fmaf_rn(x, x, zz)));
    const float R12 = __fmul_rn(2.0f, __fmaf_rn(y, z, -rx));
    const float R20 = __fmul_rn(2.0f, __fmaf_rn(-r, y, xz));
    const float R21 = __fmul_rn(2.0f, __fmaf_rn(y, z, rx));
    const float R22 = __fsub_rn(1.0f, __fmul_rn(2.0f, __fmaf_rn(x, x, yy)));
    // M = S * R : M[col][row] = s_row * R[col][row]
    const float M00 = __fmul_rn(s0, R00), M01 = __fmul_rn(s1, R01), M02 = __fmul_rn(s2, R02);
    const float M10 = __fmul_rn(s0, R10), M11 = __fmul_rn(s1, R11), M12 = __fmul_rn(s2, R12);
    const float M20 = __fmul_rn(s0, R20), M21 = __fmul_rn(s1, R21), M22 = __fmul_rn(s2, R22);
    // Sigma = M^T M : Sigma[c][r] = sum_k M[r][k] * M[c][k]
    cov[0] = dot3_ref(M00, M00, M01, M01, M02, M02);
    cov[1] = dot3_ref(M00, M10, M01, M11, M02, M12);
    cov[2] = dot3_ref(M00, M20, M01, M21, M02, M22);
    cov[3] = dot3_ref(M10, M10, M11, M11, M12, M12);
    cov[4] = dot3_ref(M10, M20, M11, M21, M12, M22);
    cov[5] = dot3_ref(M20, M20, M21, M21, M22, M22);
}

struct Cov2DTerms {
    float T00, T01, T02;  // T[0][r]
    float T10, T11, T12;  // T[1][r]
    float tx, ty, tz;     // clamped view-space mean
    float txtz, tytz;     // unclamped ratios (for the backward's clamp masks)
};

// EWA projection of the 3D covariance (reference: computeCov2D, forward.cu:75-114).  Returns
// (cov[0][0]+0.3, cov[0][1], cov[1][1]+0.3).
__device__ __forceinline__ float3 cov2d_project(float mx, float my, float mz, float focal_x,
                                                float focal_y, float tan_fovx, float tan_fovy,
                                                const float* __restrict__ c3,
                                                const float* __restrict__ view, Cov2DTerms& o) {
    float tx = affine_row(view, 0, 4, 8, 12, mx, my, mz);
    float ty = affine_row(view, 1, 5, 9, 13, mx, my, mz);
    const float tz = affine_row(view, 2, 6, 10, 14, mx, my, mz);
    const float limx = __fmul_rn(1.3f, tan_fovx);
    const float limy = __fmul_rn(1.3f, tan_fovy);
    const float txtz = __fdiv_rn(tx, tz);
    const float tytz = __fdiv_rn(ty, tz);
    tx = __fmul_rn(fminf(limx, fmaxf(-limx, txtz)), tz);
    ty = __fmul_rn(fminf(limy, fmaxf(-limy, tytz)), tz);
    const float tz2 = __fmul_rn(tz, tz);
    const float J00 = __fdiv_rn(focal_x, tz);
    const float J02 = __fdiv_rn(-__fmul_rn(focal_x, tx), tz2);
    const float J11 = __fdiv_rn(focal_y, tz);
    const float J12 = __fdiv_rn(-__fmul_rn(focal_y, ty), tz2);
    // T = W * J with W[k][r] = view[4*r + k];  T[0][r] = W[0][r]*J00 + W[2][r]*J02,
    // T[1][r] = W[1][r]*J11 + W[2][r]*J12   (the products with J's zeros add exact zeros)
    o.T00 = __fmaf_rn(view[2], J02, __fmul_rn(view[0], J00));
    o.T01 = __fmaf_rn(view[6], J02, __fmul_rn(view[4], J00));
    o.T02 = __fmaf_rn(view[10], J02, __fmul_rn(view[8], J00));
    o.T10 = __fmaf_rn(view[2], J12, __fmul_rn(view[1], J11));
    o.T11 = __fmaf_rn(view[6], J12, __fmul_rn(view[5], J11));
    o.T12 = __fmaf_rn(view[10], J12, __fmul_rn(view[9], J11));
    o.tx = tx; o.ty = ty; o.tz = tz; o.txtz = txtz; o.tytz = tytz;
    // A = T^T * Vrk^T : A[c][r] = T[r][0]*V[0][c] + T[r][1]*V[1][c] + T[r][2]*V[2][c]
    // V[0] = (c0,c1,c2), V[1] = (c1,c3,c4), V[2] = (c2,c4,c5)
    const float A00 = dot3_ref(o.T00, c3[0], o.T01, c3[1], o.T02, c3[2]);
    const float A10 = dot3_ref(o.T00, c3[1], o.T01, c3[3], o.T02, c3[4]);
    const float A20 = dot3_ref(o.T00, c3[2], o.T01, c3[4], o.T02, c3[5]);
    const float A01 = dot3_ref(o.T10, c3[0], o.T11, c3[1], o.T12, c3[2]);
    const float A11 = dot3_ref(o.T10, c3[1], o.T11, c3[3], o.T12, c3[4]);
    const float A21 = dot3_ref(o.T10, c3[2], o.T11, c3[4], o.T12, c3[5]);
    // cov[c][r] = A[0][r]*T[c][0] + A[1][r]*T[c][1] + A[2][r]*T[c][2]
    float3 cov;
    cov.x = __fadd_rn(dot3_ref(A00, o.T00, A10, o.T01, A20, o.T02), 0.3f);
    cov.y = dot3_ref(A01, o.T00, A11, o.T01, A21, o.T02);
    cov.z = __fadd_rn(dot3_ref(A01, o.T10, A11, o.T11, A21, o.T12), 0.3f);
    return cov;
}

// ndc2Pix in FP64, as the reference's double literals force (auxiliary.h:41-44; SASS:
// DADD, DFMA, DMUL, F2F).
__device__ __forceinline__ float ndc2pix_ref(float v, int S) {
    return (float)(__dmul_rn(__fma_rn(__dadd_rn((double)v, 1.0), (double)S, -1.0), 0.5));
}

// Tile rectangle (auxiliary.h:46-56): float division by 16 (== * 0.0625f), C truncation,
// clamp to [0, grid].  Note (p + r + 16.0f) - 1.0f: the reference adds BLOCK_X then subtracts 1.
__device__ __forceinline__ void tile_rect(float px, float py, int radius, int gx, int gy,
                                          int& x0, int& y0, int& x1, int& y1) {
    const float r = (float)radius;
    x0 = min(gx, max(0, (int)__fmul_rn(__fsub_rn(px, r), 0.0625f)));
    y0 = min(gy, max(0, (int)__fmul_rn(__fsub_rn(py, r), 0.0625f)));
    x1 = min(gx, max(0, (int)__fmul_rn(__fsub_rn(__fadd_rn(__fadd_rn(px, r), 16.0f), 1.0f), 0.0625f)));
    y1 = min(gy, max(0, (int)__fmul_rn(__fsub_rn(__fadd_rn(__fadd_rn(py, r), 16.0f), 1.0f), 0.0625f)));
}

// SH basis constants (auxiliary.h:21-39)
__device__ constexpr float kSH_C0 = 0.28209479177387814f;
__device__ constexpr float kSH_C1 = 0.4886025119029199f;
__device__ constexpr float kSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f,
                                        0.31539156525252005f, -1.0925484305920792f,
                                        0.5462742152960396f};
__device__ constexpr float kSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f,
                                        -0.4570457994644658f, 0.3731763325901154f,
                                        -0.4570457994644658f, 1.445305721320277f,
                                        -0.5900435899266435f};

}  // namespace vegs
