// Backward of the SH -> RGB evaluation for one Gaussian, in a streaming, per-coefficient form.
//
// Replaces computeColorFromSH's backward (reference cuda_rasterizer/backward.cu:20-139): the same
// polynomials, stated coefficient by coefficient so that a row can be processed in groups of four
// coefficients (twelve floats = three 16-byte shared-memory accesses) with a dozen live registers
// instead of the whole 48-float row:
//     colour_c = 0.5 + sum_k w_k(dir) * sh[k][c]
//     dL/dsh[k][c]  = w_k(dir) * dL/dcolour_c
//     dL/ddir      += (dw_k/ddir)(dir) * sum_c sh[k][c] * dL/dcolour_c
// and dL/dmean follows from dL/ddir through the normalisation (dnormvdv, auxiliary.h:107-117).
// The sums run in coefficient order; the reference adds the same terms band by band, a last-bit
// difference far inside the 1e-4 gradient tolerance.
#pragma once
#include <cuda_runtime.h>
#include "gsmath.cuh"

namespace vegs {

// Basis weight of coefficient K at unit direction (x, y, z) and its gradient.
template <int K>
__device__ __forceinline__ void sh_basis(float x, float y, float z, float& w, float& dx, float& dy, float& dz) {
    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
    dx = dy = dz = 0.f;
    if (K == 0) { w = kSH_C0; }
    else if (K == 1) { w = -kSH_C1 * y; dy = -kSH_C1; }
    else if (K == 2) { w = kSH_C1 * z; dz = kSH_C1; }
    else if (K == 3) { w = -kSH_C1 * x; dx = -kSH_C1; }
    else if (K == 4) { w = kSH_C2[0] * xy; dx = kSH_C2[0] * y; dy = kSH_C2[0] * x; }
    else if (K == 5) { w = kSH_C2[1] * yz; dy = kSH_C2[1] * z; dz = kSH_C2[1] * y; }
    else if (K == 6) { w = kSH_C2[2] * (2.f * zz - xx - yy); dx = kSH_C2[2] * 2.f * -x; dy = kSH_C2[2] * 2.f * -y; dz = kSH_C2[2] * 2.f * 2.f * z; }
    else if (K == 7) { w = kSH_C2[3] * xz; dx = kSH_C2[3] * z; dz = kSH_C2[3] * x; }
    else if (K == 8) { w = kSH_C2[4] * (xx - yy); dx = kSH_C2[4] * 2.f * x; dy = kSH_C2[4] * 2.f * -y; }
    else if (K == 9) { w = kSH_C3[0] * y * (3.f * xx - yy); dx = kSH_C3[0] * 3.f * 2.f * xy; dy = kSH_C3[0] * 3.f * (xx - yy); }
    else if (K == 10) { w = kSH_C3[1] * xy * z; dx = kSH_C3[1] * yz; dy = kSH_C3[1] * xz; dz = kSH_C3[1] * xy; }
    else if (K == 11) { w = kSH_C3[2] * y * (4.f * zz - xx - yy); dx = kSH_C3[2] * -2.f * xy; dy = kSH_C3[2] * (-3.f * yy + 4.f * zz - xx); dz = kSH_C3[2] * 4.f * 2.f * yz; }
    else if (K == 12) { w = kSH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy); dx = kSH_C3[3] * -3.f * 2.f * xz; dy = kSH_C3[3] * -3.f * 2.f * yz; dz = kSH_C3[3] * 3.f * (2.f * zz - xx - yy); }
    else if (K == 13) { w = kSH_C3[4] * x * (4.f * zz - xx - yy); dx = kSH_C3[4] * (-3.f * xx + 4.f * zz - yy); dy = kSH_C3[4] * -2.f * xy; dz = kSH_C3[4] * 4.f * 2.f * xz; }
    else if (K == 14) { w = kSH_C3[5] * z * (xx - yy); dx = kSH_C3[5] * 2.f * xz; dy = kSH_C3[5] * -2.f * yz; dz = kSH_C3[5] * (xx - yy); }
    else { w = kSH_C3[6] * x * (xx - 3.f * yy); dx = kSH_C3[6] * 3.f * (xx - yy); dy = kSH_C3[6] * -3.f * 2.f * xy; }
}

struct ShBwdState {
    float x, y, z;          // unit view direction
    float dRGB[3];          // dL/dcolour with clamped channels zeroed
    float gx, gy, gz;       // dL/d(unit direction), accumulated
    int used;               // (deg + 1)^2 coefficients take part
};

__device__ __forceinline__ ShBwdState sh_bwd_begin(int deg, float dox, float doy, float doz, const float dRGB[3]) {
    ShBwdState s;
    const float len = sqrtf(dox * dox + doy * doy + doz * doz);
    s.x = dox / len; s.y = doy / len; s.z = doz / len;
    s.dRGB[0] = dRGB[0]; s.dRGB[1] = dRGB[1]; s.dRGB[2] = dRGB[2];
    s.gx = s.gy = s.gz = 0.f;
    s.used = (deg + 1) * (deg + 1);
    return s;
}

// One coefficient: v[3] holds sh[K][rgb] on entry and dL/dsh[K][rgb] on exit.
template <int K>
__device__ __forceinline__ void sh_bwd_coeff(ShBwdState& s, float v[3]) {
    if (K < s.used) {
        float w, dx, dy, dz;
        sh_basis<K>(s.x, s.y, s.z, w, dx, dy, dz);
        if (K > 0) {
            const float t = v[0] * s.dRGB[0] + v[1] * s.dRGB[1] + v[2] * s.dRGB[2];
            s.gx += dx * t; s.gy += dy * t; s.gz += dz * t;
        }
        v[0] = w * s.dRGB[0]; v[1] = w * s.dRGB[1]; v[2] = w * s.dRGB[2];
    } else {
        v[0] = v[1] = v[2] = 0.f;     // coefficients above the active degree get a zero gradient
    }
}

// Four consecutive coefficients 4G .. 4G+3 = twelve floats g[0..11] in memory order [k][rgb].
template <int G>
__device__ __forceinline__ void sh_bwd_group(ShBwdState& s, float g[12]) {
    sh_bwd_coeff<4 * G + 0>(s, g + 0);
    sh_bwd_coeff<4 * G + 1>(s, g + 3);
    sh_bwd_coeff<4 * G + 2>(s, g + 6);
    sh_bwd_coeff<4 * G + 3>(s, g + 9);
}

// dL/dmean from dL/d(unit direction): dnormvdv (auxiliary.h:107-117)
__device__ __forceinline__ float3 sh_bwd_end(const ShBwdState& s, float dox, float doy, float doz) {
    const float sum2 = dox * dox + doy * doy + doz * doz;
    const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    float3 r;
    r.x = ((+sum2 - dox * dox) * s.gx - doy * dox * s.gy - doz * dox * s.gz) * invsum32;
    r.y = (-dox * doy * s.gx + (sum2 - doy * doy) * s.gy - doz * doy * s.gz) * invsum32;
    r.z = (-dox * doz * s.gx - doy * doz * s.gy + (sum2 - doz * doz) * s.gz) * invsum32;
    return r;
}

}  // namespace vegs
