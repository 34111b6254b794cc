// Helpers shared by the forward and backward compositing kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace vegs {

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Which of the tile's eight 8x4 pixel blocks (bit = by*2 + bx, block origin (X0 + 8bx, Y0 + 4by))
// can contain a pixel where this Gaussian reaches alpha >= 1/255.
//
// a = {x, y, conic.a, conic.b}, b = {conic.c, opacity, tau, depth}.  alpha >= 1/255 needs
// power >= -tau, i.e. A dx^2 + 2 B dx dy + C dy^2 <= 2 tau, an ellipse whose half extents are
// sqrt(2 tau C / det) in x and sqrt(2 tau A / det) in y (det = AC - B^2).  tau already carries a
// margin; the extents get another one, so the box is conservative and skipping the other blocks
// cannot change a result.  Whenever the box cannot be trusted (non-finite inputs, for which the
// reference's fminf(0.99, NaN) = 0.99 makes EVERY pixel contribute, or a conic that is not
// positive definite) all eight blocks are kept.
__device__ __forceinline__ uint32_t block_mask(const float4 a, const float4 b, int X0, int Y0) {
    const float tau = b.z;
    const bool finite = isfinite(a.x) && isfinite(a.y) && isfinite(a.z) && isfinite(a.w) &&
                        isfinite(b.x) && isfinite(tau);
    if (!finite) return 0xffu;
    if (tau < 0.0f) return 0u;
    const float det = a.z * b.x - a.w * a.w;
    if (!(det > 0.0f)) return 0xffu;
    const float s2 = 2.0f * tau / det;
    const float ex = sqrtf(s2 * b.x) * 1.0001f + 0.01f;
    const float ey = sqrtf(s2 * a.z) * 1.0001f + 0.01f;
    if (!isfinite(ex) || !isfinite(ey)) return 0xffu;
    const float x_lo = a.x - ex, x_hi = a.x + ex, y_lo = a.y - ey, y_hi = a.y + ey;
    uint32_t mx = 0, m = 0;
#pragma unroll
    for (int bx = 0; bx < 2; bx++) {
        const float lo = (float)(X0 + 8 * bx), hi = lo + 7.0f;
        if (x_hi >= lo && x_lo <= hi) mx |= 1u << bx;
    }
#pragma unroll
    for (int by = 0; by < 4; by++) {
        const float lo = (float)(Y0 + 4 * by), hi = lo + 3.0f;
        if (y_hi >= lo && y_lo <= hi) m |= mx << (2 * by);
    }
    return m;
}

}  // namespace vegs
