// Helpers shared by the forward and backward compositing kernels.
//
// Work decomposition of both kernels: the unit of work is one 8x4 pixel block of a 16x16 tile
// (8 blocks per tile), processed by one 8-warp CTA that walks the tile's depth-sorted instance list;
// units are independent and issued longest-list-first.  The forward CTA first SCANS the raw list
// (does the Gaussian's conservative box, half extents from the preprocess stage, reach the block?
// - one bit of the list word, or a 16-byte record per entry when ids need all 32 bits) and appends
// the survivors, in list order, to a ring of candidates; whenever 256 candidates are queued (or
// the list ends) it PROCESSES them as one working set.  The backward CTA walks the hit list the
// forward left for its unit, 256 entries per working set.  A working set alternates between two
// views of the same problem:
//   entry-parallel   thread t owns candidate t: it evaluates the reference's exact contribution
//                    test on the few pixels inside the clipped box, giving a 32-bit hit mask over
//                    the block's pixels and each hit's alpha;
//   pixel-parallel   after a 32x32 bit transpose through shuffles, lane p owns pixel p and visits
//                    only its own hits, in list order.
// At LiDAR density a Gaussian touches a handful of the tile's 256 pixels, so this replaces
// "every pixel thread evaluates every entry" by "each entry is evaluated only where it can matter",
// and the per-working-set costs (barriers, the sequential pass) are paid per 256 candidates, not
// per 256 list entries.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include "common.cuh"

namespace vegs {

constexpr int kUnitW = 8, kUnitH = 4;          // pixel block of one warp

// 32x32 bit-matrix transpose across the warp: bit t of the result in lane p = bit p of v in lane t.
__device__ __forceinline__ uint32_t transpose32(uint32_t v, int lane) {
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) {
        const uint32_t m = (s == 16) ? 0x0000ffffu : (s == 8) ? 0x00ff00ffu : (s == 4) ? 0x0f0f0f0fu
                         : (s == 2) ? 0x33333333u : 0x55555555u;
        const uint32_t o = __shfl_xor_sync(0xffffffffu, v, s);
        v = (lane & s) ? ((v & ~m) | ((o & ~m) >> s)) : ((v & m) | ((o & m) << s));
    }
    return v;
}

// Column of slot t in row `pix` of a [pixel][slot] shared matrix.  The pixel-parallel phases read
// the matrix with lane = pixel, and neighbouring pixels mostly walk the same entries: without the
// XOR all 32 lanes would hit one bank (row stride is a multiple of 32 words).  Entry-parallel
// accesses (fixed pixel, consecutive slots) stay a permutation of the banks.
__device__ __forceinline__ int swz(int t, int pix) { return t ^ pix; }

// Clip an entry's conservative box to the unit's pixel block; false when nothing is left.
// a = recA = {x, y, half2(ex, ey), depth}.
__device__ __forceinline__ bool unit_box(const float4 a, int bx0, int by0, int bx1, int by1, int& x0,
                                         int& x1, int& y0, int& y1) {
    const uint32_t packed = __float_as_uint(a.z);
    const float2 e = __half22float2(*reinterpret_cast<const __half2*>(&packed));
    x0 = (int)ceilf(fmaxf(a.x - e.x, (float)bx0));
    x1 = (int)floorf(fminf(a.x + e.x, (float)bx1));
    y0 = (int)ceilf(fmaxf(a.y - e.y, (float)by0));
    y1 = (int)floorf(fminf(a.y + e.y, (float)by1));
    return x0 <= x1 && y0 <= y1;
}

// Hit mask of one entry (slot `slot` of the working set) over the block's 32 pixels; for each hit
// (1 - alpha, depth * alpha) is stored at mat[pixel * stride + slot].
// Mask layout: (bit = (y - by0) * 8 + (x - bx0)).
// a = recA = {x, y, half2(ex, ey), depth}, b = recB = {conic.a, conic.b, conic.c, opacity}.
// The test is the reference's (cuda_rasterizer/forward.cu:371-381): power <= 0 and
// min(0.99, opacity * exp(power)) >= 1/255, in its float operation order (power as
// fma(fma(A*dx, dx, (C*dy)*dy), -0.5, -((B*dx)*dy)), read from the reference's SASS).
// Which pixels are put to that test is narrowed conservatively first: to the entry's box, and in
// each pixel row to the interval where the quadratic can reach -(log(255 opacity) + 0.01)
// (alpha >= 1/255 needs power >= -log(255 opacity)); an elongated Gaussian fills a small part of
// its box.  What has to be bounded is the REFERENCE's float evaluation of the power, whose rounding
// error grows with the size of the cancelling terms A dx^2, C dy^2, 2 B dx dy (1e5 - 1e6 for a
// long thin splat centred far outside the tile): the threshold is therefore widened by 0.01 plus
// 2e-6 (~32 float epsilons) of those terms at the unit's farthest corner.  The remaining margins
// (0.02 px, 0.05% of the half width) cover the interval arithmetic itself; a non-finite interval
// falls back to the whole row.
__device__ __forceinline__ uint32_t entry_hits_fwd(const float4 a, const float4 b, int bx0, int by0, int bx1,
                                                   int by1, float2* mat, int slot, int stride) {
    int x0, x1, y0, y1;
    if (!unit_box(a, bx0, by0, bx1, by1, x0, x1, y0, y1)) return 0u;
    const float dxm = fmaxf(fabsf(a.x - (float)bx0), fabsf(a.x - (float)bx1));
    const float dym = fmaxf(fabsf(a.y - (float)by0), fabsf(a.y - (float)by1));
    const float mag = fabsf(b.x) * dxm * dxm + fabsf(b.z) * dym * dym + 2.0f * fabsf(b.y) * dxm * dym;
    const float tau = __logf(255.0f * b.w) + 0.01f + 2e-6f * mag;
    const float bb = b.y * b.y;
    const float det = fmaf(b.x, b.z, -bb) - fmaf(b.y, b.y, -bb);      // compensated A*C - B*B
    const float inv_a = 1.0f / b.x;
    const float two_tau_a = 2.0f * tau * b.x;
    uint32_t hits = 0;
    for (int y = y0; y <= y1; y++) {
        const float dy = __fsub_rn(a.y, (float)y);
        const float disc = fmaf(-det, dy * dy, two_tau_a);
        if (disc < 0.0f) continue;
        const float root = sqrtf(disc) * 1.0005f;
        const float mid = -b.y * dy;
        const float x_lo = a.x - (mid + root) * inv_a - 0.02f;
        const float x_hi = a.x - (mid - root) * inv_a + 0.02f;
        int xs = x0, xe = x1;
        if (fabsf(x_lo) < 1e9f && fabsf(x_hi) < 1e9f) {              // false for NaN / infinity
            xs = max(x0, (int)ceilf(x_lo));
            xe = min(x1, (int)floorf(x_hi));
        }
        const float cyy = __fmul_rn(__fmul_rn(b.z, dy), dy);
        for (int x = xs; x <= xe; x++) {
            const float dx = __fsub_rn(a.x, (float)x);
            const float sum = __fmaf_rn(__fmul_rn(b.x, dx), dx, cyy);
            const float power = __fmaf_rn(sum, -0.5f, -__fmul_rn(__fmul_rn(b.y, dx), dy));
            if (power > 0.0f) continue;
            const float alpha = fminf(0.99f, __fmul_rn(b.w, expf(power)));
            if (alpha < 1.0f / 255.0f) continue;
            const int pix = (y - by0) * kUnitW + (x - bx0);
            hits |= 1u << pix;
            mat[pix * stride + slot] = make_float2(__fsub_rn(1.0f, alpha), __fmul_rn(a.w, alpha));
        }
    }
    return hits;
}

// power of one (pixel, entry) pair (same arithmetic as above)
__device__ __forceinline__ float pair_power(const float4 a, const float4 b, float pixfx, float pixfy,
                                            float& dx, float& dy) {
    dx = __fsub_rn(a.x, pixfx);
    dy = __fsub_rn(a.y, pixfy);
    const float sum = __fmaf_rn(__fmul_rn(b.x, dx), dx, __fmul_rn(__fmul_rn(b.z, dy), dy));
    return __fmaf_rn(sum, -0.5f, -__fmul_rn(__fmul_rn(b.y, dx), dy));
}

__device__ __forceinline__ float4 ldg_f4(const float4* p) { return __ldg(p); }

}  // namespace vegs
