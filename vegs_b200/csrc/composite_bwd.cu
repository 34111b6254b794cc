// Blend-stage backward: per-pixel back-to-front traversal of the tile's Gaussian list.
//
// Replaces BACKWARD::render / renderCUDA (reference cuda_rasterizer/backward.cu:417-694).
//
// What differs from the reference's kernel:
//   * the reference issues >= 10 global float atomics per (pixel, Gaussian) pair from 256 threads
//     that all hit the same Gaussian (backward.cu:612,654,682-691).  Here each warp first reduces
//     its 32 pixels with shuffles, the 8 warps of the tile combine in a shared-memory accumulator,
//     and one thread per Gaussian of the batch flushes it with three 16-byte vector reductions
//     (red.global.add.v4.f32) into a packed 48-byte per-Gaussian record — 3 global reductions per
//     (tile, Gaussian) instead of >= 2560;
//   * the traversal starts at the tile's deepest contributor (max of n_contrib over the tile)
//     instead of at the end of the tile's list: entries behind it are skipped by every pixel in
//     the reference too (backward.cu:561-563), it just fetches them first;
//   * records are staged with double-buffered cp.async like the forward, each warp owns an 8x4
//     pixel block and visits only the entries whose alpha >= 1/255 region overlaps it (same
//     conservative mask as the forward, composite_common.cuh), exp() is skipped below the
//     threshold, and tiles are taken longest-list-first;
//   * the quaternion / scale-map cotangents are not propagated: the reference accumulates them
//     into buffers that its next kernel overwrites (backward.cu:624-647 vs :322-325,:340-341), so
//     its returned gradients do not contain them.
// Accumulator record (12 floats per Gaussian, zeroed by the caller of this launcher):
//   [0..1] dL/dmean2D.xy  [2..4] dL/dconic (xx, xy, yy)  [5] dL/dopacity  [6..8] dL/drgb
//   [9] dL/ddepth  [10..11] unused
#include "common.cuh"
#include "composite_common.cuh"

namespace vegs {

namespace {

__device__ __forceinline__ void red_add_v4(float4* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b),
                 "f"(c), "f"(d)
                 : "memory");
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

constexpr int kAccStride = 11;  // 10 values + "touched" flag; odd stride -> conflict-free flush

struct BwdParams {
    const uint2* ranges;
    const uint32_t* point_list;
    const uint32_t* tile_order;
    int W, H, gx;
    const float4 *recA, *recB, *recC;
    const float* bg;
    const float* out_alpha;
    const uint32_t* n_contrib;
    const float *dL_dcolor, *dL_ddepth, *dL_dalpha;
    float4* acc;
};

__global__ void __launch_bounds__(kTilePix) composite_bwd_kernel(const BwdParams p) {
    __shared__ __align__(16) float4 sA[2][kTilePix];
    __shared__ __align__(16) float4 sB[2][kTilePix];
    __shared__ __align__(16) float4 sC[2][kTilePix];
    __shared__ int sId[2][kTilePix];
    __shared__ float sAcc[kTilePix * kAccStride];
    __shared__ uint32_t sMax[kTilePix / 32];
    __shared__ uint32_t sMask[kTilePix];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t tile = p.tile_order[blockIdx.x];
    const int tile_x = (int)(tile % (uint32_t)p.gx), tile_y = (int)(tile / (uint32_t)p.gx);
    const int X0 = tile_x * kTile, Y0 = tile_y * kTile;
    const int px = X0 + (warp & 1) * 8 + (lane & 7);
    const int py = Y0 + (warp >> 1) * 4 + (lane >> 3);
    const bool inside = px < p.W && py < p.H;
    const float pixfx = (float)px, pixfy = (float)py;
    const size_t npix = (size_t)p.W * p.H;
    const size_t pid = (size_t)py * p.W + px;

    const uint2 range = p.ranges[tile];
    const uint32_t* list = p.point_list + range.x;

    const uint32_t last_contributor = inside ? p.n_contrib[pid] : 0u;
    {
        const uint32_t m = __reduce_max_sync(0xffffffffu, last_contributor);
        if (lane == 0) sMax[warp] = m;
    }
    for (int i = tid; i < kTilePix * kAccStride; i += kTilePix) sAcc[i] = 0.0f;
    __syncthreads();
    uint32_t maxc = 0;
#pragma unroll
    for (int w = 0; w < kTilePix / 32; w++) maxc = max(maxc, sMax[w]);
    if (maxc == 0) return;  // uniform over the block
    const int n = (int)maxc;
    const int nbatches = (n + kTilePix - 1) / kTilePix;

    const float T_final = inside ? 1.0f - p.out_alpha[pid] : 0.0f;
    float T = T_final;
    float dpix0 = 0.f, dpix1 = 0.f, dpix2 = 0.f, dpix_depth = 0.f, dpix_alpha = 0.f;
    if (inside) {
        if (p.dL_dcolor) {
            dpix0 = p.dL_dcolor[pid];
            dpix1 = p.dL_dcolor[npix + pid];
            dpix2 = p.dL_dcolor[2 * npix + pid];
        }
        if (p.dL_ddepth) dpix_depth = p.dL_ddepth[pid];
        if (p.dL_dalpha) dpix_alpha = p.dL_dalpha[pid];
    }
    const float bg_dot = p.bg[0] * dpix0 + p.bg[1] * dpix1 + p.bg[2] * dpix2;
    const float ddelx_dx = 0.5f * p.W;
    const float ddely_dy = 0.5f * p.H;

    float accum_rec0 = 0.f, accum_rec1 = 0.f, accum_rec2 = 0.f, accum_red = 0.f, accum_rea = 0.f;
    float last_alpha = 0.f, last_c0 = 0.f, last_c1 = 0.f, last_c2 = 0.f, last_depth = 0.f;

    // entry index (0-based from the front of the tile list) handled by slot s of batch b
    auto entry_of = [&](int b, int s) { return n - 1 - (b * kTilePix + s); };
    auto issue = [&](int stage, int id) {
        sId[stage][tid] = id;
        if (id >= 0) {
            cp_async16(&sA[stage][tid], p.recA + id);
            cp_async16(&sB[stage][tid], p.recB + id);
            cp_async16(&sC[stage][tid], p.recC + id);
        }
        cp_async_commit();
    };

    {
        const int e = entry_of(0, tid);
        issue(0, e >= 0 ? (int)list[e] : -1);
    }
    int id_next = -1;
    {
        const int e = entry_of(1, tid);
        id_next = (e >= 0) ? (int)list[e] : -1;
    }

    for (int b = 0; b < nbatches; b++) {
        cp_async_wait_all();
        __syncthreads();  // batch b staged; previous batch's flush finished
        if (b + 1 < nbatches) {
            issue((b + 1) & 1, id_next);
            const int e = entry_of(b + 2, tid);
            id_next = (e >= 0) ? (int)list[e] : -1;
        }
        const int st = b & 1;
        const int cnt = min(kTilePix, n - b * kTilePix);
        sMask[tid] = (tid < cnt) ? block_mask(sA[st][tid], sB[st][tid], X0, Y0) : 0u;
        __syncthreads();
        for (int c0 = 0; c0 < cnt; c0 += 32) {
          uint32_t bits = __ballot_sync(0xffffffffu, (sMask[c0 + lane] >> warp) & 1u);
          while (bits) {
            const int j = c0 + __ffs(bits) - 1;
            bits &= bits - 1;
            const uint32_t e = (uint32_t)entry_of(b, j);
            bool active = inside && (e < last_contributor);
            const float4 a = sA[st][j];   // x, y, conic.a, conic.b
            const float4 bb = sB[st][j];  // conic.c, opacity, tau, depth
            const float dx = __fsub_rn(a.x, pixfx);
            const float dy = __fsub_rn(a.y, pixfy);
            const float sum = __fmaf_rn(__fmul_rn(a.z, dx), dx, __fmul_rn(__fmul_rn(bb.x, dy), dy));
            const float power = __fmaf_rn(sum, -0.5f, -__fmul_rn(__fmul_rn(a.w, dx), dy));
            if (power > 0.0f) active = false;
            if (power < -bb.z) active = false;   // alpha < 1/255 with margin
            float G = 0.f, alpha = 0.f;
            if (active) {
                G = expf(power);
                alpha = fminf(0.99f, __fmul_rn(bb.y, G));
                if (alpha < 1.0f / 255.0f) active = false;
            }
            if (!__any_sync(0xffffffffu, active)) continue;

            float g_mx = 0.f, g_my = 0.f, g_cxx = 0.f, g_cxy = 0.f, g_cyy = 0.f, g_op = 0.f;
            float g_c0 = 0.f, g_c1 = 0.f, g_c2 = 0.f, g_dep = 0.f;
            if (active) {
                const float4 c = sC[st][j];  // r, g, b, -
                T = T / (1.0f - alpha);
                const float w = alpha * T;
                float dL_dalpha = 0.0f;
                accum_rec0 = last_alpha * last_c0 + (1.f - last_alpha) * accum_rec0;
                accum_rec1 = last_alpha * last_c1 + (1.f - last_alpha) * accum_rec1;
                accum_rec2 = last_alpha * last_c2 + (1.f - last_alpha) * accum_rec2;
                last_c0 = c.x; last_c1 = c.y; last_c2 = c.z;
                dL_dalpha += (c.x - accum_rec0) * dpix0;
                dL_dalpha += (c.y - accum_rec1) * dpix1;
                dL_dalpha += (c.z - accum_rec2) * dpix2;
                g_c0 = w * dpix0; g_c1 = w * dpix1; g_c2 = w * dpix2;

                accum_red = last_alpha * last_depth + (1.f - last_alpha) * accum_red;
                last_depth = bb.w;
                dL_dalpha += (bb.w - accum_red) * dpix_depth;
                g_dep = w * dpix_depth;

                accum_rea = last_alpha + (1.f - last_alpha) * accum_rea;
                dL_dalpha += (1.f - accum_rea) * dpix_alpha;

                dL_dalpha *= T;
                last_alpha = alpha;
                dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;

                const float dL_dG = bb.y * dL_dalpha;
                const float gdx = G * dx;
                const float gdy = G * dy;
                const float dG_ddelx = -gdx * a.z - gdy * a.w;
                const float dG_ddely = -gdy * bb.x - gdx * a.w;
                g_mx = dL_dG * dG_ddelx * ddelx_dx;
                g_my = dL_dG * dG_ddely * ddely_dy;
                g_cxx = -0.5f * gdx * dx * dL_dG;
                g_cxy = -0.5f * gdx * dy * dL_dG;
                g_cyy = -0.5f * gdy * dy * dL_dG;
                g_op = G * dL_dalpha;
            }
            g_mx = warp_sum(g_mx); g_my = warp_sum(g_my);
            g_cxx = warp_sum(g_cxx); g_cxy = warp_sum(g_cxy); g_cyy = warp_sum(g_cyy);
            g_op = warp_sum(g_op);
            g_c0 = warp_sum(g_c0); g_c1 = warp_sum(g_c1); g_c2 = warp_sum(g_c2);
            g_dep = warp_sum(g_dep);
            if (lane == 0) {
                float* s = &sAcc[j * kAccStride];
                atomicAdd(s + 0, g_mx);  atomicAdd(s + 1, g_my);
                atomicAdd(s + 2, g_cxx); atomicAdd(s + 3, g_cxy); atomicAdd(s + 4, g_cyy);
                atomicAdd(s + 5, g_op);
                atomicAdd(s + 6, g_c0);  atomicAdd(s + 7, g_c1);  atomicAdd(s + 8, g_c2);
                atomicAdd(s + 9, g_dep);
                s[10] = 1.0f;
            }
          }
        }
        __syncthreads();  // all warps' shared-memory accumulation for this batch is complete
        if (tid < cnt) {
            float* s = &sAcc[tid * kAccStride];
            if (s[10] != 0.0f) {
                float4* dst = p.acc + 3 * (size_t)sId[st][tid];
                red_add_v4(dst + 0, s[0], s[1], s[2], s[3]);
                red_add_v4(dst + 1, s[4], s[5], s[6], s[7]);
                red_add_v4(dst + 2, s[8], s[9], 0.0f, 0.0f);
#pragma unroll
                for (int k = 0; k < kAccStride; k++) s[k] = 0.0f;
            }
        }
    }
    cp_async_wait_all();
}

}  // namespace

int launch_composite_bwd(const VegsBackwardArgs& a, const GeomView& g, const ImgView& im,
                         const BinView& b, float4* acc, cudaStream_t stream) {
    BwdParams p;
    p.ranges = im.ranges;
    p.point_list = b.point_list;
    p.tile_order = im.tile_order;
    p.W = a.W; p.H = a.H;
    p.gx = (a.W + kTile - 1) / kTile;
    p.recA = g.recA; p.recB = g.recB; p.recC = g.recC;
    p.bg = a.background;
    p.out_alpha = a.out_alpha;
    p.n_contrib = im.n_contrib;
    p.dL_dcolor = a.dL_dout_color; p.dL_ddepth = a.dL_dout_depth; p.dL_dalpha = a.dL_dout_alpha;
    p.acc = acc;
    const int T = p.gx * ((a.H + kTile - 1) / kTile);
    composite_bwd_kernel<<<T, kTilePix, 0, stream>>>(p);
    VEGS_LAUNCH_CHECK("composite_bwd", stream, a.debug);
    return VEGS_OK;
}

}  // namespace vegs
