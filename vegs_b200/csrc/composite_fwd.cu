// Front-to-back alpha compositing.  One CTA of 8 warps per 8x4 pixel block ("unit"), 8 units per
// 16x16 tile; a unit walks its tile's depth-sorted list in super-batches of 256 entries.
//
// Replaces FORWARD::render / renderCUDA (reference cuda_rasterizer/forward.cu:262-441), including
// the VEGS additions: depth map, alpha map, accumulated covariance quaternion map (sequential
// slerp product, glm::mix — third_party/glm/glm/ext/quaternion_common.inl:4-26) and scale map.
//
// The reference runs one 256-thread block per tile in which every pixel thread evaluates every
// list entry.  At LiDAR density that is the wrong shape: the horizon tiles hold several 1e5 tiny
// Gaussians, each touching a handful of pixels, and one block walks them while the other SMs
// idle.  The only truly sequential part of compositing is the transmittance chain
// T <- T * (1 - alpha) with its saturation test; everything else is parallel.  Per super-batch:
//   S0  entry-parallel (256 threads, one entry each): clip the entry's conservative box
//       (preprocess stage) to the unit's 32 pixels, run the reference's exact contribution test on
//       the pixels inside, keep a 32-bit hit mask and each hit's alpha (alpha matrix in shared
//       memory); entries with a hit stage their packed records
//   S1  each warp bit-transposes its 32 hit masks into per-pixel hit lists
//   S2  warp 0, lane = pixel: the sequential chain over that pixel's hits of the whole super-batch,
//       in list order and in the reference's float operation order: T before each hit, the
//       saturation point (T(1-alpha) < 1e-4 stops the pixel, that Gaussian is not blended),
//       n_contrib, plus the depth sum — a handful of instructions per hit
//   S3  pixel-parallel per 32-entry segment (8 warps, lane = pixel): colour and scale sums and the
//       slerp quaternion product of the segment's hits with the T values from S2
//   S4  warp 0 folds the eight segment partials into the pixel state, in list order
// so the expensive work (exp, 2 sin + 4 div + quaternion product per contribution, box tests) is
// spread over 8 warps per pixel block and 64 warps per tile, units are issued
// longest-list-first, and a unit stops as soon as its 32 pixels are saturated.
// Exactness: which Gaussians contribute to a pixel, alpha, T, n_contrib, depth and alpha map are
// bit-identical to the reference (same float operations in the same order).  The colour / scale
// sums and the quaternion product are associated per 32-entry segment instead of in one running
// chain, a last-bit difference far inside the 1e-5 tolerance (values of magnitude <= 1).
// Roofline: not HBM-bound (32 B per list entry per unit from L2, 48 B more per entry with a hit,
// 52 B written per pixel); the bound is instruction issue of the per-hit blend and the box tests.
#include <cfloat>
#include "common.cuh"
#include "composite_common.cuh"

namespace vegs {

namespace {

#ifndef VEGS_FWD_SUPER
#define VEGS_FWD_SUPER 256
#endif
constexpr int kSuper = VEGS_FWD_SUPER;      // entries per super-batch = threads per CTA
constexpr int kWarps = kSuper / 32;

struct FwdParams {
    const uint2* ranges;
    const uint32_t* point_list;
    const uint32_t* tile_order;
    int W, H, gx;
    const float4 *recA, *recB, *recC, *recD, *recE;
    const float* bg;
    float *out_color, *out_depth, *out_quat, *out_scale, *out_alpha;
    uint32_t* n_contrib;
    uint32_t* tile_clocks;
    uint32_t* tile_dbg;
};

struct FwdSmem {
    float alphaT[32 * kSuper];      // [pixel][slot]: alpha after S0, T-before-the-hit after S2
    float4 A[kSuper], B[kSuper], C[kSuper], D[kSuper], E[kSuper];
    uint32_t bits[kWarps][32];      // [segment][pixel]: hits of the pixel inside the segment
    uint32_t stop_at[32];           // first slot the pixel does not blend any more (kSuper = none)
    float part[kWarps][11][32];     // segment partials: rows 0-2 colour, 4-6 scale, 7-10 quaternion
    uint32_t seg_any[kWarps];       // segment has at least one hit
    int all_done;
};

#if VEGS_TILE_DEBUG
#define PHASE_INIT long long ph_t = clock64(), ph_acc[4] = {0, 0, 0, 0}
#define PHASE_MARK(k) { long long now_; asm volatile("mov.u64 %0, %%clock64;" : "=l"(now_) :: "memory"); ph_acc[k] += now_ - ph_t; ph_t = now_; }
#else
#define PHASE_INIT
#define PHASE_MARK(k)
#endif

__global__ void __launch_bounds__(kSuper) composite_fwd_kernel(const FwdParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    FwdSmem& sm = *reinterpret_cast<FwdSmem*>(smem_raw);

    const long long t_start = clock64();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t unit = blockIdx.x;
    const uint32_t tile = p.tile_order[unit >> 3];
    const int blk = (int)(unit & 7u);
    const int tile_x = (int)(tile % (uint32_t)p.gx), tile_y = (int)(tile / (uint32_t)p.gx);
    const int bx0 = tile_x * kTile + (blk & 1) * kUnitW;
    const int by0 = tile_y * kTile + (blk >> 1) * kUnitH;
    if (bx0 >= p.W || by0 >= p.H) return;                     // block entirely outside the image
    const int bx1 = min(bx0 + kUnitW - 1, p.W - 1), by1 = min(by0 + kUnitH - 1, p.H - 1);
    // lane = pixel (used by S2..S4)
    const int px = bx0 + (lane & 7), py = by0 + (lane >> 3);
    const bool inside = px < p.W && py < p.H;
    const float pixfx = (float)px, pixfy = (float)py;

    const uint2 range = p.ranges[tile];
    const int n = (int)(range.y - range.x);
    const uint32_t* list = p.point_list + range.x;

    // pixel state, meaningful in warp 0 only
    float T = 1.0f;
    float C0 = 0.f, C1 = 0.f, C2 = 0.f, Dacc = 0.f;
    float S0 = 0.f, S1 = 0.f, S2 = 0.f;
    float Qw = 1.f, Qx = 0.f, Qy = 0.f, Qz = 0.f;
    uint32_t last_contributor = 0;
    bool done = !inside;

    // software pipeline over super-batches: records of the next one are loaded during this one
    int id_cur = (tid < n) ? (int)list[tid] : -1;
    float4 a_cur = make_float4(0.f, 0.f, 0.f, 0.f), b_cur = a_cur;
    if (id_cur >= 0) { a_cur = ldg_f4(p.recA + id_cur); b_cur = ldg_f4(p.recB + id_cur); }
    int id_nxt = (kSuper + tid < n) ? (int)list[kSuper + tid] : -1;

    PHASE_INIT;
    for (int base = 0; base < n; base += kSuper) {
        float4 a_nxt = make_float4(0.f, 0.f, 0.f, 0.f), b_nxt = a_nxt;
        if (id_nxt >= 0) { a_nxt = ldg_f4(p.recA + id_nxt); b_nxt = ldg_f4(p.recB + id_nxt); }
        const int nn = base + 2 * kSuper + tid;
        const int id_nn = (nn < n) ? (int)list[nn] : -1;

        // ---- S0: thread = entry.  Hit mask over the unit's pixels, alpha of every hit.
        uint32_t h = 0;
        if (id_cur >= 0) h = entry_hits_alpha(a_cur, b_cur, bx0, by0, bx1, by1, sm.alphaT, tid, kSuper, 1);
        if (h != 0u) {
            sm.A[tid] = a_cur;
            sm.B[tid] = b_cur;
            sm.C[tid] = ldg_f4(p.recC + id_cur);
            sm.D[tid] = ldg_f4(p.recD + id_cur);
            sm.E[tid] = ldg_f4(p.recE + id_cur);
        }
        // ---- S1: per-pixel hit lists of this warp's 32-entry segment
        const uint32_t mybits = transpose32(h, lane);   // lane = pixel, bit = slot inside the segment
        sm.bits[warp][lane] = mybits;
        const bool seg_any = __any_sync(0xffffffffu, h != 0u);
        if (lane == 0) sm.seg_any[warp] = seg_any ? 1u : 0u;
        __syncthreads();
        PHASE_MARK(0);

        // ---- S2: warp 0, lane = pixel: sequential transmittance chain over the super-batch
        if (warp == 0) {
            uint32_t stop = kSuper;
            if (!done) {
                // the pixel's hits of all segments as one stream; the next hit's operands are
                // loaded before the current one is processed (the loop is latency-bound)
                int w = 0;
                uint32_t bits = sm.bits[0][lane];
                auto next_hit = [&]() -> int {
                    while (bits == 0u && w < kWarps - 1) bits = sm.bits[++w][lane];
                    if (bits == 0u) return -1;
                    const int t = w * 32 + __ffs(bits) - 1;
                    bits &= bits - 1;
                    return t;
                };
                int t = next_hit();
                float alpha = 0.f, depth = 0.f;
                if (t >= 0) { alpha = sm.alphaT[lane * kSuper + swz(t, lane)]; depth = sm.B[t].w; }
                while (t >= 0) {
                    const int t2 = next_hit();
                    float alpha2 = 0.f, depth2 = 0.f;
                    if (t2 >= 0) { alpha2 = sm.alphaT[lane * kSuper + swz(t2, lane)]; depth2 = sm.B[t2].w; }
                    const float test_T = __fmul_rn(T, __fsub_rn(1.0f, alpha));
                    if (test_T < 0.0001f) {      // not blended; the pixel is finished (forward.cu:382-386)
                        done = true;
                        stop = (uint32_t)t;
                        break;
                    }
                    sm.alphaT[lane * kSuper + swz(t, lane)] = T;   // transmittance in front of this hit
                    // depth is accumulated here, in the reference's running order, so it stays
                    // bit-identical: depth values are ~1e2, re-associating that sum per segment
                    // already costs more than the 1e-5 tolerance
                    Dacc = __fmaf_rn(__fmul_rn(depth, alpha), T, Dacc);
                    T = test_T;
                    last_contributor = (uint32_t)(base + t) + 1u;   // ordinal in the tile's list
                    t = t2; alpha = alpha2; depth = depth2;
                }
            } else {
                stop = 0u;
            }
            sm.stop_at[lane] = stop;
            const bool all = __all_sync(0xffffffffu, done);
            if (lane == 0) sm.all_done = all ? 1 : 0;
        }
        __syncthreads();
        PHASE_MARK(1);

        // ---- S3: warp = segment, lane = pixel: blend terms of the segment's hits
        uint32_t seg_mask = 0;
#pragma unroll
        for (int w = 0; w < kWarps; w++) seg_mask |= (sm.seg_any[w] != 0u ? 1u : 0u) << w;
        const bool finished = sm.all_done != 0;
        if ((seg_mask >> warp) & 1u) {
            const uint32_t stop = sm.stop_at[lane];
            uint32_t bits = mybits;
            float c0 = 0.f, c1 = 0.f, c2 = 0.f, s0 = 0.f, s1 = 0.f, s2 = 0.f;
            float qw = 1.f, qx = 0.f, qy = 0.f, qz = 0.f;
            while (bits != 0u) {
                const int t = warp * 32 + __ffs(bits) - 1;
                bits &= bits - 1;
                if ((uint32_t)t >= stop) break;
                const float4 a = sm.A[t];    // x, y, conic.a, conic.b
                const float4 bb = sm.B[t];   // conic.c, opacity, half2(ex, ey), depth
                float dx, dy;
                const float power = pair_power(a, bb, pixfx, pixfy, dx, dy);
                const float alpha = fminf(0.99f, __fmul_rn(bb.y, expf(power)));
                const float Tb = sm.alphaT[lane * kSuper + swz(t, lane)];
                const float4 c = sm.C[t];    // r, g, b, 1 / sin(theta)
                const float4 d = sm.D[t];    // scale xyz, rot.r
                const float4 e = sm.E[t];    // rot xyz, theta
                c0 = __fmaf_rn(__fmul_rn(c.x, alpha), Tb, c0);
                c1 = __fmaf_rn(__fmul_rn(c.y, alpha), Tb, c1);
                c2 = __fmaf_rn(__fmul_rn(c.z, alpha), Tb, c2);
                s0 = __fmaf_rn(__fmul_rn(d.x, alpha), Tb, s0);
                s1 = __fmaf_rn(__fmul_rn(d.y, alpha), Tb, s1);
                s2 = __fmaf_rn(__fmul_rn(d.z, alpha), Tb, s2);
                // r = mix(identity, q, w) with q = (w = rot.r, x, y, z), w = alpha * T
                const float wgt = __fmul_rn(alpha, Tb);
                const float om = __fsub_rn(1.0f, wgt);
                float rw, rx, ry, rz;
                if (d.w > 1.0f - FLT_EPSILON) {
                    rw = __fmaf_rn(d.w, wgt, om);
                    rx = __fmul_rn(e.x, wgt);
                    ry = __fmul_rn(e.y, wgt);
                    rz = __fmul_rn(e.z, wgt);
                } else {
                    // glm::mix: (sin((1-w) theta) * identity + sin(w theta) * q) / sin(theta).  With
                    // rot.r = cos(theta) the real part is cos(w theta) identically, and 1 / sin(theta)
                    // is precomputed per Gaussian: no divisions
                    // (sinf + cosf measured faster than sincosf, which returns through local memory)
                    const float wt = __fmul_rn(wgt, e.w);
                    const float sn1 = sinf(wt), cs1 = cosf(wt);
                    const float k = __fmul_rn(sn1, c.w);
                    rw = cs1;
                    rx = __fmul_rn(e.x, k);
                    ry = __fmul_rn(e.y, k);
                    rz = __fmul_rn(e.z, k);
                }
                // q = q * r (Hamilton product, glm/detail/type_quat.inl:286-296)
                const float nw = qw * rw - qx * rx - qy * ry - qz * rz;
                const float nx = qw * rx + qx * rw + qy * rz - qz * ry;
                const float ny = qw * ry + qy * rw + qz * rx - qx * rz;
                const float nz = qw * rz + qz * rw + qx * ry - qy * rx;
                qw = nw; qx = nx; qy = ny; qz = nz;
            }
            float (*pp)[32] = sm.part[warp];
            pp[0][lane] = c0; pp[1][lane] = c1; pp[2][lane] = c2;
            pp[4][lane] = s0; pp[5][lane] = s1; pp[6][lane] = s2;
            pp[7][lane] = qw; pp[8][lane] = qx; pp[9][lane] = qy; pp[10][lane] = qz;
        }
        __syncthreads();
        PHASE_MARK(2);

        // ---- S4: warp 0 folds the segment partials into the pixel state, in list order.  The other
        // warps already start the next super-batch (they only write arrays S4 does not read).
        if (warp == 0) {
            for (int w = 0; w < kWarps; w++) {
                if (((seg_mask >> w) & 1u) == 0u) continue;
                float (*pp)[32] = sm.part[w];
                C0 += pp[0][lane]; C1 += pp[1][lane]; C2 += pp[2][lane];
                S0 += pp[4][lane]; S1 += pp[5][lane]; S2 += pp[6][lane];
                const float rw = pp[7][lane], rx = pp[8][lane], ry = pp[9][lane], rz = pp[10][lane];
                const float nw = Qw * rw - Qx * rx - Qy * ry - Qz * rz;
                const float nx = Qw * rx + Qx * rw + Qy * rz - Qz * ry;
                const float ny = Qw * ry + Qy * rw + Qz * rx - Qx * rz;
                const float nz = Qw * rz + Qz * rw + Qx * ry - Qy * rx;
                Qw = nw; Qx = nx; Qy = ny; Qz = nz;
            }
        }
        PHASE_MARK(3);
        if (finished) break;
        id_cur = id_nxt; a_cur = a_nxt; b_cur = b_nxt; id_nxt = id_nn;
    }

    if (warp == 0 && inside) {
        const size_t npix = (size_t)p.W * p.H;
        const size_t pid = (size_t)py * p.W + px;
        p.out_alpha[pid] = 1.0f - T;
        p.n_contrib[pid] = last_contributor;
        p.out_color[pid] = C0 + T * p.bg[0];
        p.out_color[npix + pid] = C1 + T * p.bg[1];
        p.out_color[2 * npix + pid] = C2 + T * p.bg[2];
        p.out_depth[pid] = Dacc;
        // glm::normalize(quat): identity when the length is not positive
        const float len = sqrtf(Qw * Qw + Qx * Qx + Qy * Qy + Qz * Qz);
        float ow = 1.f, ox = 0.f, oy = 0.f, oz = 0.f;
        if (!(len <= 0.0f)) {
            const float inv = 1.0f / len;
            ow = Qw * inv; ox = Qx * inv; oy = Qy * inv; oz = Qz * inv;
        }
        p.out_quat[pid] = ow;
        p.out_quat[npix + pid] = ox;
        p.out_quat[2 * npix + pid] = oy;
        p.out_quat[3 * npix + pid] = oz;
        p.out_scale[pid] = S0;
        p.out_scale[npix + pid] = S1;
        p.out_scale[2 * npix + pid] = S2;
    }
    if (tid == 0) atomicMax(&p.tile_clocks[tile], (uint32_t)((clock64() - t_start) >> 10));
#if VEGS_TILE_DEBUG
    if (tid == 0 && blk == 0)
        for (int k = 0; k < 4; k++) p.tile_dbg[tile * 8 + k] = (uint32_t)(ph_acc[k] >> 10);
#endif
}

}  // namespace

int launch_composite_fwd(const VegsForwardArgs& a, const GeomView& g, const ScratchView& s,
                         const ImgView& im, const BinView& b, cudaStream_t stream) {
    FwdParams p;
    p.ranges = im.ranges;
    p.point_list = b.point_list;
    p.tile_order = im.tile_order;
    p.W = a.W; p.H = a.H;
    p.gx = (a.W + kTile - 1) / kTile;
    p.recA = g.recA; p.recB = g.recB; p.recC = g.recC; p.recD = s.recD; p.recE = s.recE;
    p.bg = a.background;
    p.out_color = a.out_color; p.out_depth = a.out_depth; p.out_quat = a.out_cov_quat;
    p.out_scale = a.out_cov_scale; p.out_alpha = a.out_alpha;
    p.n_contrib = im.n_contrib;
    p.tile_clocks = im.tile_clocks;
    p.tile_dbg = im.tile_dbg;
    const int T = p.gx * ((a.H + kTile - 1) / kTile);
    static bool configured = false;
    if (!configured) {
        VEGS_CUDA_OK(cudaFuncSetAttribute(composite_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)sizeof(FwdSmem)));
        configured = true;
    }
    VEGS_CUDA_OK(cudaMemsetAsync(im.tile_clocks, 0, (size_t)2 * T * sizeof(uint32_t), stream));
#if VEGS_TILE_DEBUG
    VEGS_CUDA_OK(cudaMemsetAsync(im.tile_dbg, 0, (size_t)8 * T * sizeof(uint32_t), stream));
#endif
    composite_fwd_kernel<<<T * kUnitsPerTile, kSuper, sizeof(FwdSmem), stream>>>(p);
    VEGS_LAUNCH_CHECK("composite_fwd", stream, a.debug);
    return VEGS_OK;
}

}  // namespace vegs
