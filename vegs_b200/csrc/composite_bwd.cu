// Blend-stage backward.  One CTA of 8 warps per 8x4 pixel block ("unit"), 8 units per 16x16 tile.
//
// Replaces BACKWARD::render / renderCUDA (reference cuda_rasterizer/backward.cu:417-694).
//
// The reference lets all 256 pixel threads of a tile evaluate every list entry and issues >= 10
// global float atomics per (pixel, Gaussian) pair (backward.cu:612,654,682-691).  Here the unit
// walks ITS HIT LIST, written by the forward pass (composite_fwd.cu): the entries of the tile's
// list that reached at least one of the unit's 32 pixels, with their pixel hit masks, in list
// order.  It goes back to front from the deepest contributor of its pixels in working sets of 256
// entries (composite_common.cuh):
//   S0  entry-parallel (256 threads, one entry each): alpha of every hit of the entry (same float
//       operations as the forward pass) and v = <colour, depth, 1> . <dL/dpixel colour, depth,
//       alpha>, the hit's value under the pixel's cotangent
//   S1  each warp bit-transposes its 32 hit masks into per-pixel hit lists
//   S2  the reference's sequential recurrences over a pixel's hits, back to front
//       (T /= 1-alpha, and the colour / depth / alpha accumulated behind the hit, backward.cu:
//       586-623).  The five accumulators share their coefficients, so under the cotangent they
//       collapse into one scalar recurrence  A' = alpha*v + (1-alpha)*A  with
//       dL/dalpha = T*(v - A) - T_final*<bg, dL/dcolour>/(1-alpha).  Both T and A are affine in
//       their incoming value, so the 8 warps scan the 8 segments of the working set in parallel:
//       pass 1 reduces every segment to (prod 1/(1-alpha), a, b) per pixel, pass 2 re-walks the
//       segment from the composed incoming state and leaves w = alpha*T and dL/dalpha of every
//       hit in a 32 x 256 shared matrix
//   S3  entry-parallel: thread t sums the ten gradient components of ITS entry over the pixels of
//       its hit mask in registers (recomputing dx, dy, G from the records) and issues three 16-byte
//       vector reductions (red.global.add.v4.f32, SASS REDG.E.ADD.F32x4) into the packed 48-byte
//       per-Gaussian accumulator
// so there are no per-pair atomics, no warp-shuffle reductions, every phase is spread over the 8
// warps of the pixel block, and entries that touch none of the block's pixels are never seen.
// Units are issued longest-list-first.
// The quaternion / scale-map cotangents are not propagated: the reference accumulates them into
// buffers that its next kernel overwrites (backward.cu:624-647 vs :322-325,:340-341), so its
// returned gradients do not contain them.
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

#ifndef VEGS_BWD_SUPER
#define VEGS_BWD_SUPER 256
#endif
constexpr int kSuper = VEGS_BWD_SUPER;      // entries per working set = threads per CTA
#ifdef VEGS_BWD_MINCTAS
constexpr int kMinCtas = VEGS_BWD_MINCTAS;
#else
constexpr int kMinCtas = kSuper <= 256 ? 3 : 1;
#endif
constexpr int kWarps = kSuper / 32;

struct BwdParams {
    const uint2* ranges;
    const uint2* hitlist;
    const uint32_t* last_hit;
    const uint32_t* tile_order;
    int W, H, gx;
    const float4 *recA, *recB, *recC;
    const float* bg;
    const float* out_alpha;
    const float *dL_dcolor, *dL_ddepth, *dL_dalpha;
    float4* acc;
    // opt-in live map gradients: cotangents of the scale / quaternion maps, the forward's
    // normalised quaternion map, and the per-Gaussian accumulator [2P] (scale xyz | quaternion)
    const float *dL_dscale_map, *dL_dquat_map, *out_quat;
    float4* map_acc;
    uint32_t* tile_clocks;
    uint32_t* tile_dbg;
};

struct BwdSmem {
    float2 M[32 * kSuper];          // [pixel][slot]: (alpha, v) after S0, (alpha*T, dL/dalpha) after S2
    float4 pix[32];                 // per pixel: dL/dcolor rgb, dL/ddepth
    float pixa[32];                 // per pixel: dL/dalpha map
    float4 pixs[32];                // per pixel: dL/dscale map (live map gradients only)
    float4 pixq[32];                // per pixel: dL/dquaternion map, projected onto the output's tangent space
    float agg[kWarps][3][32];       // [segment][prod 1/(1-alpha), a, b][pixel]: A_out = a + b * A_in
};

static_assert(sizeof(BwdSmem) <= 227 * 1024, "backward working set does not fit the 227 KB of shared memory per CTA");

#if VEGS_TILE_DEBUG
#define PHASE_INIT long long ph_t = clock64(), ph_acc[4] = {0, 0, 0, 0}
#define PHASE_MARK(k) { long long now_; asm volatile("mov.u64 %0, %%clock64;" : "=l"(now_) :: "memory"); ph_acc[k] += now_ - ph_t; ph_t = now_; }
#else
#define PHASE_INIT
#define PHASE_MARK(k)
#endif

__global__ void __launch_bounds__(kSuper, kMinCtas) composite_bwd_kernel(const BwdParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BwdSmem& sm = *reinterpret_cast<BwdSmem*>(smem_raw);

    const long long t_start = clock64();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t unit = blockIdx.x;
    const uint32_t tile = p.tile_order[unit >> 3];
    const int blk = (int)(unit & 7u);
    const int tile_x = (int)(tile % (uint32_t)p.gx), tile_y = (int)(tile / (uint32_t)p.gx);
    const int bx0 = tile_x * kTile + (blk & 1) * kUnitW;
    const int by0 = tile_y * kTile + (blk >> 1) * kUnitH;
    if (bx0 >= p.W || by0 >= p.H) return;
    const int bx1 = min(bx0 + kUnitW - 1, p.W - 1), by1 = min(by0 + kUnitH - 1, p.H - 1);
    // lane = pixel (every warp computes it)
    const int px = bx0 + (lane & 7), py = by0 + (lane >> 3);
    const bool inside = px < p.W && py < p.H;
    const size_t npix = (size_t)p.W * p.H;
    const size_t pid = (size_t)py * p.W + px;

    // position (+1) of the pixel's last contributor in the unit's hit list
    const uint32_t last_hit = inside ? p.last_hit[pid] : 0u;
    const int n = (int)__reduce_max_sync(0xffffffffu, last_hit);   // deepest contributor of the block
    if (n == 0) return;                                              // same value in every warp
    const uint2 range = p.ranges[tile];
    const uint2* list = p.hitlist + ((size_t)kUnitsPerTile * range.x + (size_t)blk * (size_t)(range.y - range.x));

    const float T_final = inside ? 1.0f - p.out_alpha[pid] : 0.0f;
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
    if (warp == 0) {
        sm.pix[lane] = make_float4(dpix0, dpix1, dpix2, dpix_depth);
        sm.pixa[lane] = dpix_alpha;
        if (p.map_acc != nullptr) {
            float4 ds = make_float4(0.f, 0.f, 0.f, 0.f), dq = ds;
            if (inside) {
                if (p.dL_dscale_map) ds = make_float4(p.dL_dscale_map[pid], p.dL_dscale_map[npix + pid], p.dL_dscale_map[2 * npix + pid], 0.f);
                if (p.dL_dquat_map) {
                    const float4 g = make_float4(p.dL_dquat_map[pid], p.dL_dquat_map[npix + pid], p.dL_dquat_map[2 * npix + pid], p.dL_dquat_map[3 * npix + pid]);
                    const float4 n = make_float4(p.out_quat[pid], p.out_quat[npix + pid], p.out_quat[2 * npix + pid], p.out_quat[3 * npix + pid]);
                    const float ng = n.x * g.x + n.y * g.y + n.z * g.z + n.w * g.w;
                    dq = make_float4(g.x - n.x * ng, g.y - n.y * ng, g.z - n.z * ng, g.w - n.w * ng);
                }
            }
            sm.pixs[lane] = ds;
            sm.pixq[lane] = dq;
        }
    }
    __syncthreads();
    const float bg_term = -T_final * (p.bg[0] * dpix0 + p.bg[1] * dpix1 + p.bg[2] * dpix2);
    const float ddelx_dx = 0.5f * p.W;
    const float ddely_dy = 0.5f * p.H;

    // sequential state of the pixel, replicated in every warp: transmittance in front of the last
    // processed hit and the cotangent-weighted accumulation behind it
    float T_carry = T_final, A_carry = 0.0f;

    // slot t of the working set starting at `base` holds hit-list entry n - 1 - (base + t): back to
    // front.  Software pipeline: the records of the next working set are loaded during this one.
    uint2 ent = make_uint2(0u, 0u);          // (Gaussian id, hit mask); mask 0 = empty slot
    float4 a_cur = make_float4(0.f, 0.f, 0.f, 0.f), b_cur = a_cur, c_cur = a_cur;
    if (n - 1 - tid >= 0) {
        ent = list[n - 1 - tid];
        a_cur = ldg_f4(p.recA + (size_t)kGeomRecStride * ent.x); b_cur = ldg_f4(p.recB + (size_t)kGeomRecStride * ent.x); c_cur = ldg_f4(p.recC + (size_t)kGeomRecStride * ent.x);
    }

    PHASE_INIT;
    for (int base = 0; base < n; base += kSuper) {
        uint2 ent_nxt = make_uint2(0u, 0u);
        float4 a_nxt = make_float4(0.f, 0.f, 0.f, 0.f), b_nxt = a_nxt, c_nxt = a_nxt;
        const int i_nxt = n - 1 - (base + kSuper + tid);
        if (i_nxt >= 0) {
            ent_nxt = list[i_nxt];
            a_nxt = ldg_f4(p.recA + (size_t)kGeomRecStride * ent_nxt.x); b_nxt = ldg_f4(p.recB + (size_t)kGeomRecStride * ent_nxt.x); c_nxt = ldg_f4(p.recC + (size_t)kGeomRecStride * ent_nxt.x);
        }

        // ---- S0: thread = entry.  (alpha, v) of every hit; the forward pass found the hits, the
        // same float operations reproduce their alpha.
        const uint32_t h = ent.y;
        {
            uint32_t bits = h;
            while (bits != 0u) {
                const int q = __ffs(bits) - 1;
                bits &= bits - 1;
                float dx, dy;
                const float power = pair_power(a_cur, b_cur, (float)(bx0 + (q & 7)), (float)(by0 + (q >> 3)), dx, dy);
                const float alpha = fminf(0.99f, __fmul_rn(b_cur.w, expf(power)));
                const float4 dp = sm.pix[q];
                sm.M[q * kSuper + swz(tid, q)] =
                    make_float2(alpha, c_cur.x * dp.x + c_cur.y * dp.y + c_cur.z * dp.z + a_cur.w * dp.w + sm.pixa[q]);
            }
        }
        // ---- S1: per-pixel hit lists of this warp's 32-entry segment
        const uint32_t mybits = transpose32(h, lane);
        const int any_hit = __syncthreads_or(h != 0u);
        PHASE_MARK(0);
        if (any_hit) {
            // slot t holds hit-list entry n - 1 - base - t; entries behind the pixel's last
            // contributor do not take part (backward.cu:561-563): the first slots of the set
            const int s_min = (n - 1 - base) - (int)last_hit + 1 - warp * 32;
            const uint32_t valid = s_min <= 0 ? 0xffffffffu : (s_min >= 32 ? 0u : 0xffffffffu << s_min);
            const float2* Mrow = sm.M + lane * kSuper + warp * 32;
            // ---- S2 pass 1: warp = segment, lane = pixel: the segment as an affine map
            {
                float a = 0.0f, b = 1.0f;
                uint32_t bits = mybits & valid;
                while (bits != 0u) {
                    const int sl = __ffs(bits) - 1;
                    bits &= bits - 1;
                    const float2 m = Mrow[swz(sl, lane)];
                    const float om = 1.0f - m.x;
                    a = m.x * m.y + om * a;
                    b *= om;
                }
                // transmittance ratio across the segment: one reciprocal of the product (blended
                // hits keep T >= 1e-4, so b >= 1e-4: no underflow) instead of one per hit
                const float P = __frcp_rn(b);
                sm.agg[warp][0][lane] = P;
                sm.agg[warp][1][lane] = a;
                sm.agg[warp][2][lane] = b;
            }
            __syncthreads();
            PHASE_MARK(1);
            // ---- S2 pass 2: compose the segments in front of this one, then walk it for real
            {
                float T = T_carry, A = A_carry;
#pragma unroll
                for (int w = 0; w < kWarps; w++) {
                    if (w == warp) { T = T_carry; A = A_carry; }
                    T_carry *= sm.agg[w][0][lane];
                    A_carry = sm.agg[w][1][lane] + sm.agg[w][2][lane] * A_carry;
                }
                float2* Mw = sm.M + lane * kSuper + warp * 32;
                uint32_t bits = mybits;
                while (bits != 0u) {
                    const int sl = __ffs(bits) - 1;
                    bits &= bits - 1;
                    float2 out = make_float2(0.f, 0.f);
                    if ((valid >> sl) & 1u) {
                        const float2 m = Mw[swz(sl, lane)];   // alpha, v
                        const float om = 1.0f - m.x;
                        const float r = __frcp_rn(om);
                        T *= r;                            // transmittance in front of this hit
                        out.x = m.x * T;
                        out.y = T * (m.y - A) + bg_term * r;
                        A = m.x * m.y + om * A;
                    }
                    Mw[swz(sl, lane)] = out;
                }
            }
            __syncthreads();
            PHASE_MARK(2);
            // ---- S3: thread = entry: sum the entry's gradient over the pixels it hit
            if (h != 0u) {
                const float4 a = a_cur, bb = b_cur;
                float g_mx = 0.f, g_my = 0.f, g_cxx = 0.f, g_cxy = 0.f, g_cyy = 0.f, g_op = 0.f;
                float g_c0 = 0.f, g_c1 = 0.f, g_c2 = 0.f, g_dep = 0.f;
                float g_s0 = 0.f, g_s1 = 0.f, g_s2 = 0.f, g_q0 = 0.f, g_q1 = 0.f, g_q2 = 0.f, g_q3 = 0.f;
                const bool live = p.map_acc != nullptr;
                bool any = false;
                uint32_t bits = h;
                while (bits != 0u) {
                    const int q = __ffs(bits) - 1;
                    bits &= bits - 1;
                    const float2 m = sm.M[q * kSuper + swz(tid, q)];
                    if (m.x == 0.0f && m.y == 0.0f) continue;
                    any = true;
                    const float4 dp = sm.pix[q];
                    float dx, dy;
                    const float power = pair_power(a, bb, (float)(bx0 + (q & 7)), (float)(by0 + (q >> 3)), dx, dy);
                    const float G = expf(power);
                    const float wgt = m.x, dL_dalpha = m.y;
                    g_c0 += wgt * dp.x; g_c1 += wgt * dp.y; g_c2 += wgt * dp.z;
                    g_dep += wgt * dp.w;
                    const float dL_dG = bb.w * dL_dalpha;
                    const float gdx = G * dx;
                    const float gdy = G * dy;
                    const float dG_ddelx = -gdx * bb.x - gdy * bb.y;
                    const float dG_ddely = -gdy * bb.z - gdx * bb.y;
                    g_mx += dL_dG * dG_ddelx * ddelx_dx;
                    g_my += dL_dG * dG_ddely * ddely_dy;
                    g_cxx += -0.5f * gdx * dx * dL_dG;
                    g_cxy += -0.5f * gdx * dy * dL_dG;
                    g_cyy += -0.5f * gdy * dy * dL_dG;
                    g_op += G * dL_dalpha;
                    if (live) {
                        const float4 ds = sm.pixs[q], dq = sm.pixq[q];
                        g_s0 += wgt * ds.x; g_s1 += wgt * ds.y; g_s2 += wgt * ds.z;
                        g_q0 += wgt * dq.x; g_q1 += wgt * dq.y; g_q2 += wgt * dq.z; g_q3 += wgt * dq.w;
                    }
                }
                if (any) {
                    float4* dst = p.acc + 3 * (size_t)ent.x;
                    red_add_v4(dst + 0, g_mx, g_my, g_cxx, g_cxy);
                    red_add_v4(dst + 1, g_cyy, g_op, g_c0, g_c1);
                    red_add_v4(dst + 2, g_c2, g_dep, 0.0f, 0.0f);
                    if (live) {
                        float4* md = p.map_acc + 2 * (size_t)ent.x;
                        red_add_v4(md + 0, g_s0, g_s1, g_s2, 0.0f);
                        red_add_v4(md + 1, g_q0, g_q1, g_q2, g_q3);
                    }
                }
            }
        }
        __syncthreads();   // shared arrays are reused by the next working set
        PHASE_MARK(3);
        ent = ent_nxt; a_cur = a_nxt; b_cur = b_nxt; c_cur = c_nxt;
    }
    if (tid == 0) atomicMax(&p.tile_clocks[tile], (uint32_t)((clock64() - t_start) >> 10));
#if VEGS_TILE_DEBUG
    if (tid == 0 && blk == 0)
        for (int k = 0; k < 4; k++) p.tile_dbg[tile * 16 + 4 + k] = (uint32_t)(ph_acc[k] >> 10);
#endif
}

}  // namespace

int launch_composite_bwd(const VegsBackwardArgs& a, const GeomView& g, const ImgView& im,
                         const BinView& b, float4* acc, float4* map_acc, cudaStream_t stream) {
    BwdParams p;
    p.ranges = im.ranges;
    p.hitlist = b.hitlist;
    p.tile_order = im.tile_order;
    p.W = a.W; p.H = a.H;
    p.gx = (a.W + kTile - 1) / kTile;
    p.recA = g.recA; p.recB = g.recB; p.recC = g.recC;
    p.bg = a.background;
    p.out_alpha = a.out_alpha;
    p.last_hit = im.last_hit;
    p.dL_dcolor = a.dL_dout_color; p.dL_ddepth = a.dL_dout_depth; p.dL_dalpha = a.dL_dout_alpha;
    p.acc = acc;
    p.map_acc = map_acc;
    p.dL_dscale_map = a.dL_dout_cov_scale; p.dL_dquat_map = a.dL_dout_cov_quat; p.out_quat = a.out_cov_quat;
    const int T = p.gx * ((a.H + kTile - 1) / kTile);
    p.tile_clocks = im.tile_clocks + T;
    p.tile_dbg = im.tile_dbg;
    // opt-in shared memory size is a per-device function attribute
    static DeviceOnce once;
    int dev = 0;
    VEGS_CUDA_OK(cudaGetDevice(&dev));
    if (once.need(dev)) {
        VEGS_CUDA_OK(cudaFuncSetAttribute(composite_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)sizeof(BwdSmem)));
        once.done(dev);
    }
    composite_bwd_kernel<<<T * kUnitsPerTile, kSuper, sizeof(BwdSmem), stream>>>(p);
    VEGS_LAUNCH_CHECK("composite_bwd", stream, a.debug);
    return VEGS_OK;
}

}  // namespace vegs
