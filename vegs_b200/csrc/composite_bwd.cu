// Blend-stage backward.  One CTA of 8 warps per 8x4 pixel block ("unit"), 8 units per 16x16 tile;
// a unit walks its tile's list back to front from the block's deepest contributor.
//
// Replaces BACKWARD::render / renderCUDA (reference cuda_rasterizer/backward.cu:417-694).
//
// The reference lets all 256 pixel threads of a tile evaluate every list entry and issues >= 10
// global float atomics per (pixel, Gaussian) pair (backward.cu:612,654,682-691).  Here
// (composite_common.cuh) the unit first SCANS the raw list, 256 entries per step, keeping the
// entries whose conservative box reaches the block as candidates in a ring, and then, per working
// set of 256 candidates:
//   S0  entry-parallel (256 threads, one candidate each): the block pixels the entry contributes to
//       (exact test inside the entry's conservative box) as a 32-bit hit mask, each hit's alpha
//       and v = <colour, depth, 1> . <dL/dpixel colour, depth, alpha>, the hit's value under the
//       pixel's cotangent
//   S1  each warp bit-transposes its 32 hit masks into per-pixel hit lists
//   S2  the reference's sequential recurrences over a pixel's hits, back to front
//       (T /= 1-alpha, and the colour / depth / alpha accumulated behind the hit, backward.cu:
//       586-623).  The five accumulators share their coefficients, so under the cotangent they
//       collapse into one scalar recurrence  A' = alpha*v + (1-alpha)*A  with
//       dL/dalpha = T*(v - A) - T_final*<bg, dL/dcolour>/(1-alpha).  Both T and A are affine in
//       their incoming value, so the 8 warps scan the 8 segments of the super-batch in parallel:
//       pass 1 reduces every segment to (prod 1/(1-alpha), a, b) per pixel, pass 2 re-walks the
//       segment from the composed incoming state and leaves w = alpha*T and dL/dalpha of every
//       hit in a 32 x 256 shared matrix
// so there are no per-pair atomics, no warp-shuffle reductions, the expensive parts are spread
// over 8 warps per pixel block, and entries that touch none of the block's pixels cost one box
// test.  Units are issued longest-list-first.
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

constexpr int kSuper = 256;                 // candidates per working set = threads per CTA
constexpr int kWarps = kSuper / 32;
constexpr int kRing = 768;                  // candidate ring, 3 working sets

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
    uint32_t* tile_clocks;
    uint32_t* tile_dbg;
};

struct BwdSmem {
    float2 M[32 * kSuper];          // [pixel][slot]: (alpha, v) after S0, (alpha*T, dL/dalpha) after S2
    float4 pix[32];                 // per pixel: dL/dcolor rgb, dL/ddepth
    float pixa[32];                 // per pixel: dL/dalpha map
    float agg[kWarps][3][32];       // [segment][prod 1/(1-alpha), a, b][pixel]: A_out = a + b * A_in
    uint2 ring[kRing];              // candidates, back to front: (Gaussian id, ordinal in the tile's list)
    uint32_t scan_cnt[2][kWarps];   // survivors per warp of a scan step (double-buffered)
};

#if VEGS_TILE_DEBUG
#define PHASE_INIT long long ph_t = clock64(), ph_acc[4] = {0, 0, 0, 0}
#define PHASE_MARK(k) { long long now_; asm volatile("mov.u64 %0, %%clock64;" : "=l"(now_) :: "memory"); ph_acc[k] += now_ - ph_t; ph_t = now_; }
#else
#define PHASE_INIT
#define PHASE_MARK(k)
#endif

__global__ void __launch_bounds__(kSuper, 3) composite_bwd_kernel(const BwdParams p) {
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

    const uint32_t last_contributor = inside ? p.n_contrib[pid] : 0u;
    const int n = (int)__reduce_max_sync(0xffffffffu, last_contributor);   // deepest contributor of the block
    if (n == 0) return;                                                      // same value in every warp
    const uint2 range = p.ranges[tile];
    const uint32_t* list = p.point_list + range.x;

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
    }
    __syncthreads();
    const float bg_term = -T_final * (p.bg[0] * dpix0 + p.bg[1] * dpix1 + p.bg[2] * dpix2);
    const float ddelx_dx = 0.5f * p.W;
    const float ddely_dy = 0.5f * p.H;

    // sequential state of the pixel, replicated in every warp: transmittance in front of the last
    // processed hit and the cotangent-weighted accumulation behind it
    float T_carry = T_final, A_carry = 0.0f;

    // raw-list pipeline of the scan (raw position r = list entry n - 1 - r: back to front):
    // ids two steps ahead, box records one step ahead
    int raw_base = 0;
    int rid_cur = (n - 1 - tid >= 0) ? (int)list[n - 1 - tid] : -1;
    float4 ra_cur = make_float4(0.f, 0.f, 0.f, 0.f);
    if (rid_cur >= 0) ra_cur = ldg_f4(p.recA + rid_cur);
    int rid_nxt = (n - 1 - kSuper - tid >= 0) ? (int)list[n - 1 - kSuper - tid] : -1;
    uint32_t head = 0, tail = 0;            // ring positions, same in every thread
    int scan_par = 0;
    // working-set pipeline: this thread's candidate of the next working set, when already queued
    uint32_t pre_pos = 0xffffffffu;
    uint2 pre_ent = make_uint2(0u, 0u);
    float4 pre_a = make_float4(0.f, 0.f, 0.f, 0.f), pre_b = pre_a;

    PHASE_INIT;
    for (;;) {
        // ---- SCAN: queue candidates until two working sets are ready or the list ends
        bool scanned = false;
        while (tail - head < 2u * kSuper && raw_base < n) {
            float4 ra_nxt = make_float4(0.f, 0.f, 0.f, 0.f);
            if (rid_nxt >= 0) ra_nxt = ldg_f4(p.recA + rid_nxt);
            const int e_nn = n - 1 - (raw_base + 2 * kSuper + tid);
            const int rid_nn = (e_nn >= 0) ? (int)list[e_nn] : -1;
            int x0, x1, y0, y1;
            const bool keep = rid_cur >= 0 && unit_box(ra_cur, bx0, by0, bx1, by1, x0, x1, y0, y1);
            const uint32_t bal = __ballot_sync(0xffffffffu, keep);
            if (lane == 0) sm.scan_cnt[scan_par][warp] = __popc(bal);
            __syncthreads();
            uint32_t before = 0, total = 0;
#pragma unroll
            for (int w = 0; w < kWarps; w++) {
                const uint32_t c = sm.scan_cnt[scan_par][w];
                before += (w < warp) ? c : 0u;
                total += c;
            }
            if (keep)
                sm.ring[(tail + before + __popc(bal & ((1u << lane) - 1u))) % kRing] =
                    make_uint2((uint32_t)rid_cur, (uint32_t)(n - 1 - (raw_base + tid)));
            tail += total;
            raw_base += kSuper;
            scan_par ^= 1;
            scanned = true;
            rid_cur = rid_nxt; ra_cur = ra_nxt; rid_nxt = rid_nn;
        }
        const int m = min((int)(tail - head), kSuper);
        if (m == 0) break;                   // list exhausted, nothing queued
        if (scanned) __syncthreads();        // ring entries of the last scan step
        PHASE_MARK(0);

        // this thread's candidate: prefetched during the previous working set when it was queued by then
        uint2 ent = make_uint2(0u, 0xffffffffu);
        float4 a_cur = make_float4(0.f, 0.f, 0.f, 0.f), b_cur = a_cur;
        if (tid < m) {
            if (pre_pos == head + tid) {
                ent = pre_ent; a_cur = pre_a; b_cur = pre_b;
            } else {
                ent = sm.ring[(head + tid) % kRing];
                a_cur = ldg_f4(p.recA + ent.x);
                b_cur = ldg_f4(p.recB + ent.x);
            }
        }
        pre_pos = 0xffffffffu;
        if (head + kSuper + tid < tail) {
            pre_pos = head + kSuper + tid;
            pre_ent = sm.ring[pre_pos % kRing];
            pre_a = ldg_f4(p.recA + pre_ent.x);
            pre_b = ldg_f4(p.recB + pre_ent.x);
        }

        // ---- S0: thread = candidate.  Hit mask over the unit's pixels, (alpha, v) of every hit.
        uint32_t h = 0;
        if (tid < m)
            h = entry_hits_alpha(a_cur, b_cur, bx0, by0, bx1, by1, reinterpret_cast<float*>(sm.M), tid, kSuper, 2);
        if (h != 0u) {
            const float4 c = ldg_f4(p.recC + ent.x);
            uint32_t bits = h;
            while (bits != 0u) {
                const int q = __ffs(bits) - 1;
                bits &= bits - 1;
                const float4 dp = sm.pix[q];
                sm.M[q * kSuper + swz(tid, q)].y = c.x * dp.x + c.y * dp.y + c.z * dp.z + a_cur.w * dp.w + sm.pixa[q];
            }
        }
        // ---- S1: per-pixel hit lists of this warp's 32-entry segment
        const uint32_t mybits = transpose32(h, lane);
        const int any_hit = __syncthreads_or(h != 0u);
        PHASE_MARK(1);
        if (any_hit) {
            // entries at or behind the pixel's last contributor do not take part (backward.cu:
            // 561-563).  Ordinals fall along the working set, so per pixel these are the first
            // slots of the segment: binary search over the segment's ordinals (lane = slot)
            const int my_ord = (int)ent.y;                    // -1 for an empty slot
            int lo = 0, hi = 32;
#pragma unroll
            for (int it = 0; it < 6; it++) {           // 33 possible answers
                const int mid = (lo + hi) >> 1;
                const int v = __shfl_sync(0xffffffffu, my_ord, mid & 31);
                if (lo < hi) { if (v >= (int)last_contributor) lo = mid + 1; else hi = mid; }
            }
            const uint32_t valid = lo >= 32 ? 0u : 0xffffffffu << lo;
            const float2* Mrow = sm.M + lane * kSuper + warp * 32;
            // ---- S2 pass 1: warp = segment, lane = pixel: the segment as an affine map
            {
                float P = 1.0f, a = 0.0f, b = 1.0f;
                uint32_t bits = mybits & valid;
                while (bits != 0u) {
                    const int sl = __ffs(bits) - 1;
                    bits &= bits - 1;
                    const float2 m = Mrow[swz(sl, lane)];
                    const float om = 1.0f - m.x;
                    P *= __frcp_rn(om);
                    a = m.x * m.y + om * a;
                    b *= om;
                }
                sm.agg[warp][0][lane] = P;
                sm.agg[warp][1][lane] = a;
                sm.agg[warp][2][lane] = b;
            }
            __syncthreads();
            PHASE_MARK(2);
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
                }
                if (any) {
                    float4* dst = p.acc + 3 * (size_t)ent.x;
                    red_add_v4(dst + 0, g_mx, g_my, g_cxx, g_cxy);
                    red_add_v4(dst + 1, g_cyy, g_op, g_c0, g_c1);
                    red_add_v4(dst + 2, g_c2, g_dep, 0.0f, 0.0f);
                }
            }
        }
        __syncthreads();   // shared arrays are reused by the next working set
        PHASE_MARK(3);
        head += (uint32_t)m;
    }
    if (tid == 0) atomicMax(&p.tile_clocks[tile], (uint32_t)((clock64() - t_start) >> 10));
#if VEGS_TILE_DEBUG
    if (tid == 0 && blk == 0)
        for (int k = 0; k < 4; k++) p.tile_dbg[tile * 8 + 4 + k] = (uint32_t)(ph_acc[k] >> 10);
#endif
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
    p.tile_clocks = const_cast<uint32_t*>(im.tile_clocks) + T;
    p.tile_dbg = const_cast<uint32_t*>(im.tile_dbg);
    static bool configured = false;
    if (!configured) {
        VEGS_CUDA_OK(cudaFuncSetAttribute(composite_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)sizeof(BwdSmem)));
        configured = true;
    }
    composite_bwd_kernel<<<T * kUnitsPerTile, kSuper, sizeof(BwdSmem), stream>>>(p);
    VEGS_LAUNCH_CHECK("composite_bwd", stream, a.debug);
    return VEGS_OK;
}

}  // namespace vegs
