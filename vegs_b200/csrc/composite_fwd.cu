// Front-to-back alpha compositing.  One CTA of W = 6 warps (kSuper = 192 threads = candidates per
// working set; a build parameter) per 8x4 pixel block ("unit"), 8 units per 16x16 tile; a unit walks
// its tile's depth-sorted list.
//
// Replaces FORWARD::render / renderCUDA (reference cuda_rasterizer/forward.cu:262-441), including
// the VEGS additions: depth map, alpha map, accumulated covariance quaternion map (sequential
// slerp product, glm::mix — third_party/glm/glm/ext/quaternion_common.inl:4-26) and scale map.
//
// The reference runs one 256-thread block per tile in which every pixel thread evaluates every
// list entry.  At LiDAR density that is the wrong shape: the horizon tiles hold several 1e5 tiny
// Gaussians, each touching a handful of pixels, and one block walks them while the other SMs
// idle.  The only truly sequential part of compositing is the transmittance chain
// T <- T * (1 - alpha) with its saturation test; everything else is parallel.
//   SCAN  2 x kSuper raw list entries per step, two per thread: keep the entries whose conservative box
//         (preprocess stage) reaches the unit's 32 pixels - one bit of the list word, set by the
//         emission kernel (binning.cu); survivors are appended in list order to a ring of
//         candidates (ballot + prefix over the W warps).  Repeated until two working sets are
//         queued or the list ends.
//   then, per working set of kSuper candidates:
//   S0  entry-parallel: run the reference's exact contribution test on the pixels inside the box
//       (each pixel row narrowed to the interval the quadratic can reach), keep a 32-bit hit mask
//       and (1 - alpha, depth * alpha) of each hit in a 32 x (kSuper + 1) shared matrix; entries with a hit
//       stage their packed records and are appended to the unit's hit list for the backward pass
//   S1  each warp bit-transposes its 32 hit masks into per-pixel hit lists and compacts each
//       pixel's operands to the front of its row segment
//   S2  warp 0, lane = pixel: the sequential chain over that pixel's hits of the working set, in
//       list order and in the reference's float operation order: T, the saturation point
//       (T(1-alpha) < 1e-4 stops the pixel, that Gaussian is not blended), n_contrib, plus the
//       depth sum — rounds of four hits over dense arrays, a bare multiply chain with a fallback
//       for the rare round that holds a saturation; T at the start of each segment is kept for S3
//   S3  pixel-parallel per 32-candidate segment (W warps, lane = pixel): colour and scale sums and
//       the slerp quaternion product of the segment's hits, from the stored 1 - alpha (no second
//       exp per hit) and the segment's incoming T, advanced by the same multiplies as in S2
//   S4  warps 0 and 1 fold the W segment partials into the pixel state, in list order
// so the expensive work (exp, sin, cos, quaternion product per contribution) is spread over W
// warps per pixel block and 8 W warps per tile, the fixed costs of a working set are paid per kSuper
// candidates rather than per kSuper list entries, units are issued longest-list-first, and a unit
// stops as soon as its 32 pixels are saturated.
// Exactness: which Gaussians contribute to a pixel, alpha, T, n_contrib, depth and alpha map are
// bit-identical to the reference (same float operations in the same order).  The colour / scale
// sums and the quaternion product are associated per 32-candidate segment instead of in one
// running chain, a last-bit difference far inside the 1e-5 tolerance (values of magnitude <= 1).
// Roofline: not HBM-bound (4 B per list entry per unit from L2, 36 B more per candidate, 48 B
// more per entry with a hit, 52 B written per pixel); the launch is bound by CTA
// residency (shared memory: three CTAs per SM) times the latency of a working set's phase chain;
// profiles/r02e_tile_stats.txt, DESIGN.md §8.
#include <cfloat>
#include "common.cuh"
#include "composite_common.cuh"

namespace vegs {

namespace {

// 192 candidates per working set: 73.5 KB of shared memory and 96 registers, i.e. THREE CTAs per SM;
// measured against 256 (two CTAs, 1.665 ms) and 128 (four, 1.875 ms): 1.598 ms per view at 6 M
// (profiles/r02_composite_working_set_sweep.json)
#ifndef VEGS_FWD_SUPER
#define VEGS_FWD_SUPER 192
#endif
constexpr int kSuper = VEGS_FWD_SUPER;      // candidates per working set = threads per CTA
constexpr int kWarps = kSuper / 32;
#ifdef VEGS_FWD_MINCTAS
constexpr int kMinCtas = VEGS_FWD_MINCTAS;
#else
constexpr int kMinCtas = kSuper > 256 ? 1 : (kSuper == 256 ? 2 : (kSuper == 192 ? 3 : 4));
#endif
#ifndef VEGS_FWD_SCANSUB
#define VEGS_FWD_SCANSUB 2
#endif
constexpr int kScanSub = VEGS_FWD_SCANSUB;  // list entries per thread and scan step
constexpr int kScanStep = kScanSub * kSuper;
constexpr int kRow = kSuper + 1;             // row stride of the pixel x slot matrix
constexpr int kRing = (kSuper & (kSuper - 1)) == 0 ? 4 * kSuper : 2 * kSuper + kScanStep;   // candidate ring (>= 2 * kSuper + kScanStep)

struct FwdParams {
    const uint2* ranges;
    const uint32_t* point_list;
    const uint32_t* tile_order;
    int W, H, gx;
    const float4 *recA, *recB, *recC, *recD, *recE;
    const float* bg;
    float *out_color, *out_depth, *out_quat, *out_scale, *out_alpha;
    uint32_t* n_contrib;
    uint32_t* last_hit;
    uint2* hitlist;
    uint32_t* tile_clocks;
    uint32_t* tile_dbg;
};

struct FwdSmem {
    float2 M[32 * kRow + 4];        // [pixel][slot], rows padded by one entry so that lanes reading the same
                                    // column of their own rows spread over the banks: (1 - alpha, depth * alpha)
    float4 C[kSuper], D[kSuper], E[kSuper];   // blend records of candidates with a hit (colour, scale, rotation)
    uint2 ring[kRing];              // candidates in list order: (Gaussian id, ordinal in the tile's list)
    uint32_t bits[kWarps][32];      // [segment][pixel]: hits of the pixel inside the segment
    uint32_t stop_at[32];           // first (segment * 32 + hit index) the pixel does not blend (kSuper = none)
    float part[kWarps][10][32];     // segment partials: rows 0-2 colour, 3-5 scale, 6-9 quaternion
    float tin[kWarps][32];          // [segment][pixel]: transmittance in front of the segment's first hit
    uint32_t seg_bal[kWarps];       // per segment: ballot of the slots with a hit
    uint32_t scan_cnt[2][kScanSub * kWarps];   // survivors per (sub-step, warp) of a scan step (double-buffered)
    int all_done;
#ifdef VEGS_FWD_PAD
    unsigned char pad[VEGS_FWD_PAD];    // occupancy experiments: lowers the number of resident CTAs
#endif
};

static_assert(sizeof(FwdSmem) <= 227 * 1024, "forward working set does not fit the 227 KB of shared memory per CTA");

// sin and cos of x in [0, pi] (w * theta of the slerp: w in [0, 1], theta = acos(.)), ~1 ulp: one
// Cody-Waite reduction to [-pi/4, pi/4] shared by both, Cephes single-precision polynomials.  The
// library sinf + cosf pair reduces twice and carries a large-argument path that cannot occur here.
__device__ __forceinline__ void sincos_0_pi(float x, float& s, float& c) {
    const float j = rintf(x * 0.636619772f);                 // 0, 1 or 2 quarter turns
    float r = fmaf(j, -1.5703125f, x);
    r = fmaf(j, -4.837512969970703125e-4f, r);
    r = fmaf(j, -7.54978995489188e-8f, r);
    const float r2 = r * r;
    const float sp = fmaf(fmaf(fmaf(-1.9515295891e-4f, r2, 8.3321608736e-3f), r2, -1.6666654611e-1f), r2 * r, r);
    const float cp = fmaf(fmaf(fmaf(2.443315711809948e-5f, r2, -1.388731625493765e-3f), r2, 4.166664568298827e-2f),
                          r2 * r2, fmaf(-0.5f, r2, 1.0f));
    const int q = (int)j;
    s = (q == 1) ? cp : (q == 2 ? -sp : sp);
    c = (q == 1) ? -sp : (q == 2 ? -cp : cp);
}

__device__ __forceinline__ void prefetch_l2(const void* ptr) {
    asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr));
}

#if VEGS_TILE_DEBUG
#define PHASE_INIT long long ph_t = clock64(), ph_acc[4] = {0, 0, 0, 0}; uint32_t dbg_sets = 0, dbg_cand = 0, dbg_pix = 0
#define PHASE_MARK(k) { long long now_; asm volatile("mov.u64 %0, %%clock64;" : "=l"(now_) :: "memory"); ph_acc[k] += now_ - ph_t; ph_t = now_; }
#else
#define PHASE_INIT
#define PHASE_MARK(k)
#endif

template <bool kMasks>
__global__ void __launch_bounds__(kSuper, kMinCtas) composite_fwd_kernel(const FwdParams p) {
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
#ifdef VEGS_FWD_SKIP_TOP
    if ((int)(unit >> 3) < VEGS_FWD_SKIP_TOP) {               // timing experiment: without the longest lists
        const int qx = bx0 + (threadIdx.x & 7), qy = by0 + ((threadIdx.x >> 3) & 3);
        if (threadIdx.x < 32 && qx < p.W && qy < p.H) {
            const size_t q = (size_t)qy * p.W + qx;
            p.n_contrib[q] = 0; p.last_hit[q] = 0; p.out_alpha[q] = 0.f;
        }
        return;
    }
#endif
    const int bx1 = min(bx0 + kUnitW - 1, p.W - 1), by1 = min(by0 + kUnitH - 1, p.H - 1);
    // lane = pixel (used by S2..S4)
    const int px = bx0 + (lane & 7), py = by0 + (lane >> 3);
    const bool inside = px < p.W && py < p.H;

    const uint2 range = p.ranges[tile];
    const int n = (int)(range.y - range.x);
    const uint32_t* list = p.point_list + range.x;

    // pixel state: T, depth, quaternion, contributor counts in warp 0; colour and scale sums in warp 1
    float T = 1.0f;
    float C0 = 0.f, C1 = 0.f, C2 = 0.f, Dacc = 0.f;
    float S0 = 0.f, S1 = 0.f, S2 = 0.f;
    float Qw = 1.f, Qx = 0.f, Qy = 0.f, Qz = 0.f;
    uint32_t last_contributor = 0, last_hit = 0;
    bool done = !inside;
    // entries with a hit are appended, in list order, to the unit's hit list for the backward pass
    uint2* hitlist = p.hitlist + ((size_t)kUnitsPerTile * range.x + (size_t)blk * (size_t)n);
    uint32_t hit_count = 0;

    // raw-list pipeline of the scan: the list words of the next step are loaded during this one
    int raw_base = 0;
    uint32_t w_cur[kScanSub];
#pragma unroll
    for (int j = 0; j < kScanSub; j++) w_cur[j] = (j * kSuper + tid < n) ? list[j * kSuper + tid] : 0u;
    uint32_t head = 0, tail = 0;            // ring positions, same in every thread
    int scan_par = 0;
    // working-set pipeline: this thread's candidate of the next working set, when already queued
    uint32_t pre_pos = 0xffffffffu, pre_ord = 0, pre_id = 0;
    float4 pre_a = make_float4(0.f, 0.f, 0.f, 0.f), pre_b = pre_a;

    PHASE_INIT;
    for (;;) {
        // ---- SCAN: queue candidates until two working sets are ready or the list ends.  With unit
        // masks in the list words (common.cuh) a step looks at 2 x 256 words and nothing else;
        // without them (ids beyond 24 bits) it gathers the box record of every entry.
        bool scanned = false;
        while (tail - head < 2u * kSuper && raw_base < n) {
            uint32_t w_nxt[kScanSub];
#pragma unroll
            for (int j = 0; j < kScanSub; j++) {
                const int r = raw_base + kScanStep + j * kSuper + tid;
                w_nxt[j] = (r < n) ? list[r] : 0u;
            }
            bool keep[kScanSub];
            uint32_t bal[kScanSub];
            if (kMasks) {
#pragma unroll
                for (int j = 0; j < kScanSub; j++) keep[j] = ((w_cur[j] >> (24 + blk)) & 1u) != 0u;
            } else {
                float4 ra[kScanSub];
#pragma unroll
                for (int j = 0; j < kScanSub; j++) {
                    ra[j] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (raw_base + j * kSuper + tid < n) ra[j] = ldg_f4(p.recA + (size_t)kGeomRecStride * w_cur[j]);
                }
#pragma unroll
                for (int j = 0; j < kScanSub; j++) {
                    int x0, x1, y0, y1;
                    keep[j] = raw_base + j * kSuper + tid < n && unit_box(ra[j], bx0, by0, bx1, by1, x0, x1, y0, y1);
                }
            }
#pragma unroll
            for (int j = 0; j < kScanSub; j++) {
                bal[j] = __ballot_sync(0xffffffffu, keep[j]);
                if (lane == 0) sm.scan_cnt[scan_par][j * kWarps + warp] = __popc(bal[j]);
            }
            __syncthreads();
            uint32_t run = 0;
#pragma unroll
            for (int j = 0; j < kScanSub; j++) {
#pragma unroll
                for (int w = 0; w < kWarps; w++) {
                    if (w == warp && keep[j]) {
                        const uint32_t gid = kMasks ? (w_cur[j] & kListIdMask) : w_cur[j];
                        sm.ring[(tail + run + __popc(bal[j] & ((1u << lane) - 1u))) % (uint32_t)kRing] =
                            make_uint2(gid, (uint32_t)(raw_base + j * kSuper + tid));
                        // the candidate's records will be gathered by another thread one or two
                        // working sets from now: start them on their way to L2
                        // ({A, B}, {C, pad} and {D, E} are one 32-byte sector each)
                        prefetch_l2(p.recA + (size_t)kGeomRecStride * gid);
                        prefetch_l2(p.recC + (size_t)kGeomRecStride * gid);
                        prefetch_l2(p.recD + (size_t)kScratchRecStride * gid);
                    }
                    run += sm.scan_cnt[scan_par][j * kWarps + w];
                }
            }
            tail += run;
            raw_base += kScanStep;
            scan_par ^= 1;
            scanned = true;
#pragma unroll
            for (int j = 0; j < kScanSub; j++) w_cur[j] = w_nxt[j];
        }
        const int m = min((int)(tail - head), kSuper);
        if (m == 0) break;                   // list exhausted, nothing queued
        if (scanned) __syncthreads();        // ring entries of the last scan step
        PHASE_MARK(0);

        // this thread's candidate: prefetched during the previous working set when it was queued by then
        uint32_t ord = 0, id = 0;
        float4 a_cur = make_float4(0.f, 0.f, 0.f, 0.f), b_cur = a_cur;
        if (tid < m) {
            if (pre_pos == head + tid) {
                ord = pre_ord; id = pre_id; a_cur = pre_a; b_cur = pre_b;
            } else {
                const uint2 ent = sm.ring[(head + tid) % (uint32_t)kRing];
                id = ent.x; ord = ent.y;
                a_cur = ldg_f4(p.recA + (size_t)kGeomRecStride * id);
                b_cur = ldg_f4(p.recB + (size_t)kGeomRecStride * id);
            }
        }
        pre_pos = 0xffffffffu;
        if (head + kSuper + tid < tail) {
            pre_pos = head + kSuper + tid;
            const uint2 ent = sm.ring[pre_pos % (uint32_t)kRing];
            pre_id = ent.x; pre_ord = ent.y;
            pre_a = ldg_f4(p.recA + (size_t)kGeomRecStride * pre_id);
            pre_b = ldg_f4(p.recB + (size_t)kGeomRecStride * pre_id);
        }

        // ---- S0: thread = candidate.  Hit mask over the unit's pixels, alpha of every hit.
        // The blend records are requested first so that their latency hides behind the hit test
        // (nearly every candidate has a hit).
        uint32_t h = 0;
        if (tid < m) {
            const float4 c_rec = ldg_f4(p.recC + (size_t)kGeomRecStride * id), d_rec = ldg_f4(p.recD + (size_t)kScratchRecStride * id), e_rec = ldg_f4(p.recE + (size_t)kScratchRecStride * id);
            h = entry_hits_fwd(a_cur, b_cur, bx0, by0, bx1, by1, sm.M, tid, kRow);
            if (h != 0u) {
                sm.C[tid] = c_rec;
                sm.D[tid] = d_rec;
                sm.E[tid] = e_rec;
            }
        }
        // ---- S1: per-pixel hit lists of this warp's 32-entry segment; the pixel's operands are
        // compacted to the front of the segment's 32 columns (hit j of the segment at column j), in
        // place: the write position never passes the read position.  The sequential pass then
        // walks dense arrays instead of chasing bit masks.
        const uint32_t mybits = transpose32(h, lane);   // lane = pixel, bit = slot inside the segment
        sm.bits[warp][lane] = mybits;
        __syncwarp();                                   // the segment's S0 writes come from this warp
        {
            float2* row = sm.M + lane * kRow + warp * 32;
            uint32_t bits = mybits;
            int j = 0;
            // four per round: all reads of a round come before its writes (slots >= their targets)
            while (bits != 0u) {
                int sl[4];
                float2 v[4];
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    sl[k] = bits != 0u ? __ffs(bits) - 1 : -1;
                    bits &= bits - 1;       // 0 stays 0
                    if (sl[k] >= 0) v[k] = row[sl[k]];
                }
#pragma unroll
                for (int k = 0; k < 4; k++)
                    if (sl[k] >= 0 && sl[k] != j + k) row[j + k] = v[k];
                j += 4;
            }
        }
        const uint32_t hit_bal = __ballot_sync(0xffffffffu, h != 0u);
        if (lane == 0) sm.seg_bal[warp] = hit_bal;
        __syncthreads();
        PHASE_MARK(1);
        uint32_t seg_mask = 0, hits_before = 0, hits_total = 0;
#pragma unroll
        for (int w = 0; w < kWarps; w++) {
            const uint32_t c = __popc(sm.seg_bal[w]);
            seg_mask |= (c != 0u ? 1u : 0u) << w;
            hits_before += (w < warp) ? c : 0u;
            hits_total += c;
        }
        if (h != 0u)
            hitlist[hit_count + hits_before + __popc(hit_bal & ((1u << lane) - 1u))] = make_uint2(id, h);
        const uint32_t hit_base = hit_count;
        hit_count += hits_total;
#if VEGS_TILE_DEBUG
        dbg_sets++; dbg_cand += (uint32_t)m; dbg_pix += (uint32_t)__syncthreads_count(0) + __reduce_add_sync(0xffffffffu, (uint32_t)__popc(h));
#endif

        // ---- S2: warp 0, lane = pixel: sequential transmittance chain over the working set.
        // This is the one phase the other warps of the CTA wait for, and it is a dependent chain, so
        // what counts is the number of dependent operations per hit.  Rounds of four hits: the four
        // products T(1 - alpha) are formed as a bare multiply chain (a slot past the end of the
        // pixel's list multiplies by one); T never grows, so the round contains a saturation
        // (T(1 - alpha) < 1e-4: that Gaussian is not blended and the pixel is finished,
        // forward.cu:382-386) exactly when its last product is below the threshold.  That happens at
        // most once per pixel in the whole launch: a round without one in any lane - nearly all -
        // takes the products as they are, and the depth sum runs beside them as a second,
        // independent chain of four fused multiply-adds; a round with one falls back to the
        // hit-by-hit form.  Same float operations in the same order as the reference either way.
        if (warp == 0) {
            // first (segment * 32 + hit index) the pixel does not blend; a finished pixel blends nothing
            uint32_t stop = done ? 0u : (uint32_t)kSuper;
            int last_w = -1, last_j = 0;
            bool stopped = done;             // every lane stays in the loops below (warp votes)
#pragma unroll 1
            for (int w = 0; w < kWarps; w++) {
                if (((seg_mask >> w) & 1u) == 0u) continue;     // no hit in this segment at all
                int cnt = stopped ? 0 : __popc(sm.bits[w][lane]);
                const float2* row = sm.M + lane * kRow + w * 32;
                sm.tin[w][lane] = T;         // S3 rebuilds the transmittance in front of each hit from here
                const int rounds = (__reduce_max_sync(0xffffffffu, cnt) + 3) >> 2;
                float2 m[4];
#pragma unroll
                for (int k = 0; k < 4; k++) m[k] = row[k];      // may run into the next columns: unused
                for (int r = 0; r < rounds; r++) {
                    float x[4], y[4];
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const bool active = 4 * r + k < cnt;
                        x[k] = active ? m[k].x : 1.0f;          // 1 - alpha
                        y[k] = active ? m[k].y : 0.0f;          // depth * alpha
                    }
                    if (r + 1 < rounds) {
#pragma unroll
                        for (int k = 0; k < 4; k++) m[k] = row[4 * (r + 1) + k];
                    }
                    const float t1 = __fmul_rn(T, x[0]), t2 = __fmul_rn(t1, x[1]), t3 = __fmul_rn(t2, x[2]),
                                t4 = __fmul_rn(t3, x[3]);
                    if (!__any_sync(0xffffffffu, t4 < 0.0001f)) {
                        // depth is accumulated here, in the reference's running order, so it stays
                        // bit-identical: depth values are ~1e2, re-associating that sum per segment
                        // already costs more than the 1e-5 tolerance
                        Dacc = __fmaf_rn(y[0], T, Dacc);
                        Dacc = __fmaf_rn(y[1], t1, Dacc);
                        Dacc = __fmaf_rn(y[2], t2, Dacc);
                        Dacc = __fmaf_rn(y[3], t3, Dacc);
                        T = t4;
                    } else {
#pragma unroll
                        for (int k = 0; k < 4; k++) {
                            const int j = 4 * r + k;
                            const bool active = j < cnt;
                            const float test_T = __fmul_rn(T, x[k]);
                            const bool sat = active && test_T < 0.0001f;
                            if (active && !sat) {
                                Dacc = __fmaf_rn(y[k], T, Dacc);
                                T = test_T;
                            }
                            if (sat) {           // hits j.. of this pixel are not blended, here or later
                                stop = (uint32_t)(w * 32 + j);
                                stopped = true;
                                cnt = j;
                            }
                        }
                    }
                }
                if (cnt > 0) { last_w = w; last_j = cnt - 1; }      // hits 0 .. cnt-1 of the segment were blended
            }
            done = stopped;
            if (last_w >= 0) {
                // slot of the last blended hit: the (last_j + 1)-th set bit of its segment's mask -
                // the highest one unless the pixel saturated inside that segment
                const uint32_t lb = sm.bits[last_w][lane];
                const int last_t = last_w * 32 + (last_j + 1 == __popc(lb) ? 31 - __clz(lb) : (int)__fns(lb, 0, last_j + 1));
                last_contributor = sm.ring[(head + last_t) % (uint32_t)kRing].y + 1u;   // ordinal in the tile's list
                uint32_t idx = hit_base;                                           // position in the unit's hit list
                for (int w2 = 0; w2 < last_w; w2++) idx += __popc(sm.seg_bal[w2]);
                last_hit = idx + __popc(sm.seg_bal[last_w] & ((1u << (last_t & 31)) - 1u)) + 1u;
            }
            sm.stop_at[lane] = stop;
            const bool all = __all_sync(0xffffffffu, done);
            if (lane == 0) sm.all_done = all ? 1 : 0;
        }
        __syncthreads();
        PHASE_MARK(2);

        // ---- S3: warp = segment, lane = pixel: blend terms of the segment's hits
        const bool finished = sm.all_done != 0;
        if ((seg_mask >> warp) & 1u) {
            const uint32_t stop = sm.stop_at[lane];
            uint32_t bits = mybits;
            float c0 = 0.f, c1 = 0.f, c2 = 0.f, s0 = 0.f, s1 = 0.f, s2 = 0.f;
            float qw = 1.f, qx = 0.f, qy = 0.f, qz = 0.f;
            int j = 0;
            float Tn = sm.tin[warp][lane];   // transmittance in front of the segment's first hit (S2)
            while (bits != 0u) {
                const int t = warp * 32 + __ffs(bits) - 1;
                bits &= bits - 1;
                if ((uint32_t)(warp * 32 + j) >= stop) break;
                // alpha from the stored 1 - alpha (exact to half an ulp of 1: 3e-8, far inside the 1e-5
                // tolerance of the colour / scale / quaternion maps); T by the same multiply chain as
                // the sequential pass, so it is that pass's T bit for bit
                const float om_alpha = sm.M[lane * kRow + warp * 32 + j].x;
                const float alpha = __fsub_rn(1.0f, om_alpha);
                const float Tb = Tn;
                Tn = __fmul_rn(Tn, om_alpha);
                j++;
                const float4 c = sm.C[t];    // r, g, b, 1 / sin(theta)
                const float4 d = sm.D[t];    // scale xyz, rot.r
                const float4 e = sm.E[t];    // rot xyz, theta
                // blend weight alpha * T, formed once: value * (alpha * T) instead of the reference's
                // (value * alpha) * T differs in the last bit of each term only
                const float wgt = __fmul_rn(alpha, Tb);
                c0 = __fmaf_rn(c.x, wgt, c0);
                c1 = __fmaf_rn(c.y, wgt, c1);
                c2 = __fmaf_rn(c.z, wgt, c2);
                s0 = __fmaf_rn(d.x, wgt, s0);
                s1 = __fmaf_rn(d.y, wgt, s1);
                s2 = __fmaf_rn(d.z, wgt, s2);
                // r = mix(identity, q, w) with q = (w = rot.r, x, y, z), w = alpha * T
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
                    float sn1, cs1;
                    sincos_0_pi(__fmul_rn(wgt, e.w), sn1, cs1);
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
            pp[3][lane] = s0; pp[4][lane] = s1; pp[5][lane] = s2;
            pp[6][lane] = qw; pp[7][lane] = qx; pp[8][lane] = qy; pp[9][lane] = qz;
        }
        __syncthreads();
        PHASE_MARK(3);

        // ---- S4: the segment partials are folded into the pixel state, in list order: warp 0 (which
        // also owns T and the depth sum) folds the quaternion, warp 1 the colour and scale sums.  The
        // other warps already start the next scan / working set (they only write arrays S4 does
        // not read).
        if (warp == 1) {
            for (int w = 0; w < kWarps; w++) {
                if (((seg_mask >> w) & 1u) == 0u) continue;
                float (*pp)[32] = sm.part[w];
                C0 += pp[0][lane]; C1 += pp[1][lane]; C2 += pp[2][lane];
                S0 += pp[3][lane]; S1 += pp[4][lane]; S2 += pp[5][lane];
            }
        }
        if (warp == 0) {
            for (int w = 0; w < kWarps; w++) {
                if (((seg_mask >> w) & 1u) == 0u) continue;
                float (*pp)[32] = sm.part[w];
                const float rw = pp[6][lane], rx = pp[7][lane], ry = pp[8][lane], rz = pp[9][lane];
                const float nw = Qw * rw - Qx * rx - Qy * ry - Qz * rz;
                const float nx = Qw * rx + Qx * rw + Qy * rz - Qz * ry;
                const float ny = Qw * ry + Qy * rw + Qz * rx - Qx * rz;
                const float nz = Qw * rz + Qz * rw + Qx * ry - Qy * rx;
                Qw = nw; Qx = nx; Qy = ny; Qz = nz;
            }
        }
        if (finished) break;
        head += (uint32_t)m;
    }

    // final transmittance to warp 1 (stop_at is free now), which owns the colour sums
    if (warp == 0) sm.stop_at[lane] = __float_as_uint(T);
    __syncthreads();
    if (warp == 1 && inside) {
        const size_t npix = (size_t)p.W * p.H;
        const size_t pid = (size_t)py * p.W + px;
        const float T_fin = __uint_as_float(sm.stop_at[lane]);
        p.out_color[pid] = C0 + T_fin * p.bg[0];
        p.out_color[npix + pid] = C1 + T_fin * p.bg[1];
        p.out_color[2 * npix + pid] = C2 + T_fin * p.bg[2];
        p.out_scale[pid] = S0;
        p.out_scale[npix + pid] = S1;
        p.out_scale[2 * npix + pid] = S2;
    }
    if (warp == 0 && inside) {
        const size_t npix = (size_t)p.W * p.H;
        const size_t pid = (size_t)py * p.W + px;
        p.out_alpha[pid] = 1.0f - T;
        p.n_contrib[pid] = last_contributor;
        p.last_hit[pid] = last_hit;
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
    }
    if (tid == 0) atomicMax(&p.tile_clocks[tile], (uint32_t)((clock64() - t_start) >> 10));
#if VEGS_TILE_DEBUG
    if (tid == 0 && blk == 0)
    {
        for (int k = 0; k < 4; k++) p.tile_dbg[tile * 16 + k] = (uint32_t)(ph_acc[k] >> 10);
        p.tile_dbg[tile * 16 + 8] = dbg_sets; p.tile_dbg[tile * 16 + 9] = dbg_cand;
        p.tile_dbg[tile * 16 + 10] = hit_count; p.tile_dbg[tile * 16 + 11] = dbg_pix;   // pixel hits of warp 0's segments only
    }
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
    p.last_hit = im.last_hit;
    p.hitlist = b.hitlist;
    p.tile_clocks = im.tile_clocks;
    p.tile_dbg = im.tile_dbg;
    const int T = p.gx * ((a.H + kTile - 1) / kTile);
    // opt-in shared memory size is a per-device function attribute
    static DeviceOnce once;
    int dev = 0;
    VEGS_CUDA_OK(cudaGetDevice(&dev));
    if (once.need(dev)) {
        VEGS_CUDA_OK(cudaFuncSetAttribute(composite_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)sizeof(FwdSmem)));
        VEGS_CUDA_OK(cudaFuncSetAttribute(composite_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)sizeof(FwdSmem)));
        once.done(dev);
    }
    VEGS_CUDA_OK(cudaMemsetAsync(im.tile_clocks, 0, (size_t)2 * T * sizeof(uint32_t), stream));
#if VEGS_TILE_DEBUG
    VEGS_CUDA_OK(cudaMemsetAsync(im.tile_dbg, 0, (size_t)16 * T * sizeof(uint32_t), stream));
#endif
    if (list_has_unit_masks(a.P))
        composite_fwd_kernel<true><<<T * kUnitsPerTile, kSuper, sizeof(FwdSmem), stream>>>(p);
    else
        composite_fwd_kernel<false><<<T * kUnitsPerTile, kSuper, sizeof(FwdSmem), stream>>>(p);
    VEGS_LAUNCH_CHECK("composite_fwd", stream, a.debug);
    return VEGS_OK;
}

}  // namespace vegs
