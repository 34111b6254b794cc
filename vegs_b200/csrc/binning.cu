// Tile binning: instance counting, instance emission in depth order, tile ranges.
//
// Replaces cub::DeviceScan::InclusiveSum + duplicateWithKeys + identifyTileRanges
// (reference cuda_rasterizer/rasterizer_impl.cu:70-138, :282-322).
//
// The Gaussians arrive sorted by depth (`order`).  Each block owns a chunk of kEmitChunk
// consecutive sorted positions:
//   tile_count   sums the tiles touched by the chunk                    -> chunksum[chunk]
//   tile_spine   exclusive scan over chunks, total = num_rendered       -> chunksum, *num_rendered
//   emit         scans the chunk locally, then writes the chunk's instances (tile id, Gaussian id):
//                a Gaussian of up to four tiles by the thread that scanned it, a bigger one by a
//                whole warp with one lane per INSTANCE, so big splats do not serialise a thread
//                as the reference's per-Gaussian loop does.
// Instances of one Gaussian are emitted in row-major rectangle order and Gaussians in depth
// order, so a stable sort by tile id alone reproduces the reference's (tile, depth, index) order.
// Roofline: HBM; algorithmic bytes: 12 per sorted Gaussian (index + rectangle) per pass and
// 8 per emitted instance.
#include <cuda_fp16.h>
#include "common.cuh"

namespace vegs {

namespace {

constexpr int kEmitThreads = 256;
constexpr uint32_t kEmitPart = kEmitPartInstances;   // instances emitted by one CTA
constexpr int kEmitPerThread = kEmitChunk / kEmitThreads;  // 4

__device__ __forceinline__ uint32_t rect_tiles(uint2 r) {
    const uint32_t w = (r.x >> 16) - (r.x & 0xffffu);
    const uint32_t h = (r.y >> 16) - (r.y & 0xffffu);
    return w * h;
}

__device__ __forceinline__ uint32_t block_reduce_sum(uint32_t v, uint32_t* s_warp) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = v;
    __syncthreads();
    uint32_t tot = 0;
#pragma unroll
    for (int w = 0; w < kEmitThreads / 32; w++) tot += s_warp[w];
    __syncthreads();
    return tot;
}

__global__ void __launch_bounds__(kEmitThreads)
tile_count_kernel(int P, const uint32_t* __restrict__ order, const uint2* __restrict__ rect,
                  uint32_t* __restrict__ chunksum) {
    __shared__ uint32_t s_warp[kEmitThreads / 32];
    const uint32_t base = blockIdx.x * (uint32_t)kEmitChunk;
    uint32_t sum = 0;
#pragma unroll
    for (int j = 0; j < kEmitPerThread; j++) {
        const uint32_t s = base + j * kEmitThreads + threadIdx.x;
        if (s < (uint32_t)P) sum += rect_tiles(rect[(size_t)kRectStride * order[s]]);
    }
    const uint32_t tot = block_reduce_sum(sum, s_warp);
    if (threadIdx.x == 0) chunksum[blockIdx.x] = tot;
}

// Single block: in-place exclusive scan of chunksum[0..n), chunksum[n] = total = num_rendered.
// Also the emission work list: chunk c is cut into ceil(instances / kEmitPart) parts so that a
// chunk of near, screen-filling Gaussians (millions of instances) is emitted by many CTAs;
// partbase[0..n] = exclusive scan of the part counts, counters[1] = total parts.
__global__ void __launch_bounds__(1024)
tile_spine_kernel(uint32_t* __restrict__ chunksum, uint32_t* __restrict__ partbase, uint32_t n,
                  uint32_t* __restrict__ num_rendered) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    __shared__ uint32_t s_pcarry;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { s_carry = 0; s_pcarry = 0; }
    __syncthreads();
    // pass 1: parts (needs the raw per-chunk counts, so it runs before they are overwritten)
    for (uint32_t base = 0; base < n; base += 1024) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = (i < n) ? (chunksum[i] + kEmitPart - 1) / kEmitPart : 0u;
        uint32_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += t;
            }
            s_warp[lane] = w;
        }
        __syncthreads();
        const uint32_t carry = s_pcarry;
        const uint32_t woff = warp ? s_warp[warp - 1] : 0u;
        if (i < n) partbase[i] = carry + woff + inc - v;
        __syncthreads();
        if (threadIdx.x == 1023) s_pcarry = carry + woff + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        partbase[n] = s_pcarry;
        num_rendered[1] = s_pcarry;
    }
    for (uint32_t base = 0; base < n; base += 1024) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = (i < n) ? chunksum[i] : 0u;
        uint32_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += t;
            }
            s_warp[lane] = w;  // inclusive over warps
        }
        __syncthreads();
        const uint32_t carry = s_carry;
        const uint32_t woff = warp ? s_warp[warp - 1] : 0u;
        if (i < n) chunksum[i] = carry + woff + inc - v;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = carry + woff + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        chunksum[n] = s_carry;
        num_rendered[0] = s_carry;
    }
}

#ifndef VEGS_EMIT_SMALL
#define VEGS_EMIT_SMALL 4
#endif
constexpr int kEmitSmall = VEGS_EMIT_SMALL;   // a Gaussian with at most this many tiles is emitted by its own thread

// One instance: the tile id as the sort key, the Gaussian id as the value.  With unit masks
// (ids < 2^24, common.cuh) the top byte of the value says which of the tile's eight 8x4 compositing
// units the Gaussian can reach: its pixel box (preprocess.cu; r.z, r.w) against 2 unit columns and
// 4 unit rows - the same outcome as unit_box() (composite_common.cuh).
__device__ __forceinline__ void emit_instance(uint32_t out, uint32_t tx, uint32_t ty, uint32_t g, const uint4 r,
                                              int grid_x, uint32_t* __restrict__ tkeys,
                                              uint32_t* __restrict__ tvals, uint32_t R, bool has_masks, int W, int H) {
    if (out >= R) return;
    tkeys[out] = ty * (uint32_t)grid_x + tx;
    uint32_t val = g;
    if (has_masks) {
        const int gx0 = (int)(r.z & 0xffffu), gx1 = (int)(r.z >> 16) - 1;
        const int gy0 = (int)(r.w & 0xffffu), gy1 = (int)(r.w >> 16) - 1;
        uint32_t okx = 0, oky = 0;
#pragma unroll
        for (int hx = 0; hx < 2; hx++) {
            const int bx0 = (int)tx * kTile + hx * 8;
            const int bx1 = min(bx0 + 7, W - 1);
            okx |= ((bx0 < W && max(gx0, bx0) <= min(gx1, bx1)) ? 1u : 0u) << hx;
        }
#pragma unroll
        for (int hy = 0; hy < 4; hy++) {
            const int by0 = (int)ty * kTile + hy * 4;
            const int by1 = min(by0 + 3, H - 1);
            oky |= ((by0 < H && max(gy0, by0) <= min(gy1, by1)) ? 1u : 0u) << hy;
        }
        uint32_t mask = 0;      // bit blk = (row hy) * 2 + (column hx)
#pragma unroll
        for (int hy = 0; hy < 4; hy++)
            if ((oky >> hy) & 1u) mask |= okx << (2 * hy);
        val |= mask << 24;
    }
    tvals[out] = val;
}

__global__ void __launch_bounds__(kEmitThreads)
emit_kernel(int P, const uint32_t* __restrict__ order, const uint2* __restrict__ rect,
            const uint32_t* __restrict__ chunksum, const uint32_t* __restrict__ partbase,
            uint32_t nchunks, int grid_x, uint32_t* __restrict__ tkeys,
            uint32_t* __restrict__ tvals, uint32_t R, const uint2* __restrict__ pbox, int W, int H) {
    __shared__ uint32_t s_g[kEmitChunk];
    __shared__ uint4 s_rect[kEmitChunk];   // tile rect (x, y) and pixel box (z, w) of the chunk's Gaussians
    __shared__ uint32_t s_off[kEmitChunk + 1];
    __shared__ uint32_t s_warp[kEmitThreads / 32];
    __shared__ uint16_t s_big[kEmitChunk];  // chunk positions of the Gaussians emitted by a warp
    __shared__ uint32_t s_nbig;
    const bool has_masks = pbox != nullptr;

    // the grid is an upper bound of the work items (abi.cu): surplus CTAs have nothing to do
    if (blockIdx.x >= partbase[nchunks]) return;
    // which chunk does this part belong to: last c with partbase[c] <= blockIdx.x
    uint32_t clo = 0, chi = nchunks;
    while (chi - clo > 1) {
        const uint32_t mid = (clo + chi) >> 1;
        if (partbase[mid] <= blockIdx.x) clo = mid; else chi = mid;
    }
    const uint32_t chunk = clo;
    const uint32_t part = blockIdx.x - partbase[chunk];
    const uint32_t base = chunk * (uint32_t)kEmitChunk;
    const uint32_t gbase = chunksum[chunk];
    const uint32_t n_inst = chunksum[chunk + 1] - gbase;
    const uint32_t i_begin = part * kEmitPart;
    const uint32_t i_end = min(n_inst, i_begin + kEmitPart);
    if (i_begin >= n_inst) return;
    if (threadIdx.x == 0) s_nbig = 0;

    // Thread t owns kEmitPerThread consecutive sorted positions (position order = depth order).
    uint32_t cnt[kEmitPerThread];
    uint32_t tsum = 0;
#pragma unroll
    for (int j = 0; j < kEmitPerThread; j++) {
        const uint32_t l = threadIdx.x * kEmitPerThread + j;
        const uint32_t s = base + l;
        uint32_t g = 0;
        uint4 r = make_uint4(0u, 0u, 0u, 0u);
        if (s < (uint32_t)P) {
            g = order[s];
            r = reinterpret_cast<const uint4*>(rect)[g];
        }
        s_g[l] = g;
        s_rect[l] = r;
        cnt[j] = rect_tiles(make_uint2(r.x, r.y));
        tsum += cnt[j];
    }
    // exclusive scan of the per-thread sums over the block
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = tsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t woff = 0;
#pragma unroll
    for (int w = 0; w < kEmitThreads / 32; w++)
        if (w < warp) woff += s_warp[w];
    uint32_t run = woff + inc - tsum;
#pragma unroll
    for (int j = 0; j < kEmitPerThread; j++) {
        s_off[threadIdx.x * kEmitPerThread + j] = run;
        run += cnt[j];
    }
    if (threadIdx.x == kEmitThreads - 1) s_off[kEmitChunk] = run;
    __syncthreads();

    // Emission.  Most Gaussians at LiDAR density touch one to four tiles: those are written by the
    // thread that already holds their count and offset (no search, no division; a warp's stores
    // land in one short run of the output).  A Gaussian with more tiles goes on a work list and is
    // emitted by a whole warp, one lane per instance, so a big splat does not serialise a thread.
    // Either way only the instances inside this part's range [i_begin, i_end) are written.
    {
        uint32_t off = woff + inc - tsum;
#pragma unroll
        for (int j = 0; j < kEmitPerThread; j++) {
            const uint32_t l = threadIdx.x * kEmitPerThread + j;
            if (cnt[j] > (uint32_t)kEmitSmall) {
                if (off < i_end && off + cnt[j] > i_begin) s_big[atomicAdd(&s_nbig, 1u)] = (uint16_t)l;
            } else if (cnt[j] != 0u) {
                const uint4 r = s_rect[l];
                const uint32_t x0 = r.x & 0xffffu, x1 = r.x >> 16;
                uint32_t tx = x0, ty = r.y & 0xffffu;
                for (uint32_t k = 0; k < cnt[j]; k++) {
                    const uint32_t i = off + k;
                    if (i >= i_begin && i < i_end)
                        emit_instance(gbase + i, tx, ty, s_g[l], r, grid_x, tkeys, tvals, R, has_masks, W, H);
                    if (++tx == x1) { tx = x0; ty++; }
                }
            }
            off += cnt[j];
        }
    }
    __syncthreads();
    const uint32_t nbig = s_nbig;
    for (uint32_t b = warp; b < nbig; b += kEmitThreads / 32) {
        const uint32_t l = s_big[b];
        const uint4 r = s_rect[l];
        const uint32_t g = s_g[l];
        const uint32_t first = s_off[l], n_l = s_off[l + 1] - first;
        const uint32_t x0 = r.x & 0xffffu, w = (r.x >> 16) - x0, y0 = r.y & 0xffffu;
        const uint32_t k_lo = i_begin > first ? i_begin - first : 0u;
        const uint32_t k_hi = min(n_l, i_end - first);          // first < i_end for every listed Gaussian
        for (uint32_t k = k_lo + lane; k < k_hi; k += 32) {
            const uint32_t row = k / w;
            emit_instance(gbase + first + k, x0 + (k - row * w), y0 + row, g, r, grid_x, tkeys, tvals, R, has_masks,
                          W, H);
        }
    }
}

// ranges[tile] = [first, last) positions of the tile in the sorted instance list.
__global__ void __launch_bounds__(256)
tile_ranges_kernel(uint32_t R, const uint32_t* __restrict__ tkeys, uint2* __restrict__ ranges) {
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    if (i >= R) return;
    const uint32_t cur = tkeys[i];
    if (i == 0) {
        ranges[cur].x = 0;
    } else {
        const uint32_t prev = tkeys[i - 1];
        if (cur != prev) {
            ranges[prev].y = i;
            ranges[cur].x = i;
        }
    }
    if (i == R - 1) ranges[cur].y = R;
}

// Tiles ordered by list length, longest first (33 log2 buckets, order inside a bucket arbitrary).
// The compositing kernels take their tile from this table so the few very long lists of a
// LiDAR-density scene start first instead of becoming the tail of the launch.
__global__ void __launch_bounds__(1024)
tile_order_kernel(const uint2* __restrict__ ranges, uint32_t T, uint32_t* __restrict__ order) {
    __shared__ uint32_t s_cnt[33];
    __shared__ uint32_t s_cur[33];
    if (threadIdx.x < 33) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    for (uint32_t t = threadIdx.x; t < T; t += 1024) {
        const uint2 r = ranges[t];
        const uint32_t len = r.y - r.x;
        const uint32_t key = len ? (uint32_t)__clz(len) : 32u;   // small key = long list
        atomicAdd(&s_cnt[key], 1u);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t run = 0;
        for (int k = 0; k < 33; k++) { s_cur[k] = run; run += s_cnt[k]; }
    }
    __syncthreads();
    for (uint32_t t = threadIdx.x; t < T; t += 1024) {
        const uint2 r = ranges[t];
        const uint32_t len = r.y - r.x;
        const uint32_t key = len ? (uint32_t)__clz(len) : 32u;
        order[atomicAdd(&s_cur[key], 1u)] = t;
    }
}

}  // namespace

int launch_tile_count(int P, const uint32_t* order, const uint2* rect, uint32_t* chunksum,
                      uint32_t* partbase, uint32_t* num_rendered_dev, cudaStream_t stream, bool debug) {
    const uint32_t nchunks = div_up(P, kEmitChunk);
    tile_count_kernel<<<nchunks, kEmitThreads, 0, stream>>>(P, order, rect, chunksum);
    VEGS_LAUNCH_CHECK("tile_count", stream, debug);
    tile_spine_kernel<<<1, 1024, 0, stream>>>(chunksum, partbase, nchunks, num_rendered_dev);
    VEGS_LAUNCH_CHECK("tile_spine", stream, debug);
    return VEGS_OK;
}

int launch_emit(int P, const uint32_t* order, const uint2* rect, const uint32_t* chunksum,
                const uint32_t* partbase, uint32_t nparts, int grid_x, uint32_t* tkeys,
                uint32_t* tvals, uint32_t R, const uint2* pbox, int W, int H, cudaStream_t stream,
                bool debug) {
    const uint32_t nchunks = div_up(P, kEmitChunk);
    if (nparts == 0) return VEGS_OK;
    emit_kernel<<<nparts, kEmitThreads, 0, stream>>>(P, order, rect, chunksum, partbase, nchunks,
                                                     grid_x, tkeys, tvals, R,
                                                     list_has_unit_masks(P) ? pbox : nullptr, W, H);
    VEGS_LAUNCH_CHECK("emit", stream, debug);
    return VEGS_OK;
}

int launch_tile_ranges(uint32_t R, const uint32_t* tkeys_sorted, uint2* ranges,
                       uint32_t* tile_order, int T, cudaStream_t stream, bool debug) {
    VEGS_CUDA_OK(cudaMemsetAsync(ranges, 0, (size_t)T * sizeof(uint2), stream));
    prof_mark("memset_ranges", stream);
    if (R > 0) {
        tile_ranges_kernel<<<div_up(R, 256), 256, 0, stream>>>(R, tkeys_sorted, ranges);
        VEGS_LAUNCH_CHECK("tile_ranges", stream, debug);
    }
    tile_order_kernel<<<1, 1024, 0, stream>>>(ranges, (uint32_t)T, tile_order);
    VEGS_LAUNCH_CHECK("tile_order", stream, debug);
    return VEGS_OK;
}

}  // namespace vegs
