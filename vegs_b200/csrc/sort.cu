// Stable LSD radix sort of (uint32 key, uint32 value) pairs, 8 bits per pass.
//
// Replaces the reference's cub::DeviceRadixSort::SortPairs on 64-bit (tile|depth) keys
// (cuda_rasterizer/rasterizer_impl.cu:305-313) and on Morton codes (simple-knn/simple_knn.cu:210-213).
// The rasterizer needs the order (tile, depth, Gaussian index).  Because an LSD radix sort is
// stable, that order is built in two cheaper steps instead of one 44-bit sort over R instances:
//   1. sort the P Gaussians once by their 32-bit depth key (4 passes over P items),
//   2. emit tile instances in that order and sort them by tile id only (ceil(log2 T / 8) passes
//      over R items of 8 bytes instead of 6 passes over R items of 12 bytes).
//
// One pass = three kernels:
//   radix_hist     per-block digit histogram, stored digit-major  hist[d * nblocks + b]
//   radix_rowscan  one block per digit: exclusive scan of its row + row total
//   radix_scatter  stable in-block ranking (warp ballots + per-warp counters), staging of the
//                  block's 4096 items in digit order in shared memory, coalesced scatter
// Roofline: HBM; algorithmic bytes per item per pass: 4 (hist) + 8 (read) + 8 (write).
#include "common.cuh"

namespace vegs {

namespace {

__global__ void __launch_bounds__(kSortThreads)
radix_hist_kernel(const uint32_t* __restrict__ keys, uint32_t n, int shift,
                  uint32_t* __restrict__ hist, uint32_t nblocks) {
    __shared__ uint32_t s_hist[kRadix];
    s_hist[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * (uint32_t)kSortTile;
    const uint32_t count = min((uint32_t)kSortTile, n - base);
#pragma unroll 4
    for (uint32_t i = threadIdx.x; i < count; i += kSortThreads) {
        const uint32_t d = (keys[base + i] >> shift) & (kRadix - 1);
        atomicAdd(&s_hist[d], 1u);
    }
    __syncthreads();
    hist[(size_t)threadIdx.x * nblocks + blockIdx.x] = s_hist[threadIdx.x];
}

// Exclusive scan of 256 values held one per thread (blockDim.x == 256).
__device__ __forceinline__ uint32_t block_exclusive_scan_256(uint32_t v, uint32_t* s_warp,
                                                             uint32_t* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t woff = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < kSortThreads / 32; w++) {
        const uint32_t t = s_warp[w];
        if (w < warp) woff += t;
        tot += t;
    }
    __syncthreads();
    if (total) *total = tot;
    return woff + inc - v;
}

__global__ void __launch_bounds__(kSortThreads)
radix_rowscan_kernel(uint32_t* __restrict__ hist, uint32_t nblocks,
                     uint32_t* __restrict__ rowtot) {
    __shared__ uint32_t s_warp[kSortThreads / 32];
    uint32_t* row = hist + (size_t)blockIdx.x * nblocks;
    uint32_t carry = 0;
    for (uint32_t base = 0; base < nblocks; base += kSortThreads) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = (i < nblocks) ? row[i] : 0u;
        uint32_t tot;
        const uint32_t ex = block_exclusive_scan_256(v, s_warp, &tot);
        if (i < nblocks) row[i] = carry + ex;
        carry += tot;
    }
    if (threadIdx.x == 0) rowtot[blockIdx.x] = carry;
}

template <bool kImplicitVals, bool kWriteKeys>
__global__ void __launch_bounds__(kSortThreads, 4)
radix_scatter_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                     uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, uint32_t n,
                     int shift, const uint32_t* __restrict__ hist,
                     const uint32_t* __restrict__ rowtot, uint32_t nblocks) {
    constexpr int kWarps = kSortThreads / 32;
    constexpr int kWarpItems = kSortTile / kWarps;  // 512 consecutive items per warp
    __shared__ uint32_t s_keys[kSortTile];
    __shared__ uint32_t s_vals[kSortTile];
    __shared__ uint32_t s_whist[kWarps][kRadix];
    __shared__ uint32_t s_dbase[kRadix];  // block-local first slot of digit d
    __shared__ uint32_t s_gbase[kRadix];  // global slot of local slot 0 of digit d (minus dbase)
    __shared__ uint32_t s_warp[kWarps];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t base = blockIdx.x * (uint32_t)kSortTile;
    const uint32_t count = min((uint32_t)kSortTile, n - base);

    for (int i = tid; i < kWarps * kRadix; i += kSortThreads) (&s_whist[0][0])[i] = 0;
    __syncthreads();

    // ---- phase 1: rank every item inside (warp, digit), in item order
    uint32_t key[kSortItemsPerThread];
    uint16_t rank[kSortItemsPerThread];
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < kSortItemsPerThread; r++) {
        const uint32_t li = warp * kWarpItems + r * 32 + lane;
        const bool valid = li < count;
        const uint32_t k = valid ? keys_in[base + li] : 0u;
        key[r] = k;
        const uint32_t d = (k >> shift) & (kRadix - 1);
        // lanes holding the same digit, by eight ballots (one per digit bit): measured faster than
        // match.any, whose cost grows with the number of distinct values - nearly 32 here
        uint32_t peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
        for (int b = 0; b < kRadixBits; b++) {
            const bool bit = (d >> b) & 1u;
            const uint32_t vote = __ballot_sync(0xffffffffu, bit);
            peers &= bit ? vote : ~vote;
        }
        const uint32_t before = __popc(peers & lt_mask);
        // the first lane of each group bumps the warp's digit counter and hands the old value round
        const int leader = __ffs(peers) - 1;
        uint32_t prev = 0;
        if (valid && lane == leader) prev = atomicAdd(&s_whist[warp][d], (uint32_t)__popc(peers));
        prev = __shfl_sync(0xffffffffu, prev, leader & 31);
        rank[r] = (uint16_t)(prev + before);
    }
    __syncthreads();

    // ---- phase 2: per digit, exclusive prefix over warps; digit bases (local and global)
    {
        const int d = tid;
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < kWarps; w++) {
            const uint32_t t = s_whist[w][d];
            s_whist[w][d] = run;
            run += t;
        }
        const uint32_t dbase = block_exclusive_scan_256(run, s_warp, nullptr);
        const uint32_t gdigit = block_exclusive_scan_256(rowtot[d], s_warp, nullptr);
        s_dbase[d] = dbase;
        s_gbase[d] = gdigit + hist[(size_t)d * nblocks + blockIdx.x] - dbase;
    }
    __syncthreads();

    // ---- phase 3: place the block's items in digit order in shared memory
#pragma unroll
    for (int r = 0; r < kSortItemsPerThread; r++) {
        const uint32_t li = warp * kWarpItems + r * 32 + lane;
        if (li < count) {
            const uint32_t d = (key[r] >> shift) & (kRadix - 1);
            const uint32_t pos = s_dbase[d] + s_whist[warp][d] + rank[r];
            s_keys[pos] = key[r];
            s_vals[pos] = kImplicitVals ? (base + li) : vals_in[base + li];
        }
    }
    __syncthreads();

    // ---- phase 4: coalesced runs to global memory
    for (uint32_t i = tid; i < count; i += kSortThreads) {
        const uint32_t k = s_keys[i];
        const uint32_t d = (k >> shift) & (kRadix - 1);
        const uint32_t g = s_gbase[d] + i;
        if (kWriteKeys) keys_out[g] = k;
        vals_out[g] = s_vals[i];
    }
}

}  // namespace

int radix_sort_pairs(uint32_t* k0, uint32_t* v0, uint32_t* k1, uint32_t* v1, bool implicit_vals,
                     uint32_t n, int nbits, uint32_t* hist, uint32_t* rowtot, bool write_last_keys,
                     bool* result_in_second, cudaStream_t stream, bool debug) {
    *result_in_second = false;
    if (n == 0 || nbits <= 0) return VEGS_OK;
    const uint32_t nblocks = div_up(n, kSortTile);
    const int passes = (nbits + kRadixBits - 1) / kRadixBits;
    {
        // four 43 KB blocks per SM only fit with the shared-memory carve-out at its maximum
        static DeviceOnce once;
        int dev = 0;
        VEGS_CUDA_OK(cudaGetDevice(&dev));
        if (once.need(dev)) {
            VEGS_CUDA_OK(cudaFuncSetAttribute(radix_scatter_kernel<true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
            VEGS_CUDA_OK(cudaFuncSetAttribute(radix_scatter_kernel<false, true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
            VEGS_CUDA_OK(cudaFuncSetAttribute(radix_scatter_kernel<false, false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
            once.done(dev);
        }
    }
    uint32_t *kin = k0, *vin = v0, *kout = k1, *vout = v1;
    for (int p = 0; p < passes; p++) {
        const int shift = p * kRadixBits;
        const bool first_implicit = implicit_vals && p == 0;
        const bool write_keys = write_last_keys || p + 1 < passes;
        radix_hist_kernel<<<nblocks, kSortThreads, 0, stream>>>(kin, n, shift, hist, nblocks);
        VEGS_LAUNCH_CHECK("radix_hist", stream, debug);
        radix_rowscan_kernel<<<kRadix, kSortThreads, 0, stream>>>(hist, nblocks, rowtot);
        VEGS_LAUNCH_CHECK("radix_rowscan", stream, debug);
        if (first_implicit) {
            radix_scatter_kernel<true, true><<<nblocks, kSortThreads, 0, stream>>>(
                kin, vin, kout, vout, n, shift, hist, rowtot, nblocks);
        } else if (write_keys) {
            radix_scatter_kernel<false, true><<<nblocks, kSortThreads, 0, stream>>>(
                kin, vin, kout, vout, n, shift, hist, rowtot, nblocks);
        } else {
            radix_scatter_kernel<false, false><<<nblocks, kSortThreads, 0, stream>>>(
                kin, vin, kout, vout, n, shift, hist, rowtot, nblocks);
        }
        VEGS_LAUNCH_CHECK("radix_scatter", stream, debug);
        uint32_t* t;
        t = kin; kin = kout; kout = t;
        t = vin; vin = vout; vout = t;
    }
    *result_in_second = (passes & 1) != 0;
    return VEGS_OK;
}

}  // namespace vegs
