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
// Two ways to run a pass; measured on the 6 M scene (profiles/r02d_bench*.json): the three-kernel
// pass below is the default - all radix launches of a view take 0.53 ms with it against 0.62 ms
// with the look-back scatter (which saves the histogram and row-scan launches, 0.09 ms, but whose
// scatter launches grow from 0.38 to 0.55 ms: the chained waits of 1 500 - 3 100 blocks on their
// predecessors' published counts cost more than the extra read of the keys).
// VEGS_SORT_LOOKBACK=1 selects: one histogram launch + one launch per pass ("onesweep" structure):
//   radix_hist_all global digit histograms of ALL passes in one read of the keys
//   radix_scatter  (one per pass) stable in-block ranking (warp ballots + per-warp counters),
//                  staging of the block's 4096 items in digit order in shared memory, coalesced
//                  scatter.  Where the block's items start inside each digit's output range - the
//                  number of items with that digit in all EARLIER blocks - is found by a decoupled
//                  look-back: blocks take their index from an atomic counter (so every block only
//                  waits for blocks that are already running), publish their 256 digit counts
//                  (flag "aggregate"), walk back over their predecessors' published words until
//                  they meet an "inclusive" one, and publish their own inclusive prefix.  Value
//                  and flag share one 32-bit word, so no fence is needed.
// Default: three kernels per pass -
//   radix_hist     per-block digit histogram, stored digit-major  hist[d * nblocks + b]
//   radix_rowscan  one block per digit: exclusive scan of its row + row total
//   radix_scatter  as above with the offsets read from the scanned histogram.
// Roofline: HBM; algorithmic bytes per item per pass: 8 (read) + 8 (write), + 4 once per sort.
#include <cstdlib>
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

// Global digit histograms of every pass of a sort in one read of the keys: ghist[pass * 256 + d].
__global__ void __launch_bounds__(kSortThreads)
radix_hist_all_kernel(const uint32_t* __restrict__ keys, uint32_t n, int passes, uint32_t* __restrict__ ghist) {
    __shared__ uint32_t s_hist[kSortMaxPasses * kRadix];
    for (int i = threadIdx.x; i < kSortMaxPasses * kRadix; i += kSortThreads) s_hist[i] = 0;
    __syncthreads();
    for (uint32_t i = blockIdx.x * (uint32_t)kSortThreads + threadIdx.x; i < n; i += gridDim.x * (uint32_t)kSortThreads) {
        const uint32_t k = keys[i];
#pragma unroll
        for (int p = 0; p < kSortMaxPasses; p++)
            if (p < passes) atomicAdd(&s_hist[p * kRadix + ((k >> (p * kRadixBits)) & (kRadix - 1))], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * kRadix; i += kSortThreads)
        if (s_hist[i] != 0u) atomicAdd(&ghist[i], s_hist[i]);
}

constexpr uint32_t kFlagAggregate = 1u << 30, kFlagInclusive = 2u << 30, kStatusValue = (1u << 30) - 1u;

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

// kLookback: offsets by decoupled look-back (hist = this pass's status rows [nblocks][256], rowtot =
// this pass's global digit histogram, counter = its dynamic block counter; all zero at launch);
// otherwise from the scanned block histograms (hist digit-major, rowtot = digit totals).
template <bool kImplicitVals, bool kWriteKeys, bool kLookback>
__global__ void __launch_bounds__(kSortThreads, 4)
radix_scatter_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                     uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, uint32_t n,
                     int shift, uint32_t* hist, const uint32_t* __restrict__ rowtot, uint32_t nblocks,
                     uint32_t* counter) {
    constexpr int kWarps = kSortThreads / 32;
    constexpr int kWarpItems = kSortTile / kWarps;  // 512 consecutive items per warp
    __shared__ uint32_t s_keys[kSortTile];
    __shared__ uint32_t s_vals[kSortTile];
    __shared__ uint32_t s_whist[kWarps][kRadix];
    __shared__ uint32_t s_dbase[kRadix];  // block-local first slot of digit d
    __shared__ uint32_t s_gbase[kRadix];  // global slot of local slot 0 of digit d (minus dbase)
    __shared__ uint32_t s_warp[kWarps];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // look-back needs every predecessor to be running already: blocks number themselves in the
    // order in which they start
    if (kLookback) {
        if (tid == 0) s_warp[0] = atomicAdd(counter, 1u);
    }
    for (int i = tid; i < kWarps * kRadix; i += kSortThreads) (&s_whist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t bid = kLookback ? s_warp[0] : blockIdx.x;
    const uint32_t base = bid * (uint32_t)kSortTile;
    const uint32_t count = min((uint32_t)kSortTile, n - base);
    __syncthreads();          // s_warp is reused by the scans below

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
        uint32_t before;      // items with digit d in the blocks in front of this one
        if (kLookback) {
            volatile uint32_t* st = hist;
            before = 0;
            if (bid == 0) {
                st[d] = kFlagInclusive | run;
            } else {
                st[(size_t)bid * kRadix + d] = kFlagAggregate | run;
                uint32_t j = bid - 1;
                for (;;) {
                    const uint32_t v = st[(size_t)j * kRadix + d];
                    const uint32_t flag = v & ~kStatusValue;
                    if (flag == 0u) continue;                 // not published yet: look again
                    before += v & kStatusValue;
                    if (flag == kFlagInclusive) break;
                    j--;                                      // block 0 always publishes "inclusive"
                }
                st[(size_t)bid * kRadix + d] = kFlagInclusive | (before + run);
            }
        } else {
            before = hist[(size_t)d * nblocks + bid];
        }
        s_gbase[d] = gdigit + before - dbase;
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

static bool sort_lookback_enabled() {
    static const bool on = [] { const char* e = getenv("VEGS_SORT_LOOKBACK"); return e != nullptr && e[0] == '1'; }();
    return on;
}

template <bool kLookback>
static void launch_scatter(bool first_implicit, bool write_keys, uint32_t nblocks, cudaStream_t stream,
                           const uint32_t* kin, const uint32_t* vin, uint32_t* kout, uint32_t* vout, uint32_t n,
                           int shift, uint32_t* hist, const uint32_t* rowtot, uint32_t* counter) {
    if (first_implicit)
        radix_scatter_kernel<true, true, kLookback><<<nblocks, kSortThreads, 0, stream>>>(kin, vin, kout, vout, n, shift, hist, rowtot, nblocks, counter);
    else if (write_keys)
        radix_scatter_kernel<false, true, kLookback><<<nblocks, kSortThreads, 0, stream>>>(kin, vin, kout, vout, n, shift, hist, rowtot, nblocks, counter);
    else
        radix_scatter_kernel<false, false, kLookback><<<nblocks, kSortThreads, 0, stream>>>(kin, vin, kout, vout, n, shift, hist, rowtot, nblocks, counter);
}

int radix_sort_pairs(uint32_t* k0, uint32_t* v0, uint32_t* k1, uint32_t* v1, bool implicit_vals,
                     uint32_t n, int nbits, uint32_t* hist, uint32_t* rowtot, bool write_last_keys,
                     bool* result_in_second, cudaStream_t stream, bool debug) {
    *result_in_second = false;
    if (n == 0 || nbits <= 0) return VEGS_OK;
    const uint32_t nblocks = div_up(n, kSortTile);
    const int passes = (nbits + kRadixBits - 1) / kRadixBits;
    if (passes > kSortMaxPasses || n >= (1u << 30)) {
        set_error("radix sort: at most %d passes and 2^30 items", kSortMaxPasses);
        return VEGS_EINVAL;
    }
    {
        // four 43 KB blocks per SM only fit with the shared-memory carve-out at its maximum
        static DeviceOnce once;
        int dev = 0;
        VEGS_CUDA_OK(cudaGetDevice(&dev));
        if (once.need(dev)) {
            VEGS_CUDA_OK(cudaFuncSetAttribute(radix_scatter_kernel<true, true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
            VEGS_CUDA_OK(cudaFuncSetAttribute(radix_scatter_kernel<false, true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
            VEGS_CUDA_OK(cudaFuncSetAttribute(radix_scatter_kernel<false, false, true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
            VEGS_CUDA_OK(cudaFuncSetAttribute(radix_scatter_kernel<true, true, false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
            VEGS_CUDA_OK(cudaFuncSetAttribute(radix_scatter_kernel<false, true, false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
            VEGS_CUDA_OK(cudaFuncSetAttribute(radix_scatter_kernel<false, false, false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
            once.done(dev);
        }
    }
    const bool lookback = sort_lookback_enabled();
    // layout of `hist` (sort_hist_words): status rows of every pass, global histograms, counters
    const size_t status_words = (size_t)kRadix * nblocks;
    uint32_t* ghist = hist + status_words * kSortMaxPasses;
    uint32_t* counters = ghist + (size_t)kSortMaxPasses * kRadix;
    uint32_t *kin = k0, *vin = v0, *kout = k1, *vout = v1;
    if (lookback) {
        VEGS_CUDA_OK(cudaMemsetAsync(hist, 0, sort_hist_words(n) * sizeof(uint32_t), stream));
        prof_mark("memset_sort", stream);
        const uint32_t hb = nblocks < 148u * 8u ? nblocks : 148u * 8u;
        radix_hist_all_kernel<<<hb, kSortThreads, 0, stream>>>(kin, n, passes, ghist);
        VEGS_LAUNCH_CHECK("radix_hist", stream, debug);
    }
    for (int p = 0; p < passes; p++) {
        const int shift = p * kRadixBits;
        const bool first_implicit = implicit_vals && p == 0;
        const bool write_keys = write_last_keys || p + 1 < passes;
        if (lookback) {
            launch_scatter<true>(first_implicit, write_keys, nblocks, stream, kin, vin, kout, vout, n, shift,
                                 hist + status_words * p, ghist + (size_t)p * kRadix, counters + p);
        } else {
            radix_hist_kernel<<<nblocks, kSortThreads, 0, stream>>>(kin, n, shift, hist, nblocks);
            VEGS_LAUNCH_CHECK("radix_hist", stream, debug);
            radix_rowscan_kernel<<<kRadix, kSortThreads, 0, stream>>>(hist, nblocks, rowtot);
            VEGS_LAUNCH_CHECK("radix_rowscan", stream, debug);
            launch_scatter<false>(first_implicit, write_keys, nblocks, stream, kin, vin, kout, vout, n, shift, hist,
                                  rowtot, nullptr);
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
