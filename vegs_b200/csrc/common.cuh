// Shared declarations for libvegs_b200.so (sm_100a only).
//
// Data layout in HBM (all sub-arrays 256-byte aligned inside caller-provided byte buffers):
//
//   geom    (kept for backward)  recA,recB,recC : float4, interleaved per Gaussian (64-byte stride)
//                                clamped        : uint8[P]    SH clamp bits (bit c = channel c)
//                                counters       : uint32[64]  [0] = num_rendered
//   scratch (forward only)       recD,recE      : float4, interleaved per Gaussian (32-byte stride)
//                                rect           : uint2[P]    tile rectangle (min | max<<16)
//                                keys0/1,vals0/1: uint32[P]   depth-sort ping-pong
//                                hist, rowtot, chunk sums
//   img     (kept for backward)  n_contrib      : uint32[W*H]
//                                ranges         : uint2[T]
//                                tile_order     : uint32[T]   tiles, longest list first
//   binning (kept for backward)  point_list     : uint32[R]
//   bin_scratch (forward only)   tile keys ping-pong, value ping, hist
//
// Record contents (only written for Gaussians with radii > 0):
//   recA = { mean2D.x, mean2D.y, half2(ex, ey), depth(view z) }   ex, ey: half extents of the box
//                                                           outside which alpha < 1/255 (preprocess.cu);
//                                                           all a compositing unit needs to reject an entry
//   recB = { conic.a, conic.b, conic.c, opacity }
//   recC = { rgb.r, rgb.g, rgb.b, 1/sin(theta) }            theta = acos(rot.r)
//   recD = { scale.x, scale.y, scale.z, rot.r }            raw kernel inputs (no modifier)
//   recE = { rot.x, rot.y, rot.z, theta }
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <atomic>
#include "../../include/vegs_b200.h"

namespace vegs {

constexpr int kTile = VEGS_TILE;
constexpr int kTilePix = kTile * kTile;

// ---------------------------------------------------------------- radix sort configuration
constexpr int kSortThreads = 256;
constexpr int kSortItemsPerThread = 16;
constexpr int kSortTile = kSortThreads * kSortItemsPerThread;  // 4096 items per block
constexpr int kRadixBits = 8;
constexpr int kRadix = 1 << kRadixBits;
constexpr int kSortMaxPasses = 4;           // 32-bit keys
// Words of the sort's `hist` scratch for n items: per pass one 256-word status row per 4096-item
// block (decoupled look-back; the three-kernel fallback uses the first pass's part as its block
// histograms), then the global digit histograms of all passes and the dynamic block counters.
static inline size_t sort_hist_words(size_t n) {
    const size_t nblocks = (n + (size_t)kSortTile - 1) / (size_t)kSortTile;
    return (size_t)kRadix * nblocks * kSortMaxPasses + (size_t)kSortMaxPasses * kRadix + 64;
}

// ---------------------------------------------------------------- emission configuration
constexpr int kUnitsPerTile = 8;  // compositing units (8x4 pixel blocks) per 16x16 tile
// Packed records are interleaved per Gaussian so that a gather touches whole 32-byte sectors:
// geom holds {recA, recB, recC, pad} (64 B), scratch holds {recD, recE} (32 B).  recX[g] lives at
// recX + stride * g with these strides (in float4 units).
constexpr int kGeomRecStride = 4;
constexpr int kScratchRecStride = 2;
constexpr int kRectStride = 2;    // {tile rect, pixel box} per Gaussian, in uint2 units
// Instance list words: Gaussian id, plus - when every id fits 24 bits - in the top byte the mask of
// the tile's compositing units the Gaussian's conservative box reaches (written by the emission
// kernel), so a unit rejects an entry from the list word alone.
constexpr uint32_t kListIdMask = 0x00ffffffu;
bool sh_bulk_disabled();             // abi.cu: VEGS_SH_BULK=0 in the environment selects the 16-byte-load paths (A/B measurements)
bool list_has_unit_masks(int P);   // abi.cu: P < 2^24 unless disabled through vegs_debug_unit_masks
constexpr int kEmitChunk = 1024;  // depth-sorted Gaussians per emission block
constexpr uint32_t kEmitPartInstances = 8192;   // instances emitted by one CTA

static inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }
static inline uint32_t div_up(uint64_t a, uint64_t b) { return (uint32_t)((a + b - 1) / b); }

struct GeomView {
    float4* recA;
    float4* recB;
    float4* recC;
    uint8_t* clamped;
    uint32_t* counters;
    size_t bytes;
};

struct ScratchView {
    float4* recD;
    float4* recE;
    uint2* rect;
    uint2* pbox;         // pixel box of the Gaussian clipped to the image: (x0 | (x1+1) << 16, y0 | (y1+1) << 16)
    uint32_t* keys0;
    uint32_t* keys1;
    uint32_t* vals0;
    uint32_t* vals1;
    uint32_t* hist;      // kRadix * nblocks(P)
    uint32_t* rowtot;    // kRadix
    uint32_t* chunksum;  // nchunks(P) + 1
    uint32_t* partbase;  // nchunks(P) + 1
    size_t bytes;
};

struct ImgView {
    uint32_t* n_contrib;
    uint32_t* last_hit;      // [W*H] 1 + index, in the pixel's unit hit list, of its last contributor
    uint2* ranges;
    uint32_t* tile_order;
    uint32_t* tile_clocks;   // [2*T] SM cycles/1024 spent on each tile by composite fwd / bwd (diagnostics)
    uint32_t* tile_dbg;      // [16*T] extra counters, only written when built with -DVEGS_TILE_DEBUG=1
    size_t bytes;
};

struct BinView {
    uint32_t* point_list;
    uint2* hitlist;          // [8*R] per compositing unit: (Gaussian id, pixel hit mask) of the entries
                             // that reached one of its pixels, in list order; unit (tile, blk) owns
                             // [8*first + blk*len, +len) with (first, len) the tile's range
    size_t bytes;
};

struct BinScratchView {
    uint32_t* tkeys0;
    uint32_t* tkeys1;
    uint32_t* tvals1;
    uint32_t* hist;
    uint32_t* rowtot;
    size_t bytes;
};

template <typename T>
static inline T* carve(char*& cur, size_t count) {
    T* p = reinterpret_cast<T*>(cur);
    cur += align_up(count * sizeof(T));
    return p;
}

static inline GeomView geom_view(void* base, int P) {
    GeomView v;
    char* cur = reinterpret_cast<char*>(base);
    size_t n = (size_t)(P > 0 ? P : 1);
    v.recA = carve<float4>(cur, (size_t)kGeomRecStride * n);   // interleaved {A, B, C, pad}
    v.recB = v.recA + 1;
    v.recC = v.recA + 2;
    v.clamped = carve<uint8_t>(cur, n);
    v.counters = carve<uint32_t>(cur, 64);
    v.bytes = (size_t)(cur - reinterpret_cast<char*>(base));
    return v;
}

static inline ScratchView scratch_view(void* base, int P) {
    ScratchView v;
    char* cur = reinterpret_cast<char*>(base);
    size_t n = (size_t)(P > 0 ? P : 1);
    v.recD = carve<float4>(cur, (size_t)kScratchRecStride * n);   // interleaved {D, E}
    v.recE = v.recD + 1;
    v.rect = carve<uint2>(cur, (size_t)kRectStride * n);   // interleaved {rect, pbox}: one 16-byte record
    v.pbox = v.rect + 1;
    v.keys0 = carve<uint32_t>(cur, n);
    v.keys1 = carve<uint32_t>(cur, n);
    v.vals0 = carve<uint32_t>(cur, n);
    v.vals1 = carve<uint32_t>(cur, n);
    v.hist = carve<uint32_t>(cur, sort_hist_words(n));
    v.rowtot = carve<uint32_t>(cur, kRadix);
    v.chunksum = carve<uint32_t>(cur, (size_t)div_up(n, kEmitChunk) + 1);
    v.partbase = carve<uint32_t>(cur, (size_t)div_up(n, kEmitChunk) + 1);
    v.bytes = (size_t)(cur - reinterpret_cast<char*>(base));
    return v;
}

static inline ImgView img_view(void* base, int W, int H) {
    ImgView v;
    char* cur = reinterpret_cast<char*>(base);
    size_t npix = (size_t)W * H;
    size_t T = (size_t)((W + kTile - 1) / kTile) * ((H + kTile - 1) / kTile);
    v.n_contrib = carve<uint32_t>(cur, npix ? npix : 1);
    v.last_hit = carve<uint32_t>(cur, npix ? npix : 1);
    v.ranges = carve<uint2>(cur, T ? T : 1);
    v.tile_order = carve<uint32_t>(cur, T ? T : 1);
    v.tile_clocks = carve<uint32_t>(cur, 2 * (T ? T : 1));
    v.tile_dbg = carve<uint32_t>(cur, 16 * (T ? T : 1));
    v.bytes = (size_t)(cur - reinterpret_cast<char*>(base));
    return v;
}

static inline BinView bin_view(void* base, int64_t R) {
    BinView v;
    char* cur = reinterpret_cast<char*>(base);
    size_t n = (size_t)(R > 0 ? R : 1);
    v.point_list = carve<uint32_t>(cur, n);
    v.hitlist = carve<uint2>(cur, (size_t)kUnitsPerTile * n);
    v.bytes = (size_t)(cur - reinterpret_cast<char*>(base));
    return v;
}

static inline BinScratchView bin_scratch_view(void* base, int64_t R) {
    BinScratchView v;
    char* cur = reinterpret_cast<char*>(base);
    size_t n = (size_t)(R > 0 ? R : 1);
    v.tkeys0 = carve<uint32_t>(cur, n);
    v.tkeys1 = carve<uint32_t>(cur, n);
    v.tvals1 = carve<uint32_t>(cur, n);
    v.hist = carve<uint32_t>(cur, sort_hist_words(n));
    v.rowtot = carve<uint32_t>(cur, kRadix);
    v.bytes = (size_t)(cur - reinterpret_cast<char*>(base));
    return v;
}

// ---------------------------------------------------------------- per-device one-time set-up (host)
// Opt-in shared-memory sizes and carve-outs are per-device function attributes.  Thread-safe: one
// atomic bit per device ordinal; two threads that race on the first use both set the (idempotent)
// attributes, nothing else.
struct DeviceOnce {
    std::atomic<uint64_t> bits{0};
    bool need(int dev) const { return ((bits.load(std::memory_order_acquire) >> (dev & 63)) & 1ull) == 0ull; }
    void done(int dev) { bits.fetch_or(1ull << (dev & 63), std::memory_order_release); }
};

// ---------------------------------------------------------------- error handling (host)
void set_error(const char* fmt, ...);
void count_launch(int n = 1);
void prof_mark(const char* name, cudaStream_t stream);

#define VEGS_CUDA_OK(expr)                                                              \
    do {                                                                                \
        cudaError_t _e = (expr);                                                        \
        if (_e != cudaSuccess) {                                                        \
            vegs::set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__,        \
                            cudaGetErrorString(_e));                                    \
            return VEGS_ECUDA;                                                          \
        }                                                                               \
    } while (0)

// After a kernel launch: always check the launch; in debug mode synchronise too, mirroring the
// reference's CHECK_CUDA (cuda_rasterizer/auxiliary.h:166-173).
#define VEGS_LAUNCH_CHECK(name, stream, debug)                                          \
    do {                                                                                \
        vegs::count_launch();                                                           \
        vegs::prof_mark(name, stream);                                                  \
        cudaError_t _e = cudaGetLastError();                                            \
        if (_e == cudaSuccess && (debug)) _e = cudaStreamSynchronize(stream);           \
        if (_e != cudaSuccess) {                                                        \
            vegs::set_error("kernel %s failed: %s", name, cudaGetErrorString(_e));      \
            return VEGS_ECUDA;                                                          \
        }                                                                               \
    } while (0)

// ---------------------------------------------------------------- host-side stage launchers
struct Camera {
    float view[16];
    float proj[16];
    float campos[3];
};

// sort.cu — stable LSD radix sort on bits [0, nbits) of 32-bit keys, 8 bits per pass.
// Pass p reads (kin,vin) and writes (kout,vout), then the pair is swapped.  If vin_first == nullptr the
// first pass uses the item index as value.  Returns (via *result_in_second) whether the sorted
// data ended in the (k1,v1) pair.  write_last_keys=false skips the key store of the last pass.
int radix_sort_pairs(uint32_t* k0, uint32_t* v0, uint32_t* k1, uint32_t* v1, bool implicit_vals,
                     uint32_t n, int nbits, uint32_t* hist, uint32_t* rowtot, bool write_last_keys,
                     bool* result_in_second, cudaStream_t stream, bool debug);

// preprocess.cu
int launch_preprocess_fwd(const VegsForwardArgs& a, const GeomView& g, const ScratchView& s,
                          cudaStream_t stream);
int launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present,
                        cudaStream_t stream);

// binning.cu
int launch_tile_count(int P, const uint32_t* order, const uint2* rect, uint32_t* chunksum,
                      uint32_t* partbase, uint32_t* num_rendered_dev, cudaStream_t stream, bool debug);
int launch_emit(int P, const uint32_t* order, const uint2* rect, const uint32_t* chunksum,
                const uint32_t* partbase, uint32_t nparts, int grid_x, uint32_t* tkeys,
                uint32_t* tvals, uint32_t R, const uint2* pbox, int W, int H, cudaStream_t stream,
                bool debug);
int launch_tile_ranges(uint32_t R, const uint32_t* tkeys_sorted, uint2* ranges,
                       uint32_t* tile_order, int T, cudaStream_t stream, bool debug);

// composite_fwd.cu
int launch_composite_fwd(const VegsForwardArgs& a, const GeomView& g, const ScratchView& s,
                         const ImgView& im, const BinView& b, cudaStream_t stream);

// composite_bwd.cu
int launch_composite_bwd(const VegsBackwardArgs& a, const GeomView& g, const ImgView& im,
                         const BinView& b, float4* acc, float4* map_acc, cudaStream_t stream);

// preprocess_bwd.cu
int launch_preprocess_bwd(const VegsBackwardArgs& a, const GeomView& g, const float4* acc,
                          const float4* map_acc, cudaStream_t stream);

// export.cu
int launch_export(const VegsForwardArgs& a, const GeomView& g, const ScratchView& s,
                  const ImgView& im, const BinView& b, uint32_t R, uint64_t* keys,
                  uint32_t* point_list, uint32_t* ranges, uint32_t* tiles_touched, float* depths,
                  float* means2D, float* conic_opacity, float* rgb, uint32_t* n_contrib,
                  cudaStream_t stream);

}  // namespace vegs
