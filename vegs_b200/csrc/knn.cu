// distCUDA2: mean of the three smallest squared distances from every point to the other points.
//
// Replaces SimpleKNN::knn (reference submodules/simple-knn/simple_knn.cu:45-221, spatial.cu:15-26).
//
// The result is the exact 3-NN answer, so any exact search gives the reference's values; what is
// pinned is the float arithmetic of a distance, d = fma(dz,dz, fma(dx,dx, dy*dy)) with
// d* = other - self (SASS of the reference's boxMeanDist), and the final (b0 + b1 + b2) / 3.0f on
// the ascending triple.
//
// Search structure (built per call, no host synchronisation, no allocation):
//   1. bounding box by atomic min/max on order-preserving integer images of the floats
//   2. 30-bit Morton codes, radix sort (sort.cu), points gathered into sorted order
//   3. three levels of AABBs over runs of 32 / 1024 / 32768 consecutive sorted points
//   4. one thread per point (in sorted order, so a warp's 32 queries are neighbours and traverse
//      the same boxes): seed the bound with the +-3 sorted neighbours as the reference does, then
//      descend L3 -> L2 -> L1 -> points, pruning a box when its distance exceeds the current third
//      best.  The box distance uses the same fma order on component-wise smaller magnitudes, so it
//      never exceeds the computed distance of a point inside the box: pruning is exact in float.
//      (A warp-cooperative variant - one traversal per 32 queries with their common box and the
//      largest of their bounds, leaf points handed round by shuffles, knn3_warp_kernel,
//      VEGS_KNN_WARP=1 - gives identical results but measured 4x slower at LiDAR density: the
//      shared bound is much looser than each query's own, and every visited leaf costs all 32
//      lanes 32 evaluations.  ncu of the per-thread kernel: 86 % issue utilisation, 16 of 32 lanes
//      active, 12.6 G warp instructions for 6 M queries.)
// The reference tests every one of the P/1024 boxes from every thread (3.5e10 tests at 6 M points).
// Roofline: latency / L2-gather bound; algorithmic HBM bytes 16 per point.
#include <algorithm>
#include <cfloat>
#include <cstdlib>
#include <mutex>
#include "common.cuh"

namespace vegs {

namespace {

constexpr int kFan = 32;

struct KnnView {
    uint32_t* bbox;  // 6 ordered-uint words: min xyz, max xyz
    uint32_t *codes0, *codes1, *idx0, *idx1;
    uint32_t *hist, *rowtot;
    float4* pts;   // sorted points (x, y, z, original index bits)
    float4* box1;  // 2 float4 per box: min, max
    float4* box2;
    float4* box3;
    size_t bytes;
};

KnnView knn_view(void* base, int P) {
    KnnView v;
    char* cur = reinterpret_cast<char*>(base);
    const size_t n = (size_t)(P > 0 ? P : 1);
    const size_t n1 = div_up(n, kFan), n2 = div_up(n1, kFan), n3 = div_up(n2, kFan);
    v.bbox = carve<uint32_t>(cur, 8);
    v.codes0 = carve<uint32_t>(cur, n);
    v.codes1 = carve<uint32_t>(cur, n);
    v.idx0 = carve<uint32_t>(cur, n);
    v.idx1 = carve<uint32_t>(cur, n);
    v.hist = carve<uint32_t>(cur, sort_hist_words(n));
    v.rowtot = carve<uint32_t>(cur, kRadix);
    v.pts = carve<float4>(cur, n);
    v.box1 = carve<float4>(cur, 2 * n1);
    v.box2 = carve<float4>(cur, 2 * n2);
    v.box3 = carve<float4>(cur, 2 * n3);
    v.bytes = (size_t)(cur - reinterpret_cast<char*>(base));
    return v;
}

__device__ __forceinline__ uint32_t f2ord(float f) {
    const uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

__global__ void bbox_init_kernel(uint32_t* bbox) {
    if (threadIdx.x < 3) bbox[threadIdx.x] = 0xffffffffu;
    else if (threadIdx.x < 6) bbox[threadIdx.x] = 0u;
}

__global__ void __launch_bounds__(256) bbox_kernel(int P, const float* __restrict__ pts,
                                                   uint32_t* __restrict__ bbox) {
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = blockIdx.x * 256 + threadIdx.x; i < P; i += gridDim.x * 256) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const float v = pts[3 * (size_t)i + k];
            mn[k] = fminf(mn[k], v);
            mx[k] = fmaxf(mx[k], v);
        }
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
            mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            atomicMin(&bbox[k], f2ord(mn[k]));
            atomicMax(&bbox[3 + k], f2ord(mx[k]));
        }
    }
}

__device__ __forceinline__ uint32_t spread10(uint32_t x) {
    x = (x | (x << 16)) & 0x030000FFu;
    x = (x | (x << 8)) & 0x0300F00Fu;
    x = (x | (x << 4)) & 0x030C30C3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}

__global__ void __launch_bounds__(256) morton_kernel(int P, const float* __restrict__ pts,
                                                     const uint32_t* __restrict__ bbox,
                                                     uint32_t* __restrict__ codes) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= P) return;
    uint32_t q[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const float lo = ord2f(bbox[k]), hi = ord2f(bbox[3 + k]);
        const float ext = hi - lo;
        float t = ext > 0.f ? (pts[3 * (size_t)i + k] - lo) / ext : 0.f;
        t = fminf(fmaxf(t, 0.f), 1.f);
        q[k] = (uint32_t)(t * 1023.0f);
    }
    codes[i] = spread10(q[0]) | (spread10(q[1]) << 1) | (spread10(q[2]) << 2);
}

__global__ void __launch_bounds__(256) gather_points_kernel(int P, const float* __restrict__ pts,
                                                            const uint32_t* __restrict__ order,
                                                            float4* __restrict__ sorted) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= P) return;
    const uint32_t g = order[i];
    sorted[i] = make_float4(pts[3 * (size_t)g], pts[3 * (size_t)g + 1], pts[3 * (size_t)g + 2],
                            __uint_as_float(g));
}

// One warp per output box: AABB of up to 32 children (points when kLeaf, else child boxes).
template <bool kLeaf>
__global__ void __launch_bounds__(256) build_boxes_kernel(uint32_t n_in, const float4* __restrict__ in,
                                                          uint32_t n_out, float4* __restrict__ out) {
    const uint32_t w = (blockIdx.x * 256u + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= n_out) return;
    const uint32_t c = w * kFan + lane;
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (c < n_in) {
        if (kLeaf) {
            const float4 p = in[c];
            mn[0] = mx[0] = p.x; mn[1] = mx[1] = p.y; mn[2] = mx[2] = p.z;
        } else {
            const float4 a = in[2 * (size_t)c], b = in[2 * (size_t)c + 1];
            mn[0] = a.x; mn[1] = a.y; mn[2] = a.z;
            mx[0] = b.x; mx[1] = b.y; mx[2] = b.z;
        }
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
            mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
        }
    }
    if (lane == 0) {
        out[2 * (size_t)w] = make_float4(mn[0], mn[1], mn[2], 0.f);
        out[2 * (size_t)w + 1] = make_float4(mx[0], mx[1], mx[2], 0.f);
    }
}

__device__ __forceinline__ float point_dist2(const float4& self, const float4& other) {
    const float dx = __fsub_rn(other.x, self.x);
    const float dy = __fsub_rn(other.y, self.y);
    const float dz = __fsub_rn(other.z, self.z);
    return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
}

// Lower bound (in the same float arithmetic) of point_dist2 to any point inside the box.
__device__ __forceinline__ float box_dist2(const float4& p, const float4& mn, const float4& mx) {
    float dx = 0.f, dy = 0.f, dz = 0.f;
    if (p.x < mn.x || p.x > mx.x) dx = fminf(fabsf(__fsub_rn(mn.x, p.x)), fabsf(__fsub_rn(mx.x, p.x)));
    if (p.y < mn.y || p.y > mx.y) dy = fminf(fabsf(__fsub_rn(mn.y, p.y)), fabsf(__fsub_rn(mx.y, p.y)));
    if (p.z < mn.z || p.z > mx.z) dz = fminf(fabsf(__fsub_rn(mn.z, p.z)), fabsf(__fsub_rn(mx.z, p.z)));
    return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
}

__device__ __forceinline__ void update3(float best[3], float d) {
#pragma unroll
    for (int j = 0; j < 3; j++) {
        if (best[j] > d) {
            const float t = best[j];
            best[j] = d;
            d = t;
        }
    }
}

__global__ void __launch_bounds__(128)
knn3_kernel(uint32_t P, const float4* __restrict__ pts, const float4* __restrict__ box1,
            uint32_t n1, const float4* __restrict__ box2, uint32_t n2,
            const float4* __restrict__ box3, uint32_t n3, float* __restrict__ out) {
    const uint32_t i = blockIdx.x * 128u + threadIdx.x;
    if (i >= P) return;
    const float4 self = pts[i];
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    {
        const uint32_t lo = i >= 3 ? i - 3 : 0u;
        const uint32_t hi = min(P - 1, i + 3);
        for (uint32_t k = lo; k <= hi; k++)
            if (k != i) update3(best, point_dist2(self, pts[k]));
    }
    // The seed only provides the pruning bound; the full search recounts every point once.
    float bound = best[2];
    best[0] = best[1] = best[2] = FLT_MAX;
    for (uint32_t b3 = 0; b3 < n3; b3++) {
        if (box_dist2(self, box3[2 * b3], box3[2 * b3 + 1]) > fminf(bound, best[2])) continue;
        const uint32_t e2 = min(n2, (b3 + 1) * kFan);
        for (uint32_t b2 = b3 * kFan; b2 < e2; b2++) {
            if (box_dist2(self, box2[2 * b2], box2[2 * b2 + 1]) > fminf(bound, best[2])) continue;
            const uint32_t e1 = min(n1, (b2 + 1) * kFan);
            for (uint32_t b1 = b2 * kFan; b1 < e1; b1++) {
                if (box_dist2(self, box1[2 * (size_t)b1], box1[2 * (size_t)b1 + 1]) >
                    fminf(bound, best[2]))
                    continue;
                const uint32_t e0 = min(P, (b1 + 1) * kFan);
                for (uint32_t k = b1 * kFan; k < e0; k++)
                    if (k != i) update3(best, point_dist2(self, pts[k]));
            }
        }
    }
    out[__float_as_uint(self.w)] = (best[0] + best[1] + best[2]) / 3.0f;
}

// Lower bound (same float arithmetic) of point_dist2 between ANY point of box g and ANY point of
// box c: per axis the gap between the boxes (0 when they overlap); float subtraction, squaring and
// fma are monotone, so the bound never exceeds the computed distance of a pair it covers.
__device__ __forceinline__ float box_box_dist2(const float4& gmn, const float4& gmx, const float4& cmn, const float4& cmx) {
    const float dx = fmaxf(0.f, fmaxf(__fsub_rn(cmn.x, gmx.x), __fsub_rn(gmn.x, cmx.x)));
    const float dy = fmaxf(0.f, fmaxf(__fsub_rn(cmn.y, gmx.y), __fsub_rn(gmn.y, cmx.y)));
    const float dz = fmaxf(0.f, fmaxf(__fsub_rn(cmn.z, gmx.z), __fsub_rn(gmn.z, cmx.z)));
    return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
}

// Warp-cooperative search (opt-in experiment, slower than knn3_kernel on the benchmark scene, see
// the header): one warp answers 32 consecutive queries of the sorted order (spatial neighbours)
// TOGETHER.  The hierarchy is walked once per warp with the queries' common bounding
// box and the largest of their current third-best distances as the pruning bound, 32 child boxes
// tested by the 32 lanes at a time; a surviving leaf's 32 points are loaded by one coalesced read
// and handed round with shuffles, every lane updating its own three best.  Compared with one
// private traversal per thread (knn3_kernel) the lanes never diverge and a box test is paid once
// per warp instead of once per query.  The pairs evaluated are a superset of each query's own, the
// distance arithmetic is the same, so the three smallest values per query are identical.
__global__ void __launch_bounds__(128)
knn3_warp_kernel(uint32_t P, const float4* __restrict__ pts, const float4* __restrict__ box1,
                 uint32_t n1, const float4* __restrict__ box2, uint32_t n2,
                 const float4* __restrict__ box3, uint32_t n3, float* __restrict__ out) {
    const uint32_t i = blockIdx.x * 128u + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const uint32_t warp_first = i - lane;
    if (warp_first >= P) return;
    const bool active = i < P;
    const float4 self = pts[active ? i : P - 1];
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    if (active) {
        const uint32_t lo = i >= 3 ? i - 3 : 0u;
        const uint32_t hi = min(P - 1, i + 3);
        for (uint32_t k = lo; k <= hi; k++)
            if (k != i) update3(best, point_dist2(self, pts[k]));
    }
    // the seed only provides the pruning bound; the search below counts every point once
    const float seed = active ? best[2] : 0.f;
    best[0] = best[1] = best[2] = FLT_MAX;
    // the queries' common box (inactive lanes of a partial last warp stay out of it)
    float4 gmn = make_float4(active ? self.x : FLT_MAX, active ? self.y : FLT_MAX, active ? self.z : FLT_MAX, 0.f);
    float4 gmx = make_float4(active ? self.x : -FLT_MAX, active ? self.y : -FLT_MAX, active ? self.z : -FLT_MAX, 0.f);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        gmn.x = fminf(gmn.x, __shfl_xor_sync(0xffffffffu, gmn.x, o)); gmx.x = fmaxf(gmx.x, __shfl_xor_sync(0xffffffffu, gmx.x, o));
        gmn.y = fminf(gmn.y, __shfl_xor_sync(0xffffffffu, gmn.y, o)); gmx.y = fmaxf(gmx.y, __shfl_xor_sync(0xffffffffu, gmx.y, o));
        gmn.z = fminf(gmn.z, __shfl_xor_sync(0xffffffffu, gmn.z, o)); gmx.z = fmaxf(gmx.z, __shfl_xor_sync(0xffffffffu, gmx.z, o));
    }
    // pruning bound of the warp: the largest "third best so far" of its queries (non-negative floats
    // order like their bit patterns)
    auto group_bound = [&]() {
        const float mine = active ? fminf(seed, best[2]) : 0.f;
        return __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(mine)));
    };
    float bound = group_bound();
    for (uint32_t c3 = 0; c3 < n3; c3 += 32) {
        const uint32_t b3 = c3 + lane;
        bool hit3 = b3 < n3 && !(box_box_dist2(gmn, gmx, box3[2 * b3], box3[2 * b3 + 1]) > bound);
        uint32_t m3 = __ballot_sync(0xffffffffu, hit3);
        while (m3 != 0u) {
            const uint32_t p3 = c3 + (uint32_t)(__ffs(m3) - 1);
            m3 &= m3 - 1;
            const uint32_t b2 = p3 * kFan + lane;
            bool hit2 = b2 < n2 && !(box_box_dist2(gmn, gmx, box2[2 * b2], box2[2 * b2 + 1]) > bound);
            uint32_t m2 = __ballot_sync(0xffffffffu, hit2);
            while (m2 != 0u) {
                const uint32_t p2 = p3 * kFan + (uint32_t)(__ffs(m2) - 1);
                m2 &= m2 - 1;
                const uint32_t b1 = p2 * kFan + lane;
                bool hit1 = b1 < n1 && !(box_box_dist2(gmn, gmx, box1[2 * (size_t)b1], box1[2 * (size_t)b1 + 1]) > bound);
                uint32_t m1 = __ballot_sync(0xffffffffu, hit1);
                while (m1 != 0u) {
                    const uint32_t p1 = p2 * kFan + (uint32_t)(__ffs(m1) - 1);
                    m1 &= m1 - 1;
                    // the bound may have tightened since the ballot: test this leaf again
                    if (box_box_dist2(gmn, gmx, box1[2 * (size_t)p1], box1[2 * (size_t)p1 + 1]) > bound) continue;
                    const uint32_t k0 = p1 * kFan;
                    const uint32_t cnt = min((uint32_t)kFan, P - k0);
                    const float4 mine = pts[min(k0 + (uint32_t)lane, P - 1)];
                    for (uint32_t t = 0; t < cnt; t++) {
                        float4 o;
                        o.x = __shfl_sync(0xffffffffu, mine.x, (int)t);
                        o.y = __shfl_sync(0xffffffffu, mine.y, (int)t);
                        o.z = __shfl_sync(0xffffffffu, mine.z, (int)t);
                        if (active && k0 + t != i) update3(best, point_dist2(self, o));
                    }
                    bound = group_bound();
                }
            }
        }
    }
    if (active) out[__float_as_uint(self.w)] = (best[0] + best[1] + best[2]) / 3.0f;
}

}  // namespace

}  // namespace vegs

extern "C" size_t vegs_dist2_workspace_bytes(int32_t P) {
    return vegs::knn_view(nullptr, P).bytes + 256;
}

extern "C" int vegs_dist2(int32_t P, const float* points, float* out, void* workspace,
                          size_t workspace_bytes, void* stream_) {
    using namespace vegs;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    if (P < 0 || (P > 0 && (!points || !out || !workspace))) {
        set_error("vegs_dist2: bad argument");
        return VEGS_EINVAL;
    }
    if (P == 0) return VEGS_OK;
    if (workspace_bytes < vegs_dist2_workspace_bytes(P)) {
        set_error("vegs_dist2: workspace too small (%zu < %zu)", workspace_bytes,
                  vegs_dist2_workspace_bytes(P));
        return VEGS_ESIZE;
    }
    void* base = reinterpret_cast<void*>(align_up(reinterpret_cast<uintptr_t>(workspace)));
    const KnnView v = knn_view(base, P);
    const uint32_t n = (uint32_t)P;
    const uint32_t n1 = div_up(n, kFan), n2 = div_up(n1, kFan), n3 = div_up(n2, kFan);

    bbox_init_kernel<<<1, 32, 0, stream>>>(v.bbox);
    VEGS_LAUNCH_CHECK("bbox_init", stream, false);
    const int bb = (int)std::min<uint32_t>(div_up(n, 256), 148u * 8u);
    bbox_kernel<<<bb, 256, 0, stream>>>(P, points, v.bbox);
    VEGS_LAUNCH_CHECK("bbox", stream, false);
    morton_kernel<<<div_up(n, 256), 256, 0, stream>>>(P, points, v.bbox, v.codes0);
    VEGS_LAUNCH_CHECK("morton", stream, false);
    bool second = false;
    int rc = radix_sort_pairs(v.codes0, v.idx0, v.codes1, v.idx1, true, n, 30, v.hist, v.rowtot,
                              false, &second, stream, false);
    if (rc) return rc;
    const uint32_t* order = second ? v.idx1 : v.idx0;
    gather_points_kernel<<<div_up(n, 256), 256, 0, stream>>>(P, points, order, v.pts);
    VEGS_LAUNCH_CHECK("gather_points", stream, false);
    build_boxes_kernel<true><<<div_up((uint64_t)n1 * 32, 256), 256, 0, stream>>>(n, v.pts, n1, v.box1);
    VEGS_LAUNCH_CHECK("build_boxes1", stream, false);
    build_boxes_kernel<false><<<div_up((uint64_t)n2 * 32, 256), 256, 0, stream>>>(n1, v.box1, n2, v.box2);
    VEGS_LAUNCH_CHECK("build_boxes2", stream, false);
    build_boxes_kernel<false><<<div_up((uint64_t)n3 * 32, 256), 256, 0, stream>>>(n2, v.box2, n3, v.box3);
    VEGS_LAUNCH_CHECK("build_boxes3", stream, false);
    // measured at 6 M points: per-thread traversal 13.4 ms, warp-cooperative 52 ms (below) - default per thread
    static const bool per_thread = [] { const char* e = getenv("VEGS_KNN_WARP"); return !(e != nullptr && e[0] == '1'); }();
    if (per_thread) knn3_kernel<<<div_up(n, 128), 128, 0, stream>>>(n, v.pts, v.box1, n1, v.box2, n2, v.box3, n3, out);
    else knn3_warp_kernel<<<div_up(n, 128), 128, 0, stream>>>(n, v.pts, v.box1, n1, v.box2, n2, v.box3, n3, out);
    VEGS_LAUNCH_CHECK("knn3", stream, false);
    return VEGS_OK;
}

// ------------------------------------------------------------------------------------------------
// vegs_dist2_noalloc: distCUDA2 for callers that must not allocate device memory.
//
// Who this is for: code that runs where cudaMalloc / cudaFree are off limits — a host application
// that hands this library a fixed memory budget and owns every allocation itself, or a real-time /
// latency-sensitive loop where cudaMalloc and cudaFree are unacceptable because they serialise with
// all work on the device.  Such a caller still passes its own input and output arrays; what it
// cannot provide is the search structure's scratch memory, so that lives in the array below.
//
// The scratch is one file-scope __device__ array.  It is part of the library image: every GPU that
// loads libvegs_b200.so holds these bytes from load until the process exits, whether or not this
// entry point is ever called, which is why the size (and so the largest point count) is modest.
// In host code the array's name is only a handle for the runtime, not an address the GPU can use;
// the device address is looked up once per device with cudaGetSymbolAddress and then passed to the
// kernels like any other workspace pointer.
constexpr int32_t kNoAllocMaxPoints = VEGS_DIST2_NOALLOC_MAX_POINTS;
constexpr size_t kNoAllocScratchBytes = VEGS_DIST2_NOALLOC_SCRATCH_BYTES;
// the carve-out of knn_view() for the largest accepted input: 4 sort arrays + histograms + sorted
// points + three box levels (each region rounded up to 256 B) must fit, plus the alignment slack
constexpr size_t kNoAllocNeed =
    256 + (size_t)VEGS_DIST2_NOALLOC_MAX_POINTS * (4 * 4 + 16) +
    4 * (size_t)vegs::kRadix * vegs::kSortMaxPasses * ((VEGS_DIST2_NOALLOC_MAX_POINTS + vegs::kSortTile - 1) / vegs::kSortTile) +
    4 * ((size_t)vegs::kSortMaxPasses * vegs::kRadix + 64) + vegs::kRadix * 4 +
    32 * ((size_t)VEGS_DIST2_NOALLOC_MAX_POINTS / 32 + VEGS_DIST2_NOALLOC_MAX_POINTS / 1024 + 16) + 16 * 256;
static_assert(kNoAllocNeed <= kNoAllocScratchBytes,
              "VEGS_DIST2_NOALLOC_SCRATCH_BYTES cannot hold the workspace of VEGS_DIST2_NOALLOC_MAX_POINTS points");
static __device__ __align__(256) unsigned char g_dist2_noalloc_scratch[kNoAllocScratchBytes];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;
void* g_noalloc_ptr[kMaxDevices] = {nullptr};
}  // namespace

extern "C" int vegs_dist2_noalloc(int32_t P, const float* points, float* out, void* stream_) {
    using namespace vegs;
    if (P < 0 || (P > 0 && (!points || !out))) {
        set_error("vegs_dist2_noalloc: bad argument");
        return VEGS_EINVAL;
    }
    if (P > kNoAllocMaxPoints) {
        set_error("vegs_dist2_noalloc: %d points exceed the built-in scratch (at most %d points); "
                  "use vegs_dist2 with a caller-provided workspace", P, kNoAllocMaxPoints);
        return VEGS_ESIZE;
    }
    if (P == 0) return VEGS_OK;
    int dev = -1;
    VEGS_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) {
        set_error("vegs_dist2_noalloc: device ordinal %d not supported", dev);
        return VEGS_EINVAL;
    }
    void* scratch = nullptr;
    {
        // Set-up, once per device, with that device current (it is: we just asked which one is;
        // the header requires the current device to be the stream's).  The cached address stays
        // valid until the process exits or the device is reset.
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        if (g_noalloc_ptr[dev] == nullptr) {
            void* p = nullptr;
            VEGS_CUDA_OK(cudaGetSymbolAddress(&p, g_dist2_noalloc_scratch));
            g_noalloc_ptr[dev] = p;
        }
        scratch = g_noalloc_ptr[dev];
    }
    if (vegs_dist2_workspace_bytes(P) > kNoAllocScratchBytes) {
        set_error("vegs_dist2_noalloc: internal error, scratch array too small for %d points", P);
        return VEGS_ESIZE;
    }
    // Nothing in the scratch has to start at zero: the first kernel re-initialises the bounding-box
    // words and every other region is fully written before it is read, so the previous call's
    // contents are harmless.  The existing entry point does the rest; it allocates nothing.
    return vegs_dist2(P, points, out, scratch, kNoAllocScratchBytes, stream_);
}
