// Front-to-back alpha compositing.  One warp per 8x4 pixel block ("unit"), 8 units per 16x16 tile.
//
// Replaces FORWARD::render / renderCUDA (reference cuda_rasterizer/forward.cu:262-441), including
// the VEGS additions: depth map, alpha map, accumulated covariance quaternion map (sequential
// slerp product, glm::mix — third_party/glm/glm/ext/quaternion_common.inl:4-26) and scale map.
//
// What differs from the reference's kernel (one 256-thread block per tile, every thread evaluates
// every list entry, __syncthreads per 256 entries):
//   * entry-parallel hit detection, pixel-parallel blending (composite_common.cuh): per batch of
//     32 list entries each lane tests ITS entry against the few block pixels inside the entry's
//     conservative box, a warp bit-transpose turns the 32 hit masks into per-pixel hit lists, and
//     each pixel lane then runs the reference's sequential arithmetic over its own hits only;
//   * no block-level synchronisation: units are independent warps, so the very long lists of a
//     LiDAR-density scene (several 1e5 entries in the horizon tiles) are shared by 8 warps that can
//     run on 8 SMs, units are issued longest-list-first, and a unit stops as soon as ITS 32 pixels
//     are saturated;
//   * colour, depth, scale and rotation travel in packed 16-byte records (no per-contribution
//     re-reads of several arrays, forward.cu:391-392); records of the next batch are loaded while
//     the current one is processed;
//   * acos(q.r) and sin(acos(q.r)) of the slerp are per-Gaussian constants from the preprocess
//     stage instead of being recomputed per (pixel, Gaussian).
// Every decision (power > 0, alpha >= 1/255, T(1-alpha) < 1e-4) is evaluated with the reference's
// float operation order on exactly the (pixel, Gaussian) pairs where the reference evaluates it
// to "contribute", so n_contrib and all outputs match.
// Roofline: not HBM-bound (32 B per list entry per unit from L2, 48 B more per entry with a hit,
// 52 B written per pixel); the bound is instruction issue of the per-hit blend (slerp: 2 sin,
// 4 div, quaternion product) and of the box tests.
#include <cfloat>
#include "common.cuh"
#include "composite_common.cuh"

namespace vegs {

namespace {

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
};

__global__ void __launch_bounds__(32) composite_fwd_kernel(const FwdParams p) {
    __shared__ __align__(16) float4 sA[32];
    __shared__ __align__(16) float4 sB[32];
    __shared__ __align__(16) float4 sC[32];
    __shared__ __align__(16) float4 sD[32];
    __shared__ __align__(16) float4 sE[32];

    const long long t_start = clock64();
    const int lane = threadIdx.x;
    const uint32_t unit = blockIdx.x;
    const uint32_t tile = p.tile_order[unit >> 3];
    const int blk = (int)(unit & 7u);
    const int tile_x = (int)(tile % (uint32_t)p.gx), tile_y = (int)(tile / (uint32_t)p.gx);
    const int bx0 = tile_x * kTile + (blk & 1) * kUnitW;
    const int by0 = tile_y * kTile + (blk >> 1) * kUnitH;
    if (bx0 >= p.W || by0 >= p.H) return;                     // block entirely outside the image
    const int bx1 = min(bx0 + kUnitW - 1, p.W - 1), by1 = min(by0 + kUnitH - 1, p.H - 1);
    const int px = bx0 + (lane & 7), py = by0 + (lane >> 3);
    const bool inside = px < p.W && py < p.H;
    const float pixfx = (float)px, pixfy = (float)py;

    const uint2 range = p.ranges[tile];
    const int n = (int)(range.y - range.x);
    const uint32_t* list = p.point_list + range.x;

    float T = 1.0f;
    float C0 = 0.f, C1 = 0.f, C2 = 0.f, Dacc = 0.f;
    float S0 = 0.f, S1 = 0.f, S2 = 0.f;
    float Qw = 1.f, Qx = 0.f, Qy = 0.f, Qz = 0.f;
    uint32_t last_contributor = 0;
    bool done = !inside;

    // software pipeline: records of batch k+1 and indices of batch k+2 are in flight during batch k
    int id_cur = (lane < n) ? (int)list[lane] : -1;
    float4 a_cur = make_float4(0.f, 0.f, 0.f, 0.f), b_cur = a_cur;
    if (id_cur >= 0) { a_cur = ldg_f4(p.recA + id_cur); b_cur = ldg_f4(p.recB + id_cur); }
    int id_nxt = (32 + lane < n) ? (int)list[32 + lane] : -1;

    for (int base = 0; base < n; base += 32) {
        float4 a_nxt = make_float4(0.f, 0.f, 0.f, 0.f), b_nxt = a_nxt;
        if (id_nxt >= 0) { a_nxt = ldg_f4(p.recA + id_nxt); b_nxt = ldg_f4(p.recB + id_nxt); }
        const int nn = base + 64 + lane;
        const int id_nn = (nn < n) ? (int)list[nn] : -1;

        // ---- entry-parallel: lane t tests entry base + t against the block's pixels
        const uint32_t h = (id_cur >= 0) ? entry_hits(a_cur, b_cur, bx0, by0, bx1, by1) : 0u;
        if (__any_sync(0xffffffffu, h != 0u)) {
            if (h != 0u) {
                sA[lane] = a_cur;
                sB[lane] = b_cur;
                sC[lane] = ldg_f4(p.recC + id_cur);
                sD[lane] = ldg_f4(p.recD + id_cur);
                sE[lane] = ldg_f4(p.recE + id_cur);
            }
            uint32_t mybits = transpose32(h, lane);   // lane p: entries of this batch that hit pixel p
            __syncwarp();
            // ---- pixel-parallel: the reference's per-pixel sequence over this pixel's hits
            while (mybits != 0u && !done) {
                const int t = __ffs(mybits) - 1;
                mybits &= mybits - 1;
                const float4 a = sA[t];    // x, y, conic.a, conic.b
                const float4 bb = sB[t];   // conic.c, opacity, half2(ex, ey), depth
                float dx, dy;
                const float power = pair_power(a, bb, pixfx, pixfy, dx, dy);
                const float alpha = fminf(0.99f, __fmul_rn(bb.y, expf(power)));
                const float test_T = __fmul_rn(T, __fsub_rn(1.0f, alpha));
                if (test_T < 0.0001f) {      // this Gaussian is not blended (forward.cu:382-386)
                    done = true;
                    break;
                }
                const float4 c = sC[t];    // r, g, b, sin(theta)
                const float4 d = sD[t];    // scale xyz, rot.r
                const float4 e = sE[t];    // rot xyz, theta
                C0 = __fmaf_rn(__fmul_rn(c.x, alpha), T, C0);
                C1 = __fmaf_rn(__fmul_rn(c.y, alpha), T, C1);
                C2 = __fmaf_rn(__fmul_rn(c.z, alpha), T, C2);
                Dacc = __fmaf_rn(__fmul_rn(bb.w, alpha), T, Dacc);
                S0 = __fmaf_rn(__fmul_rn(d.x, alpha), T, S0);
                S1 = __fmaf_rn(__fmul_rn(d.y, alpha), T, S1);
                S2 = __fmaf_rn(__fmul_rn(d.z, alpha), T, S2);
                // r = mix(identity, q, w) with q = (w = rot.r, x, y, z), w = alpha * T
                const float wgt = __fmul_rn(alpha, T);
                const float om = __fsub_rn(1.0f, wgt);
                float rw, rx, ry, rz;
                if (d.w > 1.0f - FLT_EPSILON) {
                    rw = __fmaf_rn(d.w, wgt, om);
                    rx = __fmul_rn(e.x, wgt);
                    ry = __fmul_rn(e.y, wgt);
                    rz = __fmul_rn(e.z, wgt);
                } else {
                    const float s0 = sinf(__fmul_rn(om, e.w));
                    const float s1 = sinf(__fmul_rn(wgt, e.w));
                    rw = __fdiv_rn(__fmaf_rn(d.w, s1, s0), c.w);
                    rx = __fdiv_rn(__fmul_rn(e.x, s1), c.w);
                    ry = __fdiv_rn(__fmul_rn(e.y, s1), c.w);
                    rz = __fdiv_rn(__fmul_rn(e.z, s1), c.w);
                }
                // Q = Q * r (Hamilton product, glm/detail/type_quat.inl:286-296)
                const float nw = Qw * rw - Qx * rx - Qy * ry - Qz * rz;
                const float nx = Qw * rx + Qx * rw + Qy * rz - Qz * ry;
                const float ny = Qw * ry + Qy * rw + Qz * rx - Qx * rz;
                const float nz = Qw * rz + Qz * rw + Qx * ry - Qy * rx;
                Qw = nw; Qx = nx; Qy = ny; Qz = nz;
                T = test_T;
                last_contributor = (uint32_t)(base + t) + 1u;   // ordinal in the tile's list
            }
            __syncwarp();
            if (__all_sync(0xffffffffu, done)) break;   // every pixel of the block is saturated
        }
        id_cur = id_nxt; a_cur = a_nxt; b_cur = b_nxt; id_nxt = id_nn;
    }

    if (inside) {
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
    if (lane == 0) atomicMax(&p.tile_clocks[tile], (uint32_t)((clock64() - t_start) >> 10));
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
    const int T = p.gx * ((a.H + kTile - 1) / kTile);
    VEGS_CUDA_OK(cudaMemsetAsync(im.tile_clocks, 0, (size_t)2 * T * sizeof(uint32_t), stream));
    composite_fwd_kernel<<<T * kUnitsPerTile, 32, 0, stream>>>(p);
    VEGS_LAUNCH_CHECK("composite_fwd", stream, a.debug);
    return VEGS_OK;
}

}  // namespace vegs
