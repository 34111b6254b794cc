// Front-to-back alpha compositing, one 16x16 tile per CTA, one pixel per thread.
//
// Replaces FORWARD::render / renderCUDA (reference cuda_rasterizer/forward.cu:262-441), including
// the VEGS additions: depth map, alpha map, accumulated covariance quaternion map (sequential
// slerp product, glm::mix — third_party/glm/glm/ext/quaternion_common.inl:4-26) and scale map.
//
// What differs from the reference's kernel:
//   * the tile's Gaussians are staged in shared memory with asynchronous 16-byte copies
//     (cp.async, SASS LDGSTS) of packed records, double buffered: batch b+1 is in flight while
//     batch b is composited, and the index run for batch b+2 is already loaded;
//   * colour and depth travel in the staged records instead of being re-read from global memory
//     inside the per-pixel loop (forward.cu:391-392);
//   * acos(q.r) and sin(acos(q.r)) of the slerp are per-Gaussian constants from the preprocess
//     stage instead of being recomputed per (pixel, Gaussian);
//   * per-block culling: each warp owns an 8x4 pixel block.  For every staged entry one thread
//     computes the bounding box of the region where alpha can reach 1/255
//     (power >= -tau, tau = log(255*opacity) + margin) and an 8-bit mask of the pixel blocks it
//     overlaps; a warp then visits only its own entries (ballot + find-first-set keeps their
//     order).  At LiDAR density a Gaussian covers a few pixels, so a warp skips most of the list;
//     the decisions of the entries it does visit are the reference's, so results are unchanged;
//   * exp() is skipped when power < -tau (alpha would be below 1/255 with margin);
//   * tiles are taken from a longest-list-first table so the few very long lists start first.
// The per-pixel arithmetic that decides which Gaussians contribute (power, alpha, test_T) uses the
// reference's exact float operation order so n_contrib matches bit for bit.
// Roofline: not HBM-bound (gathers 80 B per fetched instance, writes 52 B per pixel); the bound is
// the issue rate / latency of the per-(pixel, Gaussian) loop and MUFU for exp / sin.
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
};

__global__ void __launch_bounds__(kTilePix) composite_fwd_kernel(const FwdParams p) {
    __shared__ __align__(16) float4 sA[2][kTilePix];
    __shared__ __align__(16) float4 sB[2][kTilePix];
    __shared__ __align__(16) float4 sC[2][kTilePix];
    __shared__ __align__(16) float4 sD[2][kTilePix];
    __shared__ __align__(16) float4 sE[2][kTilePix];
    __shared__ uint32_t sMask[kTilePix];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t tile = p.tile_order[blockIdx.x];
    const int tile_x = (int)(tile % (uint32_t)p.gx), tile_y = (int)(tile / (uint32_t)p.gx);
    const int X0 = tile_x * kTile, Y0 = tile_y * kTile;
    // warp w owns the 8x4 pixel block (w & 1, w >> 1)
    const int px = X0 + (warp & 1) * 8 + (lane & 7);
    const int py = Y0 + (warp >> 1) * 4 + (lane >> 3);
    const bool inside = px < p.W && py < p.H;
    const float pixfx = (float)px, pixfy = (float)py;

    const uint2 range = p.ranges[tile];
    const int n = (int)(range.y - range.x);
    const int nbatches = (n + kTilePix - 1) / kTilePix;
    const uint32_t* list = p.point_list + range.x;

    float T = 1.0f;
    float C0 = 0.f, C1 = 0.f, C2 = 0.f, Dacc = 0.f;
    float S0 = 0.f, S1 = 0.f, S2 = 0.f;
    float Qw = 1.f, Qx = 0.f, Qy = 0.f, Qz = 0.f;
    uint32_t last_contributor = 0;
    bool done = !inside;

    auto issue = [&](int stage, int id) {
        if (id >= 0) {
            cp_async16(&sA[stage][tid], p.recA + id);
            cp_async16(&sB[stage][tid], p.recB + id);
            cp_async16(&sC[stage][tid], p.recC + id);
            cp_async16(&sD[stage][tid], p.recD + id);
            cp_async16(&sE[stage][tid], p.recE + id);
        }
        cp_async_commit();
    };

    int id_next = -1;
    if (nbatches > 0) {
        const int id0 = (tid < n) ? (int)list[tid] : -1;
        issue(0, id0);
        id_next = (kTilePix + tid < n) ? (int)list[kTilePix + tid] : -1;
    }

    for (int b = 0; b < nbatches; b++) {
        cp_async_wait_all();
        // Barrier: batch b is visible to all threads, and every thread has finished reading the
        // other stage and the mask array.  The vote ends the tile early when every pixel is
        // saturated (forward.cu:334-336).
        const int num_done = __syncthreads_count(done);
        if (num_done == kTilePix) break;
        if (b + 1 < nbatches) {
            issue((b + 1) & 1, id_next);
            const int nn = (b + 2) * kTilePix + tid;
            id_next = (nn < n) ? (int)list[nn] : -1;
        }
        const int st = b & 1;
        const int cnt = min(kTilePix, n - b * kTilePix);
        sMask[tid] = (tid < cnt) ? block_mask(sA[st][tid], sB[st][tid], X0, Y0) : 0u;
        __syncthreads();

        const uint32_t base_ord = (uint32_t)(b * kTilePix);
        for (int c0 = 0; c0 < cnt; c0 += 32) {
            if (__all_sync(0xffffffffu, done)) break;
            uint32_t bits = __ballot_sync(0xffffffffu, (sMask[c0 + lane] >> warp) & 1u);
            while (bits) {
                const int j = c0 + __ffs(bits) - 1;
                bits &= bits - 1;
                if (done) continue;
                const float4 a = sA[st][j];   // x, y, conic.a, conic.b
                const float4 bb = sB[st][j];  // conic.c, opacity, tau, depth
                const float dx = __fsub_rn(a.x, pixfx);
                const float dy = __fsub_rn(a.y, pixfy);
                // power = -0.5f*(A*dx*dx + C*dy*dy) - B*dx*dy in the reference's operation order
                const float sum = __fmaf_rn(__fmul_rn(a.z, dx), dx, __fmul_rn(__fmul_rn(bb.x, dy), dy));
                const float power = __fmaf_rn(sum, -0.5f, -__fmul_rn(__fmul_rn(a.w, dx), dy));
                if (power > 0.0f) continue;
                if (power < -bb.z) continue;   // alpha < 1/255 with margin: same outcome, no exp
                const float alpha = fminf(0.99f, __fmul_rn(bb.y, expf(power)));
                if (alpha < 1.0f / 255.0f) continue;
                const float test_T = __fmul_rn(T, __fsub_rn(1.0f, alpha));
                if (test_T < 0.0001f) {
                    done = true;
                    continue;
                }
                const float4 c = sC[st][j];   // r, g, b, sin(theta)
                const float4 d = sD[st][j];   // scale xyz, rot.r
                const float4 e = sE[st][j];   // rot xyz, theta
                C0 = __fmaf_rn(__fmul_rn(c.x, alpha), T, C0);
                C1 = __fmaf_rn(__fmul_rn(c.y, alpha), T, C1);
                C2 = __fmaf_rn(__fmul_rn(c.z, alpha), T, C2);
                Dacc = __fmaf_rn(__fmul_rn(bb.w, alpha), T, Dacc);
                S0 = __fmaf_rn(__fmul_rn(d.x, alpha), T, S0);
                S1 = __fmaf_rn(__fmul_rn(d.y, alpha), T, S1);
                S2 = __fmaf_rn(__fmul_rn(d.z, alpha), T, S2);

                // r = mix(identity, q, w) with q = (w=rot.r, x, y, z), w = alpha*T
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
                last_contributor = base_ord + (uint32_t)j + 1u;   // ordinal in the tile's list
            }
        }
    }
    cp_async_wait_all();

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
    const int T = p.gx * ((a.H + kTile - 1) / kTile);
    composite_fwd_kernel<<<T, kTilePix, 0, stream>>>(p);
    VEGS_LAUNCH_CHECK("composite_fwd", stream, a.debug);
    return VEGS_OK;
}

}  // namespace vegs
