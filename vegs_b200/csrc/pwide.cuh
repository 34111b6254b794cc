// Shared by the two per-Gaussian ("P-wide") kernels, preprocess.cu and preprocess_bwd.cu: how a
// warp's 32 consecutive Gaussians get their inputs into shared memory, and the optional fused
// activations.
//
// Shared-memory staging area of one warp (WarpStage):
//   sh   SH coefficients of the warp's Gaussians, one row per lane.  Three layouts:
//          kShBulkRows  [P,M,3] input whose rows are a multiple of 16 bytes (M = 4, 8, 12, 16): each
//                       row is bulk-copied to a row of kShRowStride = 52 floats (208 B: 16-byte
//                       aligned rows, and 16-byte reads of a quarter warp fall into 8 distinct
//                       4-bank groups because 13 is odd) and read with 16-byte loads;
//          kShSplit     f_dc [P,1,3] + f_rest [P,M-1,3] (the fused "cat"): the warp's rest block
//                       (32 x 12(M-1) bytes, contiguous) and dc block (384 B) are bulk-copied
//                       densely; row strides 3(M-1) (45 at M = 16) and 3 floats are odd, so scalar
//                       reads with lane = row are conflict-free;
//          kShLdg       anything else: warp-cooperative 16-byte loads into rows of 3M|1 floats.
//   in   the warp's means (96 floats), scales (96), rotations (128), opacities (32), as they lie
//        in global memory.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "common.cuh"
#include "bulk.cuh"

namespace vegs {

constexpr int kPwThreads = 128;
constexpr int kPwWarps = kPwThreads / 32;
constexpr int kMaxStageCoeffs = 16;                   // SH degree <= 3
constexpr int kShRowStride = 52;                      // floats
constexpr int kShWarpFloats = 32 * kShRowStride;      // 1664 floats = 6656 B
constexpr int kInMeans = 0, kInScales = 96, kInRots = 192, kInOpac = 320, kInWarpFloats = 352;

enum : int { kShLdg = 0, kShBulkRows = 1, kShSplit = 2 };

struct __align__(16) WarpStage {
    float sh[kShWarpFloats];
    float in[kInWarpFloats];
};

__device__ __forceinline__ float4 ldg_stream4(const float4* p) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p));
    return v;
}

// ---- fused activations (VEGS_ACT_* in include/vegs_b200.h), float operations as torch's CUDA
// kernels perform them: sigmoid = 1 / (1 + exp(-x)), exp, and F.normalize = x / max(||x||, 1e-12)
__device__ __forceinline__ float act_sigmoid(float x) { return __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-x))); }

__device__ __forceinline__ float4 act_normalize(float4 q, float& denom) {
    const float n = sqrtf(__fmaf_rn(q.w, q.w, __fmaf_rn(q.z, q.z, __fmaf_rn(q.y, q.y, __fmul_rn(q.x, q.x)))));
    denom = fmaxf(n, 1e-12f);
    return make_float4(__fdiv_rn(q.x, denom), __fdiv_rn(q.y, denom), __fdiv_rn(q.z, denom), __fdiv_rn(q.w, denom));
}

// The raw per-Gaussian inputs of one lane.
struct LaneInputs {
    float mx, my, mz;
    float sx, sy, sz;      // as stored (raw when VEGS_ACT_SCALE_EXP)
    float4 q;              // as stored (raw when VEGS_ACT_ROT_NORMALIZE)
    float op;              // as stored (raw when VEGS_ACT_OPACITY_SIGMOID)
};

// Start the bulk copies of the warp's input blocks (one lane issues; `bar` must expect 1 arrival).
// Only for a full warp (32 Gaussians in range) with 16-byte aligned arrays.
// `in` = the warp's kInWarpFloats-float input area; `extra_bytes` = what the caller copies to the
// same barrier on top of these blocks (its arrive.expect_tx is done here, once).
__device__ __forceinline__ void issue_input_copies(float* in, uint64_t* bar, const float* means,
                                                   const float* scales, const float* rots, const float* opac,
                                                   int warp_base, uint32_t extra_bytes = 0u) {
    uint32_t bytes = 384u + (opac ? 128u : 0u) + (scales ? 384u + 512u : 0u) + extra_bytes;
    mbar_arrive_expect_tx(bar, bytes);
    bulk_g2s(in + kInMeans, means + 3 * (size_t)warp_base, 384u, bar);
    if (opac) bulk_g2s(in + kInOpac, opac + (size_t)warp_base, 128u, bar);
    if (scales) {
        bulk_g2s(in + kInScales, scales + 3 * (size_t)warp_base, 384u, bar);
        bulk_g2s(in + kInRots, rots + 4 * (size_t)warp_base, 512u, bar);
    }
}

__device__ __forceinline__ LaneInputs read_inputs_smem(const float* in, int lane, bool have_sr, bool have_op) {
    LaneInputs r;
    r.mx = in[kInMeans + 3 * lane + 0];
    r.my = in[kInMeans + 3 * lane + 1];
    r.mz = in[kInMeans + 3 * lane + 2];
    r.sx = r.sy = r.sz = 0.f;
    r.q = make_float4(1.f, 0.f, 0.f, 0.f);
    r.op = 0.f;
    if (have_sr) {
        r.sx = in[kInScales + 3 * lane + 0];
        r.sy = in[kInScales + 3 * lane + 1];
        r.sz = in[kInScales + 3 * lane + 2];
        r.q = *reinterpret_cast<const float4*>(&in[kInRots + 4 * lane]);
    }
    if (have_op) r.op = in[kInOpac + lane];
    return r;
}

__device__ __forceinline__ LaneInputs read_inputs_global(const float* means, const float* scales, const float* rots,
                                                         const float* opac, int idx, bool rots_vec_ok) {
    LaneInputs r;
    r.mx = means[3 * (size_t)idx + 0];
    r.my = means[3 * (size_t)idx + 1];
    r.mz = means[3 * (size_t)idx + 2];
    r.sx = r.sy = r.sz = 0.f;
    r.q = make_float4(1.f, 0.f, 0.f, 0.f);
    r.op = 0.f;
    if (scales != nullptr) {
        r.sx = scales[3 * (size_t)idx + 0];
        r.sy = scales[3 * (size_t)idx + 1];
        r.sz = scales[3 * (size_t)idx + 2];
        if (rots_vec_ok) r.q = reinterpret_cast<const float4*>(rots)[idx];
        else r.q = make_float4(rots[4 * (size_t)idx + 0], rots[4 * (size_t)idx + 1], rots[4 * (size_t)idx + 2],
                               rots[4 * (size_t)idx + 3]);
    }
    if (opac != nullptr) r.op = opac[idx];
    return r;
}

// kShLdg staging: rows of `rowlen` floats of the warp's Gaussians into srow with stride rowlen|1;
// only rows of lanes in `vis` are fetched (16-byte chunks that touch such a row).
__device__ __forceinline__ void stage_sh_ldg(float* srow, const float* shs, int rowlen, int warp_base, int P,
                                             unsigned vis, bool visible, int idx, int lane, bool vec_ok, int need) {
    const int stride = rowlen | 1;
    const bool full = (warp_base + 32 <= P);
    if (full && vec_ok) {
        const float4* src = reinterpret_cast<const float4*>(shs + (size_t)warp_base * rowlen);
        const int nvec = 8 * rowlen;  // 32 rows * rowlen floats / 4
        for (int f = lane; f < nvec; f += 32) {
            const int e0 = 4 * f;
            const int r0 = e0 / rowlen;
            const int r1 = (e0 + 3) / rowlen;
            if (((vis >> r0) | (vis >> r1)) & 1u) {
                const float4 v = ldg_stream4(src + f);
                const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const int e = e0 + k;
                    const int r = e / rowlen;
                    srow[r * stride + (e - r * rowlen)] = vv[k];
                }
            }
        }
    } else if (visible) {
        const float* src = shs + (size_t)idx * rowlen;
        for (int k = 0; k < need && k < rowlen; k++) srow[lane * stride + k] = src[k];
    }
}

}  // namespace vegs
