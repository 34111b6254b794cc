// Optimiser step on the view-parallel step's buffers, one pass.
//
// No counterpart inside the reference's extensions: it replaces what the reference's CALLER does
// after the backward pass with separate torch kernels over all Gaussians:
//   * `gaussians.optimizer.step()` = torch.optim.Adam over six parameter tensors with their own
//     learning rates (scene/gaussian_model.py:154-172, train.py:319) - dozens of element-wise
//     kernels (or torch's multi-tensor kernels) reading and writing param / grad / exp_avg /
//     exp_avg_sq;
//   * `optimizer.zero_grad()` (train.py:320);
//   * folding the step's densification statistics into the running ones
//     (scene/gaussian_model.py:411-413, train.py:300-301) after the all-reduce.
// Here one launch walks a table of up to VEGS_ADAM_MAX_SEGMENTS contiguous segments.  A parameter
// segment reads grad, exp_avg, exp_avg_sq, param once and writes the three back (plus, on
// request, a zero into grad): 28 B per element, against 60+ B for the unfused sequence.  A
// statistics segment adds (or maxes) the step's reduced value into the running accumulator.
// Arithmetic is torch.optim.Adam's (betas, eps, no weight decay, no amsgrad), in the operation
// forms of torch's kernels:
//   m += (1 - b1)(g - m);  v = b2 v + (1 - b2) g g
//   p -= (lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
// Roofline: HBM, 28 B per parameter element.
#include <cmath>
#include "common.cuh"

namespace vegs {

namespace {

struct AdamSeg {
    float* param;        // kind 0: parameter; kinds 1, 2: the running accumulator
    float* grad;         // kind 0: gradient; kinds 1, 2: this step's (reduced) value
    float* exp_avg;
    float* exp_avg_sq;
    unsigned long long begin;   // first flat element index of the segment
    float step_size;     // lr / (1 - b1^t)
    int kind;            // 0 Adam, 1 accumulate (+=), 2 maximum
};

struct AdamParams {
    AdamSeg seg[VEGS_ADAM_MAX_SEGMENTS];
    unsigned long long group_begin[VEGS_ADAM_MAX_SEGMENTS];   // first 4-element group of each segment
    int nseg;
    unsigned long long total;
    float beta1, beta2, eps, sqrt_bc2;
    float omb1, omb2;    // 1 - beta, formed in double as torch forms them
    int zero_grad;
};

__device__ __forceinline__ void adam_elem(float& p, float g, float& m, float& v, const AdamParams& a, float step_size) {
    // lerp_(g, 1 - b1); mul_(b2).addcmul_(g, g, 1 - b2); sqrt() / sqrt(bc2) + eps; addcdiv_(m, denom, -step_size)
    m = __fmaf_rn(a.omb1, g - m, m);
    v = __fmaf_rn(a.omb2 * g, g, __fmul_rn(v, a.beta2));
    const float denom = __fadd_rn(__fdiv_rn(sqrtf(v), a.sqrt_bc2), a.eps);
    p = __fmaf_rn(-step_size, __fdiv_rn(m, denom), p);
}

// One thread per group of four consecutive elements of a segment (16-byte accesses when every
// array is 16-byte aligned; a group never straddles two segments: groups restart per segment).
template <bool kVec>
__global__ void __launch_bounds__(256) adam_kernel(const AdamParams a, unsigned long long total_groups) {
    for (unsigned long long gidx = blockIdx.x * 256ull + threadIdx.x; gidx < total_groups;
         gidx += (unsigned long long)gridDim.x * 256ull) {
        int s = 0;
#pragma unroll
        for (int k = 1; k < VEGS_ADAM_MAX_SEGMENTS; k++)
            if (k < a.nseg && gidx >= a.group_begin[k]) s = k;
        const AdamSeg& sg = a.seg[s];
        const unsigned long long e0 = (gidx - a.group_begin[s]) * 4ull;
        const unsigned long long n = ((s + 1 < a.nseg) ? a.seg[s + 1].begin : a.total) - sg.begin;
        const int cnt = (int)((n - e0) < 4ull ? (n - e0) : 4ull);
        if (kVec && cnt == 4) {
            float4 g = *reinterpret_cast<const float4*>(sg.grad + e0);
            float4 p = *reinterpret_cast<const float4*>(sg.param + e0);
            if (sg.kind == 0) {
                float4 m = *reinterpret_cast<const float4*>(sg.exp_avg + e0);
                float4 v = *reinterpret_cast<const float4*>(sg.exp_avg_sq + e0);
                adam_elem(p.x, g.x, m.x, v.x, a, sg.step_size);
                adam_elem(p.y, g.y, m.y, v.y, a, sg.step_size);
                adam_elem(p.z, g.z, m.z, v.z, a, sg.step_size);
                adam_elem(p.w, g.w, m.w, v.w, a, sg.step_size);
                *reinterpret_cast<float4*>(sg.exp_avg + e0) = m;
                *reinterpret_cast<float4*>(sg.exp_avg_sq + e0) = v;
            } else if (sg.kind == 1) {
                p.x += g.x; p.y += g.y; p.z += g.z; p.w += g.w;
            } else {
                p.x = fmaxf(p.x, g.x); p.y = fmaxf(p.y, g.y); p.z = fmaxf(p.z, g.z); p.w = fmaxf(p.w, g.w);
            }
            *reinterpret_cast<float4*>(sg.param + e0) = p;
            if (a.zero_grad) *reinterpret_cast<float4*>(sg.grad + e0) = make_float4(0.f, 0.f, 0.f, 0.f);
        } else {
            for (int k = 0; k < cnt; k++) {
                const float g = sg.grad[e0 + k];
                float p = sg.param[e0 + k];
                if (sg.kind == 0) {
                    float m = sg.exp_avg[e0 + k], v = sg.exp_avg_sq[e0 + k];
                    adam_elem(p, g, m, v, a, sg.step_size);
                    sg.exp_avg[e0 + k] = m;
                    sg.exp_avg_sq[e0 + k] = v;
                } else if (sg.kind == 1) {
                    p += g;
                } else {
                    p = fmaxf(p, g);
                }
                sg.param[e0 + k] = p;
                if (a.zero_grad) sg.grad[e0 + k] = 0.0f;
            }
        }
    }
}

}  // namespace

}  // namespace vegs

extern "C" int vegs_adam_step(const VegsAdamArgs* args, void* stream_) {
    using namespace vegs;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    if (!args || args->num_segments < 0 || args->num_segments > VEGS_ADAM_MAX_SEGMENTS || args->step < 1) {
        set_error("vegs_adam_step: bad argument (segments 0..%d, step >= 1)", VEGS_ADAM_MAX_SEGMENTS);
        return VEGS_EINVAL;
    }
    AdamParams a;
    a.nseg = 0;
    a.beta1 = (float)args->beta1; a.beta2 = (float)args->beta2; a.eps = (float)args->eps;
    const double bc1 = 1.0 - std::pow(args->beta1, (double)args->step);
    const double bc2 = 1.0 - std::pow(args->beta2, (double)args->step);
    a.sqrt_bc2 = (float)std::sqrt(bc2);
    a.omb1 = (float)(1.0 - args->beta1);
    a.omb2 = (float)(1.0 - args->beta2);
    a.zero_grad = args->zero_grad;
    for (int i = 0; i < VEGS_ADAM_MAX_SEGMENTS; i++) a.group_begin[i] = 0;
    unsigned long long flat = 0, groups = 0;
    bool vec = true;
    for (int i = 0; i < args->num_segments; i++) {
        const VegsAdamSegment& s = args->segments[i];
        if (s.count == 0) continue;
        if (!s.param || !s.grad || (s.kind == 0 && (!s.exp_avg || !s.exp_avg_sq)) || s.kind < 0 || s.kind > 2 || s.count < 0) {
            set_error("vegs_adam_step: segment %d has a null pointer, a negative count or a bad kind", i);
            return VEGS_EINVAL;
        }
        AdamSeg& d = a.seg[a.nseg];
        d.param = s.param; d.grad = s.grad; d.exp_avg = s.exp_avg; d.exp_avg_sq = s.exp_avg_sq;
        d.begin = flat; d.kind = s.kind;
        d.step_size = (float)((double)s.lr / bc1);
        a.group_begin[a.nseg] = groups;
        auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };
        vec = vec && al16(s.param) && al16(s.grad) && al16(s.exp_avg) && al16(s.exp_avg_sq);
        flat += (unsigned long long)s.count;
        groups += ((unsigned long long)s.count + 3ull) / 4ull;
        a.nseg++;
    }
    a.total = flat;
    if (groups == 0) return VEGS_OK;
    const unsigned long long want = (groups + 255ull) / 256ull;
    const int blocks = (int)(want < 148ull * 16ull ? want : 148ull * 16ull);
    if (vec) adam_kernel<true><<<blocks, 256, 0, stream>>>(a, groups);
    else adam_kernel<false><<<blocks, 256, 0, stream>>>(a, groups);
    VEGS_LAUNCH_CHECK("adam_step", stream, false);
    return VEGS_OK;
}
