// Asynchronous bulk copies (the 1-D form of TMA: PTX cp.async.bulk, SASS UBLKCP) and the
// mbarrier that tracks their completion.  sm_100a only; no tensor map is needed for the 1-D form.
//
// Used where a kernel moves CONTIGUOUS bytes between global and shared memory:
//   * the spherical-harmonics rows of a warp's 32 Gaussians in the per-Gaussian kernels
//     (preprocess.cu, preprocess_bwd.cu): 192-byte rows (or one 5760-byte block for the split
//     f_dc / f_rest layout), and the dL/dSH rows on the way back - as a plain bulk store or,
//     for fused gradient accumulation, as a bulk REDUCTION (cp.reduce.async.bulk ... .add.f32)
//     that performs the read-modify-write in L2 without bringing the old value to the SM;
//   * the contiguous list slices of the compositing kernels (composite_fwd.cu, composite_bwd.cu).
// Requirements of the instruction: source, destination 16-byte aligned, size a multiple of 16.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace vegs {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}

// One barrier object = 8 bytes of shared memory, 8-byte aligned.
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t arrivals) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(arrivals) : "memory");
}

// After mbar_init, before the barrier is used by the async proxy (bulk copies).
__device__ __forceinline__ void mbar_init_fence() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// Arrive and announce `tx_bytes` more bytes that bulk copies will deliver to this barrier.
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t tx_bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(tx_bytes)
                 : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// Block until the phase with the given parity has completed (hardware sleep, not a spin on memory).
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

// global -> shared, completion (bytes) signalled on `bar`.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// Generic-proxy writes to shared memory (st.shared) must be made visible to the async proxy
// before a bulk copy reads them.
__device__ __forceinline__ void fence_smem_to_async_proxy() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// shared -> global (plain store), tracked by the thread's bulk async-group.
__device__ __forceinline__ void bulk_s2g(void* gmem_dst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst),
                 "r"(smem_u32(smem_src)), "r"(bytes)
                 : "memory");
}

// shared -> global, element-wise float ADD into the destination (performed at L2).
__device__ __forceinline__ void bulk_s2g_add_f32(void* gmem_dst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(gmem_dst),
                 "r"(smem_u32(smem_src)), "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }

// Wait until the issuing thread's bulk groups have finished READING their shared-memory source
// (the source may then be reused or the CTA may exit; the global writes complete on their own).
__device__ __forceinline__ void bulk_wait_read_all() {
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

}  // namespace vegs
