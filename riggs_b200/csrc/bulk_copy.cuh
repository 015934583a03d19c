// 1-D bulk copies through the TMA engine (cp.async.bulk, sm_90+/sm_100a): one elected thread moves a contiguous
// block between global and shared memory while the other threads of the CTA compute.  Used to stage the per-CTA
// block of SH coefficients (256 Gaussians x 180 B = 45 KB contiguous) that per-thread loads would fetch with a
// 180-byte stride (~27 sectors per request measured by ncu), and to write the SH gradients back the same way.
// Addresses and sizes must be multiples of 16 bytes.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace rg {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

// Arms the barrier for `bytes` of incoming data and starts the global -> shared bulk copy (one thread calls this).
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// Upper bound (ns) on how long the hardware may keep a waiting thread suspended inside one try_wait before it returns
// false.  Without the hint the limit is short: the warps of the MLP chain kernel that wait for their next quarter polled
// ~26 times per wait (16 M TRYWAIT instructions per launch) and took issue slots from the warps on the critical path.
constexpr uint32_t kMbarSuspendHint = 0x989680;
// Potentially-blocking probe and plain arrival, for warp-to-warp hand-offs inside a CTA (one lane arrives after a
// __syncwarp; arrive has release, try_wait acquire semantics at CTA scope).
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity), "r"(kMbarSuspendHint)
        : "memory");
    return ok != 0;
}
// All threads that read the staged block wait here (phase parity of the barrier's current use).  A plain loop over the
// probe: no asm labels, so any number of inlined copies can live in one kernel.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {}
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// Shared -> global bulk store (one thread calls this after a __syncthreads that follows the writes to shared memory).
__device__ __forceinline__ void bulk_store(void* gmem_dst, const void* smem_src, uint32_t bytes) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)),
                 "r"(bytes)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

// Shared -> global bulk REDUCTION: global[i] += shared[i] for `bytes / 4` floats, performed by the copy engine at the
// L2 (no read-back into the SM).  Same calling convention as bulk_store.
__device__ __forceinline__ void bulk_reduce_add_f32(void* gmem_dst, const void* smem_src, uint32_t bytes) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(gmem_dst),
                 "r"(smem_u32(smem_src)), "r"(bytes)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

// The issuing thread must not let shared memory go away (CTA exit) before the engine has read it.
__device__ __forceinline__ void bulk_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }

}  // namespace rg
