// Blackwell / Hopper asynchronous bulk copies (PTX `cp.async.bulk`, SASS UBLKCP) with mbarrier completion: the 1-D
// form of TMA. A thread hands the copy engine a contiguous run of bytes (global -> shared or shared -> global) and goes
// on; loads signal an mbarrier by byte count (SASS SYNCS), stores are tracked in the issuing thread's bulk groups.
// Addresses and sizes must be multiples of 16 bytes.
#pragma once
#include <stdint.h>

namespace bulk {

__device__ __forceinline__ unsigned smem_addr(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned bar, unsigned arrivals) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(arrivals) : "memory");
}
// makes freshly initialised barriers visible to the copy engine (async proxy); follow with a CTA / warp barrier
__device__ __forceinline__ void mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

// one arrival that also announces `bytes` of copy traffic the phase has to wait for
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(unsigned bar, unsigned parity) {
    unsigned ok;
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0u;
}
// waits for the phase with the given parity; a copy that never lands traps instead of hanging the GPU
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    unsigned spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spins > (1u << 20)) __trap();
    }
}

// global -> shared, completion counted on `bar` in bytes
__device__ __forceinline__ void load(unsigned dst_smem, const void *src_global, unsigned bytes, unsigned bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_smem),
                 "l"(src_global), "r"(bytes), "r"(bar)
                 : "memory");
}
// shared -> global, joins the calling thread's current bulk group
__device__ __forceinline__ void store(void *dst_global, unsigned src_smem, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_global), "r"(src_smem), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// until the calling thread's bulk stores have finished READING shared memory (the source may then be overwritten)
__device__ __forceinline__ void wait_stores_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// orders this thread's ordinary shared-memory writes before copies it issues afterwards (generic -> async proxy)
__device__ __forceinline__ void fence_smem_writes() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ float4 lds128(unsigned addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts128(unsigned addr, float a, float b, float c, float d) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

}  // namespace bulk
