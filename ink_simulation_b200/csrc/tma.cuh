// tma.cuh — the few sm_90+/sm_100a primitives the tile-staging kernels need: an mbarrier in shared memory and 1-D bulk
// asynchronous copies global -> shared executed by the TMA unit (cp.async.bulk, SASS UBLKCP), which deposit their byte
// count on the mbarrier when they land.  No tensor maps: every staged piece here is a CONTIGUOUS run of z-rows of a grid
// component ((j*X + i)*Z + k layout: x-adjacent rows are adjacent in memory), 16-byte aligned when Z*sizeof(Real) % 16 == 0.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ink {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned int arrivals) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(arrivals) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");   // make the initialised barrier visible to the async proxy
}
// one arrival that also announces `bytes` of asynchronous traffic still to land on this barrier
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, unsigned int bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned int parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "INK_MBAR_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra INK_MBAR_DONE;\n"
        "bra INK_MBAR_WAIT;\n"
        "INK_MBAR_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}
// TMA 1-D bulk copy: `bytes` (multiple of 16) from global `src` (16-byte aligned) to shared `dst` (16-byte aligned);
// completion is signalled on `bar` as `bytes` of transaction count
__device__ __forceinline__ void tma_load_1d(void* dst_smem, const void* src_gmem, unsigned int bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)), "l"(src_gmem),
                 "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

}  // namespace ink
