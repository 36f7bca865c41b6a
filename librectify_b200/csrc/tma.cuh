// tma.cuh — thin inline-PTX wrappers for the sm_100a bulk-tensor copy engine (TMA) and mbarriers.
// Used by the stencil kernels to stage halo tiles: one elected thread issues cp.async.bulk.tensor, the copy engine
// zero-fills everything outside the image (which is exactly the border rule of the reference's filters), and the
// CTA waits on an mbarrier phase instead of spending LSU issue slots and address arithmetic on the halo.
#pragma once
#include <cuda.h>
#include <stdint.h>

namespace lrb {

__device__ __forceinline__ unsigned smem_u32(const void * p)
{
    return static_cast<unsigned>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t * bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}

// make the initialised barrier visible to the async proxy (the copy engine)
__device__ __forceinline__ void mbar_fence_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t * bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}

__device__ __forceinline__ bool mbar_try_wait(uint64_t * bar, unsigned parity)
{
    unsigned ok;
    asm volatile(
        "{\n"
        " .reg .pred p;\n"
        " mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        " selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}

__device__ __forceinline__ void mbar_wait(uint64_t * bar, unsigned parity)
{
    while (!mbar_try_wait(bar, parity)) { }
}

// 3-D tiled load: coordinates are (x = column, y = row, z = frame slot), innermost first; out-of-range elements
// arrive as zeros.
__device__ __forceinline__ void tma_load_3d(void * smem_dst, const CUtensorMap * map, int x, int y, int z, uint64_t * bar)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        :: "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar))
        : "memory");
}

__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap * map)
{
    asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(map)) : "memory");
}

} // namespace lrb
