// orbx_tma.cuh — thin PTX wrappers for the Blackwell/Hopper bulk-tensor copy engine (TMA) and the
// mbarrier it signals. Device side only; tensor maps are encoded on the host (engine.cu) with
// cuTensorMapEncodeTiled and live in a small device array indexed by pyramid level.
//
// Usage in a kernel (one elected thread issues, everybody waits):
//     __shared__ __align__(128) uint8_t tile[ROWS * 128];
//     __shared__ __align__(8) uint64_t bar;
//     if (tid == 0) { orbx_mbar_init(&bar, 1); orbx_fence_mbar_init(); }
//     __syncthreads();
//     if (tid == 0) { orbx_mbar_expect_tx(&bar, bytes); orbx_tma_load_3d(tile, map, x, y, img, &bar); }
//     orbx_mbar_wait(&bar, 0);
// Out-of-bounds box elements (negative coordinates, beyond the level) arrive as zeros.
#pragma once
#include <cuda.h>
#include <stdint.h>

__device__ __forceinline__ uint32_t orbx_smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void orbx_mbar_init(uint64_t* bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(orbx_smem_addr(bar)), "r"(count));
}

// Make the barrier initialisation (a generic-proxy shared-memory write) visible to the async (TMA) proxy.
// Deliberately NOT fence.mbarrier_init.release.cluster: no other CTA touches these barriers, and a
// cluster-scope fence compiles to CCTL.IVALL, which invalidates the SM's whole L1 on every call (seen in ncu).
__device__ __forceinline__ void orbx_fence_mbar_init()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void orbx_mbar_expect_tx(uint64_t* bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(orbx_smem_addr(bar)), "r"(bytes) : "memory");
}

__device__ __forceinline__ void orbx_mbar_wait(uint64_t* bar, unsigned phase)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(orbx_smem_addr(bar)),
        "r"(phase)
        : "memory");
}

// 3-D tiled load: box of the tensor map at element coordinates (x, y, z) -> shared memory; completes on `bar`.
__device__ __forceinline__ void orbx_tma_load_3d(void* dst, const CUtensorMap* map, int x, int y, int z, uint64_t* bar)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
            orbx_smem_addr(dst)),
        "l"(map), "r"(orbx_smem_addr(bar)), "r"(x), "r"(y), "r"(z)
        : "memory");
}

// The same operations on 32-bit shared-window addresses (orbx_smem_addr taken ONCE by the caller): every conversion of a
// generic pointer costs a read of the CTA's window base (S2R SR_CgaCtaId + LEA) where the compiler cannot fold it away.
__device__ __forceinline__ void orbx_mbar_init_a(uint32_t bar, unsigned count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)); }
__device__ __forceinline__ void orbx_mbar_expect_tx_a(uint32_t bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void orbx_mbar_wait_a(uint32_t bar, unsigned phase)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP_A:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE_A;\n"
        "bra WAIT_LOOP_A;\n"
        "WAIT_DONE_A:\n"
        "}\n" ::"r"(bar),
        "r"(phase)
        : "memory");
}
__device__ __forceinline__ void orbx_tma_load_3d_a(uint32_t dst, const CUtensorMap* map, int x, int y, int z, uint32_t bar)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
                 "l"(map), "r"(bar), "r"(x), "r"(y), "r"(z)
                 : "memory");
}

__device__ __forceinline__ void orbx_tma_prefetch_desc(const CUtensorMap* map)
{
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
