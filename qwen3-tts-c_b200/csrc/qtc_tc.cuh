/*
 * qtc_tc.cuh - tcgen05 / TMA / mbarrier building blocks shared by the tensor-core kernels of this library
 * (qtc_gemm_tc.cu: batched projections; qtc_conv_tc.cu: the vocoder's convolutions as implicit GEMMs).
 * Operands are K-major bf16 tiles in SWIZZLE_128B layout (rows of 64 bf16 = 128 bytes, 8-row groups 1024 bytes apart),
 * filled by cp.async.bulk.tensor; accumulators are fp32 in TMEM; M = 128.
 */
#pragma once
#include <cuda.h>
#include <stdint.h>

#define TC_BM 128           /* UMMA M */

__device__ __forceinline__ uint32_t tc_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void tc_mbar_init(uint64_t *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tc_smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void tc_mbar_expect_tx(uint64_t *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tc_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tc_mbar_wait(uint64_t *bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "TCW_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra TCD_%=;\n"
        "bra TCW_%=;\n"
        "TCD_%=:\n"
        "}\n" ::"r"(tc_smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tc_tma_2d(void *dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(tc_smem_u32(dst)), "l"(map), "r"(tc_smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
/* K-major operand tile in SWIZZLE_128B layout: rows of 128 bytes, 8-row groups 1024 bytes apart */
__device__ __forceinline__ uint64_t tc_smem_desc(uint32_t addr) {
    uint64_t d = 0;
    d |= (uint64_t)((addr & 0x3ffffu) >> 4);          /* start address, 16-byte units           [0,14)  */
    d |= (uint64_t)(1024 >> 4) << 32;                 /* stride byte offset (8 rows x 128 B)    [32,46) */
    d |= 1ull << 46;                                  /* descriptor version (sm_100)            [46,48) */
    d |= 2ull << 61;                                  /* layout: SWIZZLE_128B                   [61,64) */
    return d;
}
/* kind::f16 instruction descriptor: D f32, A/B bf16, both K-major, M = 128, N = n (multiple of 16, <= 256) */
__device__ __forceinline__ uint32_t tc_idesc(int n) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
}
__device__ __forceinline__ void tc_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(tc_smem_u32(bar)) : "memory");
}


/* bf16 [rows, cols] row-major -> TMA boxes of [box_rows x 64 columns], 128-byte swizzle, zero fill out of bounds (cached per
 * pointer / shape; qtc_gemm_tc_forget drops the maps of a freed buffer) */
const CUtensorMap *qtc_tmap_bf16(const void *ptr, int rows, int cols, int box_rows);
