// bulk_tile.cuh -- PTX helpers of the streaming kernels (softmax_ce.cu, finalize_map in estep_kernel.cu): a
// tile of whole pixels moves global -> shared and shared -> global as ONE bulk copy of the copy engine
// (cp.async.bulk, completion on an mbarrier / a bulk group), so that 84-byte pixels cost no 4-byte accesses.
#pragma once
#include <stdint.h>

namespace wl {

__device__ __forceinline__ uint32_t ce_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void ce_mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(ce_smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void ce_mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(ce_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void ce_mbar_wait(uint64_t *bar, uint32_t parity) {
    const uint32_t a = ce_smem_u32(bar);
    uint32_t ok;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                     : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void ce_bulk_g2s(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
                 ::"r"(ce_smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(ce_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void ce_bulk_s2g(void *gdst, const void *smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n"
                 ::"l"(gdst), "r"(ce_smem_u32(smem_src)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;\n" ::: "memory");
}
__device__ __forceinline__ float ce_exp(float x) {           // exp(x) for x <= 0: ex2.approx, ~2^-22 relative
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;\n" : "=f"(y) : "f"(x * 1.4426950408889634f));
    return y;
}


}  // namespace wl
