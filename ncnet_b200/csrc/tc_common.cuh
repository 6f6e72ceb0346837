// sm_100a building blocks shared by the tensor-core kernels: mbarrier, TMA (bulk + tensor tile),
// UMMA shared-memory descriptors, tcgen05.mma / commit / ld, and the host-side tensor-map encoder.
#pragma once
#include "common.cuh"

namespace ncnet {
namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// A shared-window address the compiler must keep in a register: the plain conversion is "pure", so it gets
// re-materialised (S2UR SR_CgaCtaId + 3 uniform ops, ~60 cycles of latency) next to every use in unrolled code.
__device__ __forceinline__ uint32_t smem_u32_pinned(const void *p) {
    uint32_t r;
    asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"((uint32_t)__cvta_generic_to_shared(p)));
    return r;
}

// true in exactly one lane of a converged warp; ptxas knows code under it runs single-lane, so
// tcgen05/TMA operands move to uniform registers without a per-value "waterfall" loop
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred P;\n\t"
        "elect.sync _|P, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t}"
        : "=r"(pred));
    return pred != 0;
}

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// bounded wait: a protocol bug traps (the launch reports an error) instead of hanging the GPU.  The first probe
// carries no clock read: the waiting lanes of the tensor-core kernels run ~10 cycles per instruction.
__device__ __forceinline__ bool mbar_try(uint32_t addr, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait_u32(uint32_t addr, uint32_t parity) {
    if (mbar_try(addr, parity)) return;
    const long long t0 = clock64();
    for (;;) {
        if (mbar_try(addr, parity)) return;
        if (clock64() - t0 > 8000000000ll) __trap();
    }
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) { mbar_wait_u32(smem_u32(bar), parity); }
// long waits (an epilogue waiting for a whole main loop): back off between probes -- every probe is a shared-memory
// pipeline transaction, and four spinning warps slow the co-resident CTA's epilogue loads/stores severalfold
__device__ __forceinline__ void mbar_wait_sleep(uint64_t *bar, uint32_t parity, unsigned ns) {
    const uint32_t addr = smem_u32(bar);
    if (mbar_try(addr, parity)) return;
    const long long t0 = clock64();
    for (;;) {
        __nanosleep(ns);
        if (mbar_try(addr, parity)) return;
        if (clock64() - t0 > 8000000000ll) __trap();
    }
}
// Shared-memory accesses by 32-bit shared-window address.  Through a generic pointer the compiler re-derives the
// window base (S2UR SR_CgaCtaId, ~60 cycles) next to every access of an unrolled epilogue; these forms cost one LDS/STS.
__device__ __forceinline__ void sts_f32(uint32_t addr, float v) {
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ float lds_f32(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void sts_v4(uint32_t addr, float4 v) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float4 lds_v4(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ uint4 lds_u4(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}
// 1-D bulk copy global -> shared through the TMA engine (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_g2s_u32(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// UMMA shared-memory descriptor, SWIZZLE_NONE, version 1 (sm_100).
//   K-major : SBO = stride between 8-row groups, LBO = stride between the 16-byte K chunks
//   MN-major: SBO = stride between 8-element M/N groups, LBO = stride between groups of 8 K rows
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
// The A-operand collector: consecutive MMAs that read the SAME A tile (different B / D) can keep it in the tensor core's
// collector buffer -- the first one fetches it (fill), the others reuse it (use ... lastuse).  Measured on B200
// (scripts/umma_rate.cu, groups of 5): M128 x N x K16 drops from 39/40/44/52 cycles to 16/22/28/41 at N = 16/32/48/80,
// the CTA-pair M256 x N80 from 46 to 41.  MODE: 0 = no hint, 1 = fill, 2 = use, 3 = lastuse.
template <int MODE, bool ACC1>      // ACC1: the accumulate flag is the constant 1 (no predicate set-up in the single issuing lane)
__device__ __forceinline__ void umma_f16_coll(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
#define NC_UMMA_COLL(Q)                                                                                                                     \
    do {                                                                                                                                    \
        if (ACC1)                                                                                                                           \
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::1.kind::f16" Q " [%0], %1, %2, %3, p;\n\t}"  \
                         ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc) : "memory");                                                     \
        else                                                                                                                                \
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16" Q " [%0], %1, %2, %3, p;\n\t}" \
                         ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");                                           \
    } while (0)
    if (MODE == 1) NC_UMMA_COLL(".collector::a::fill");
    else if (MODE == 2) NC_UMMA_COLL(".collector::a::use");
    else if (MODE == 3) NC_UMMA_COLL(".collector::a::lastuse");
    else NC_UMMA_COLL("");
#undef NC_UMMA_COLL
}
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
// split form: issue several loads, then one wait per destination array (the wait names the registers so the
// compiler cannot move their first use above it)
__device__ __forceinline__ void tmem_ld16_issue(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld16_wait(uint32_t (&r)[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]),
                   "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
                 :: "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float (&v)[8]) {
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}

}  // namespace tc
}  // namespace ncnet
