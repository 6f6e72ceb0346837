// Micro-benchmark: issue rate of tcgen05.mma (kind::f16, cta_group::1) as a function of tile shape, operand layout /
// swizzle mode, operand source (smem descriptor or TMEM) and the A collector hint.  Development tool for the
// Conv4d kernels' tiling decisions; prints cycles per MMA.  Data is all zeros, results are unused.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>
#include "../ncnet_b200/csrc/tc_common.cuh"

using namespace ncnet::tc;

struct RateArgs {
    uint32_t idesc;
    uint64_t adesc, bdesc;        // start address field filled in by the kernel
    uint32_t a_step, b_step;      // descriptor start-address advance (16 B units) per MMA, cycled over `cyc` MMAs
    int cyc, iters, nacc, N, coll, a_tmem, tmem_cols;
    long long *out;
};

template <int COLL>
__device__ __forceinline__ void mma_desc(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc) {
    if (COLL == 0)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                     ::"r"(d), "l"(a), "l"(b), "r"(idesc) : "memory");
    else if (COLL == 1)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::1.kind::f16.collector::a::fill [%0], %1, %2, %3, p;\n\t}"
                     ::"r"(d), "l"(a), "l"(b), "r"(idesc) : "memory");
    else
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::1.kind::f16.collector::a::use [%0], %1, %2, %3, p;\n\t}"
                     ::"r"(d), "l"(a), "l"(b), "r"(idesc) : "memory");
}
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a_tmem, uint64_t b, uint32_t idesc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                 ::"r"(d), "r"(a_tmem), "l"(b), "r"(idesc) : "memory");
}

// A fetched by the first MMA of each group of GRP (collector::a::fill) and reused by the others (collector::a::use)
template <int GRP>
__device__ __forceinline__ void run_coll(const RateArgs &p, uint32_t sa4, uint32_t sb4, uint32_t tb) {
    uint64_t bv[2 * GRP];
    uint32_t dv[2 * GRP];
    const uint64_t a = p.adesc | (uint64_t)sa4;
#pragma unroll
    for (int u = 0; u < 2 * GRP; ++u) {
        bv[u] = p.bdesc | (uint64_t)(sb4 + (uint32_t)(u % p.cyc) * p.b_step);
        dv[u] = tb + (uint32_t)((u % p.nacc) * p.N);
    }
    for (int it = 0; it < p.iters; it += 2 * GRP) {
#pragma unroll
        for (int u = 0; u < 2 * GRP; ++u) {
            if (u % GRP == 0) mma_desc<1>(dv[u], a, bv[u], p.idesc);
            else mma_desc<2>(dv[u], a, bv[u], p.idesc);
        }
    }
}

__global__ void __launch_bounds__(128) rate_kernel(const RateArgs p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    for (int e = threadIdx.x; e < 96 * 1024 / 16; e += blockDim.x) reinterpret_cast<uint4 *>(smem)[e] = make_uint4(0, 0, 0, 0);
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(p.tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tb = tmem_base_s;
    if (warp == 1) {
        const uint32_t sa4 = smem_u32(smem) >> 4, sb4 = (smem_u32(smem) + 64 * 1024) >> 4;
        long long best = 1ll << 60;
        for (int rep = 0; rep < 3; ++rep) {
            const long long t0 = clock64();
            if (elect_one()) {
                // the 8 operand sets of one unrolled round are loop-invariant: the loop body is 8 bare MMAs
                uint64_t av[8], bv[8];
                uint32_t dv[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const uint32_t c = (uint32_t)(u % p.cyc);
                    av[u] = p.adesc | (uint64_t)(sa4 + c * p.a_step);
                    bv[u] = p.bdesc | (uint64_t)(sb4 + c * p.b_step);
                    dv[u] = tb + (uint32_t)((u % p.nacc) * p.N);
                }
                const uint32_t at = tb + (uint32_t)(p.tmem_cols - 16);
                if (p.a_tmem) {
                    for (int it = 0; it < p.iters; it += 8) {
#pragma unroll
                        for (int u = 0; u < 8; ++u) mma_ts(dv[u], at, bv[u], p.idesc);
                    }
                } else if (p.coll == 0) {
                    for (int it = 0; it < p.iters; it += 8) {
#pragma unroll
                        for (int u = 0; u < 8; ++u) mma_desc<0>(dv[u], av[u], bv[u], p.idesc);
                    }
                } else if (p.coll == 1) {
                    for (int it = 0; it < p.iters; it += 8) {       // A refilled every 4th MMA, reused by the next three
#pragma unroll
                        for (int u = 0; u < 8; ++u) {
                            if ((u & 3) == 0) mma_desc<1>(dv[u], av[u], bv[u], p.idesc);
                            else mma_desc<2>(dv[u], av[u], bv[u], p.idesc);
                        }
                    }
                } else if (p.coll == 3) run_coll<3>(p, sa4, sb4, tb);
                else if (p.coll == 5) run_coll<5>(p, sa4, sb4, tb);
                else run_coll<2>(p, sa4, sb4, tb);
                umma_commit(&bar);
            }
            __syncwarp();
            mbar_wait(&bar, rep & 1);
            const long long dt = clock64() - t0;
            if (dt < best) best = dt;
        }
        if ((threadIdx.x & 31) == 0) p.out[blockIdx.x] = best;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tb), "r"(p.tmem_cols) : "memory");
}


// ---- CTA pair (cta_group::2): M = 256 across two SMs, each CTA supplies its 128 rows of A and HALF of B ----
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128) rate2_kernel(const RateArgs p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    for (int e = threadIdx.x; e < 96 * 1024 / 16; e += blockDim.x) reinterpret_cast<uint4 *>(smem)[e] = make_uint4(0, 0, 0, 0);
    const int warp = threadIdx.x >> 5;
    uint32_t rank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    if (threadIdx.x == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(p.tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tb = tmem_base_s;
    long long best = 1ll << 60;
    for (int rep = 0; rep < 3; ++rep) {
        const long long t0 = clock64();
        if (warp == 1 && rank == 0) {
            if (elect_one()) {
                const uint32_t sa4 = smem_u32(smem) >> 4, sb4 = (smem_u32(smem) + 64 * 1024) >> 4;
                uint64_t av[8], bv[8];
                uint32_t dv[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const uint32_t c = (uint32_t)(u % p.cyc);
                    av[u] = p.adesc | (uint64_t)(sa4 + c * p.a_step);
                    bv[u] = p.bdesc | (uint64_t)(sb4 + c * p.b_step);
                    dv[u] = tb + (uint32_t)((u % p.nacc) * p.N);
                }
                if (p.coll == 5) {      // groups of 5 MMAs sharing A: fill, then four uses
                    for (int it = 0; it < p.iters; it += 10) {
#pragma unroll
                        for (int u = 0; u < 10; ++u) {
                            const uint32_t dcol = tb + (uint32_t)((u % p.nacc) * p.N);
                            if (u % 5 == 0)
                                asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, 1, 0;\n\ttcgen05.mma.cta_group::2.kind::f16.collector::a::fill [%0], %1, %2, %3, q;\n\t}"
                                             ::"r"(dcol), "l"(av[0]), "l"(bv[u % 8]), "r"(p.idesc) : "memory");
                            else
                                asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, 1, 0;\n\ttcgen05.mma.cta_group::2.kind::f16.collector::a::use [%0], %1, %2, %3, q;\n\t}"
                                             ::"r"(dcol), "l"(av[0]), "l"(bv[u % 8]), "r"(p.idesc) : "memory");
                        }
                    }
                } else
                for (int it = 0; it < p.iters; it += 8) {
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                        asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, 1, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, q;\n\t}"
                                     ::"r"(dv[u]), "l"(av[u]), "l"(bv[u]), "r"(p.idesc) : "memory");
                }
                asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                             ::"r"(smem_u32(&bar)), "h"((uint16_t)3) : "memory");
            }
            __syncwarp();
        }
        if (warp == 1) {
            mbar_wait(&bar, rep & 1);
            const long long dt = clock64() - t0;
            if (dt < best) best = dt;
        }
    }
    if (warp == 1 && (threadIdx.x & 31) == 0) p.out[blockIdx.x] = best;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tb), "r"(p.tmem_cols) : "memory");
}

static uint64_t desc_hi(uint32_t lbo, uint32_t sbo, int layout) {
    uint64_t d = 0;
    d |= (uint64_t)((lbo >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)layout << 61;       // 0 none, 2 128B, 4 64B, 6 32B
    return d;
}

struct Cfg { const char *name; int M, N, layout, a_mn, b_mn; uint32_t lbo_a, sbo_a, lbo_b, sbo_b; uint32_t a_step, b_step; int cyc, coll, a_tmem, ctas_per_sm; };

int main() {
    long long *out;
    cudaMalloc(&out, 4096 * sizeof(long long));
    cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
    std::vector<Cfg> cfgs;
    const int Ns[] = {16, 32, 48, 80, 128, 160, 240, 256};
    for (int N : Ns) cfgs.push_back({"Kmaj none  M128", 128, N, 0, 0, 0, 8192, 128, (uint32_t)N * 16, 128, 29, 0, 5, 0, 0, 1});
    for (int N : Ns) cfgs.push_back({"Kmaj sw128 M128", 128, N, 2, 0, 0, 16, 1024, 16, 1024, 2, 0, 4, 0, 0, 1});
    for (int N : {80, 256}) cfgs.push_back({"Kmaj sw32  M128", 128, N, 6, 0, 0, 16, 256, 16, 256, 16, 0, 4, 0, 0, 1});
    for (int N : {80, 256}) cfgs.push_back({"Kmaj sw64  M128", 128, N, 4, 0, 0, 16, 512, 16, 512, 32, 0, 4, 0, 0, 1});
    for (int N : {80, 160, 240}) cfgs.push_back({"Kmaj none  M128 sameA", 128, N, 0, 0, 0, 8192, 128, (uint32_t)N * 16, 128, 0, 160, 5, 0, 0, 1});
    for (int N : {80, 160, 240}) cfgs.push_back({"Kmaj none  M128 collector", 128, N, 0, 0, 0, 8192, 128, (uint32_t)N * 16, 128, 0, 160, 5, 1, 0, 1});
    for (int N : {16, 32, 48, 80}) cfgs.push_back({"Kmaj none  M128 coll x5", 128, N, 0, 0, 0, 8192, 128, (uint32_t)N * 16, 128, 0, 160, 5, 5, 0, 1});
    for (int N : {16, 48, 96}) cfgs.push_back({"Kmaj none  M128 coll x3", 128, N, 0, 0, 0, 8192, 128, (uint32_t)N * 16, 128, 0, 160, 3, 3, 0, 1});
    for (int N : {80}) cfgs.push_back({"Kmaj none  M128 coll x2", 128, N, 0, 0, 0, 8192, 128, (uint32_t)N * 16, 128, 0, 160, 2, 2, 0, 1});
    for (int N : {80}) cfgs.push_back({"MNmaj sw32 M128 coll x5", 128, N, 6, 1, 1, 32, 256, 29 * 32, 256, 0, 32, 5, 5, 0, 1});
    for (int N : {80}) cfgs.push_back({"MNmaj sw32 M128", 128, N, 6, 1, 1, 32, 256, 29 * 32, 256, 32, 32, 5, 0, 0, 1});
    for (int N : {80, 160, 240, 256}) cfgs.push_back({"A in TMEM  M128", 128, N, 0, 0, 0, 8192, 128, (uint32_t)N * 16, 128, 0, 0, 1, 0, 1, 1});
    for (int N : {80, 240}) cfgs.push_back({"Kmaj none  M128 2cta/sm", 128, N, 0, 0, 0, 8192, 128, (uint32_t)N * 16, 128, 29, 0, 5, 0, 0, 2});
    for (int N : {40, 80, 128, 256}) cfgs.push_back({"Kmaj none  M64", 64, N, 0, 0, 0, 8192, 128, (uint32_t)N * 16, 128, 29, 0, 5, 0, 0, 1});
    for (int N : {40, 80, 256}) cfgs.push_back({"MNmaj none M64", 64, N, 0, 1, 1, 2048, 16, 2048, 464, 1, 0, 5, 0, 0, 1});
    for (int N : {80, 256}) cfgs.push_back({"MNmaj none M128", 128, N, 0, 1, 1, 2048, 128, 4096, 128, 1, 0, 5, 0, 0, 1});
    for (int N : {80, 256}) cfgs.push_back({"MNmaj sw128 M128", 128, N, 2, 1, 1, 2048, 1024, 4096, 1024, 0, 0, 1, 0, 0, 1});
    printf("%-28s %4s %4s %10s %8s\n", "config", "M", "N", "cyc/MMA", "math");
    for (const Cfg &c : cfgs) {
        RateArgs a{};
        a.idesc = (1u << 4) | ((uint32_t)c.a_mn << 15) | ((uint32_t)c.b_mn << 16) | ((uint32_t)(c.N >> 3) << 17) | ((uint32_t)(c.M >> 4) << 24);
        a.adesc = desc_hi(c.lbo_a, c.sbo_a, c.layout);
        a.bdesc = desc_hi(c.lbo_b, c.sbo_b, c.layout);
        a.a_step = c.a_step; a.b_step = c.b_step; a.cyc = c.cyc; a.iters = c.coll > 1 ? 2040 : 2048; a.N = c.N;
        a.tmem_cols = c.ctas_per_sm == 2 ? 256 : 512;
        a.nacc = (a.tmem_cols - 16) / c.N; if (a.nacc > 3) a.nacc = 3; if (a.nacc < 1) a.nacc = 1;
        a.coll = c.coll; a.a_tmem = c.a_tmem; a.out = out;
        const int grid = 148 * c.ctas_per_sm;
        rate_kernel<<<grid, 128, 96 * 1024>>>(a);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%-28s %4d %4d  ERROR %s\n", c.name, c.M, c.N, cudaGetErrorString(e)); return 1; }
        std::vector<long long> h(grid);
        cudaMemcpy(h.data(), out, grid * sizeof(long long), cudaMemcpyDeviceToHost);
        double s = 0; for (long long v : h) s += (double)v;
        printf("%-28s %4d %4d %10.1f %8.1f\n", c.name, c.M, c.N, s / grid / a.iters, c.M * c.N / 256.0);
        fflush(stdout);
    }

    // CTA pairs: N is the TOTAL N of the M256 MMA, each CTA's B descriptor addresses its own N/2 columns
    cudaFuncSetAttribute(rate2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
    for (int per_sm = 1; per_sm <= 3; ++per_sm)
        for (int N : {80, 160, 256}) {
            RateArgs a{};
            if (per_sm == 3) { if (N != 80) continue; a.coll = 5; }
            a.idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
            a.adesc = desc_hi(8192, 128, 0);
            a.bdesc = desc_hi((uint32_t)(N / 2) * 16, 128, 0);
            a.a_step = 29; a.b_step = 0; a.cyc = 5; a.iters = 2040; a.N = N;
            a.tmem_cols = per_sm == 2 ? 256 : 512;
            a.nacc = (a.tmem_cols - 16) / N; if (a.nacc > 3) a.nacc = 3; if (a.nacc < 1) a.nacc = 1;
            a.out = out;
            const int grid = 148 * (per_sm == 2 ? 2 : 1);
            rate2_kernel<<<grid, 128, 96 * 1024>>>(a);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("pair N=%d: ERROR %s\n", N, cudaGetErrorString(e)); return 1; }
            std::vector<long long> h(grid);
            cudaMemcpy(h.data(), out, grid * sizeof(long long), cudaMemcpyDeviceToHost);
            double sum = 0; int cnt = 0;
            for (int i = 0; i < grid; i += 2) { sum += (double)h[i]; ++cnt; }       // leaders
            printf("%-28s %4d %4d %10.1f %8.1f\n", per_sm == 3 ? "cta_group::2 Kmaj coll x5" : per_sm == 2 ? "cta_group::2 Kmaj 2pair/2sm" : "cta_group::2 Kmaj none", 256, N,
                   sum / cnt / a.iters, 128 * N / 256.0);
            fflush(stdout);
        }
    return 0;
}
