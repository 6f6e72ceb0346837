// Development check: tcgen05.mma operand addressing with SWIZZLE_32B when the descriptor start address is a
// "tap shift" (multiple of 32 B, not aligned to the 256-byte swizzle period).  Activations are position-major,
// [pos][16 ch] fp16 = 32 B per position, stored with the 16-byte halves XOR-swapped by ADDRESS bit 7.
//   test 0 (MN-major, weight-gradient form): A[m=(atom,ch)][k] = X[s0+atom+k][ch], B[n=(atom,o)][k] = G[r0+atom*P+k][o]
//   test 1 (K-major, forward form):          A[m][k=ch] = X[s0+m][ch],            B[n][k=c] = W[n][c]
// Prints the max error against the CPU for base_offset = 0 and base_offset = (start >> 7) & 1.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cmath>
#include <vector>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include "../ncnet_b200/csrc/tc_common.cuh"
using namespace ncnet::tc;

constexpr int NPOS = 1024;      // positions per operand buffer (32 KB each)
struct Args { int test, s0, r0, P, use_bo, M, N; const __half *x, *g; float *d; };

__global__ void __launch_bounds__(128) check_kernel(const Args p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    // verbatim copy: the global images are already swizzled by (offset bit 7), the buffers are 1024-aligned
    for (int e = threadIdx.x; e < 2 * NPOS * 2; e += blockDim.x) {
        const uint4 *src = reinterpret_cast<const uint4 *>(e < NPOS * 2 ? p.x : p.g) + (e % (NPOS * 2));
        reinterpret_cast<uint4 *>(smem)[e] = *src;
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(128) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tb = tmem_base_s;
    if (warp == 1) {
        if (elect_one()) {
            const uint32_t xa = smem_u32(smem) + (uint32_t)p.s0 * 32u, ga = smem_u32(smem) + NPOS * 32u + (uint32_t)p.r0 * 32u;
            uint64_t ad, bd;
            uint32_t idesc;
            auto mk = [&](uint32_t addr, uint32_t lbo, uint32_t sbo) {
                uint64_t d = 0;
                d |= (uint64_t)((addr >> 4) & 0x3fff);
                d |= (uint64_t)((lbo >> 4) & 0x3fff) << 16;
                d |= (uint64_t)((sbo >> 4) & 0x3fff) << 32;
                d |= (uint64_t)1 << 46;
                if (p.use_bo) d |= (uint64_t)((addr >> 7) & 1) << 49;
                d |= (uint64_t)6 << 61;      // SWIZZLE_32B
                return d;
            };
            if (p.test == 0) {      // MN-major: LBO = stride between 16-element MN atoms, SBO = stride between 8-row K atoms
                ad = mk(xa, 32, 256);
                bd = mk(ga, (uint32_t)p.P * 32u, 256);
                idesc = (1u << 4) | (1u << 15) | (1u << 16) | ((uint32_t)(p.N >> 3) << 17) | ((uint32_t)(p.M >> 4) << 24);
            } else {                // K-major: SBO = stride between 8-row groups (256 B), LBO unused
                ad = mk(xa, 16, 256);
                bd = mk(ga, 16, 256);
                idesc = (1u << 4) | ((uint32_t)(p.N >> 3) << 17) | ((uint32_t)(p.M >> 4) << 24);
            }
            umma_f16(tb, ad, bd, idesc, 0u);
            umma_commit(&bar);
        }
        __syncwarp();
    }
    mbar_wait(&bar, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    for (int c0 = 0; c0 < p.N; c0 += 16) {
        float v[16];
        tmem_ld16(tb + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
        for (int u = 0; u < 16; ++u) p.d[(size_t)(warp * 32 + lane) * p.N + c0 + u] = v[u];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tb), "r"(128) : "memory");
}

int main() {
    std::vector<float> X(NPOS * 16), G(NPOS * 16);
    srand(7);
    for (auto &v : X) v = (float)(rand() % 17 - 8);
    for (auto &v : G) v = (float)(rand() % 13 - 6);
    auto image = [&](const std::vector<float> &src) {
        std::vector<__half> img(NPOS * 16);
        for (int pos = 0; pos < NPOS; ++pos)
            for (int ch = 0; ch < 16; ++ch) {
                const int half = (ch >> 3) ^ ((pos >> 2) & 1);          // byte-offset bit 4 ^= bit 7
                img[(size_t)pos * 16 + half * 8 + (ch & 7)] = __float2half(src[(size_t)pos * 16 + ch]);
            }
        return img;
    };
    auto xi = image(X), gi = image(G);
    __half *dx, *dg; float *dd;
    cudaMalloc(&dx, NPOS * 32); cudaMalloc(&dg, NPOS * 32); cudaMalloc(&dd, 128 * 128 * sizeof(float));
    cudaMemcpy(dx, xi.data(), NPOS * 32, cudaMemcpyHostToDevice);
    cudaMemcpy(dg, gi.data(), NPOS * 32, cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(check_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * NPOS * 32);
    const int P = 29, M = 128, N = 80;
    for (int test = 0; test < 2; ++test)
        for (int use_bo = 0; use_bo < 2; ++use_bo)
            for (int s0 : {0, 1, 3, 4, 5, 8, 11}) {
                const int r0 = (s0 * 3) % 7;
                Args a{test, s0, r0, P, use_bo, M, N, dx, dg, dd};
                cudaMemset(dd, 0, 128 * 128 * sizeof(float));
                check_kernel<<<1, 128, 2 * NPOS * 32>>>(a);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("test %d: CUDA error %s\n", test, cudaGetErrorString(e)); return 1; }
                std::vector<float> D(M * N);
                cudaMemcpy(D.data(), dd, M * N * sizeof(float), cudaMemcpyDeviceToHost);
                double maxerr = 0; int bad = 0;
                for (int m = 0; m < M; ++m)
                    for (int n = 0; n < N; ++n) {
                        double ref = 0;
                        for (int k = 0; k < 16; ++k) {
                            if (test == 0) ref += X[(size_t)(s0 + m / 16 + k) * 16 + m % 16] * G[(size_t)(r0 + (n / 16) * P + k) * 16 + n % 16];
                            else ref += X[(size_t)(s0 + m) * 16 + k] * G[(size_t)(r0 + n) * 16 + k];
                        }
                        const double err = fabs(ref - D[(size_t)m * N + n]);
                        if (err > maxerr) maxerr = err;
                        if (err > 1e-3) ++bad;
                    }
                printf("test %d (%s) base_offset=%d s0=%2d r0=%d : max err %.1f, bad %d / %d\n", test, test ? "K-major" : "MN-major", use_bo, s0, r0, maxerr, bad, M * N);
            }
    return 0;
}
