// FeatureCorrelation on the tensor cores (reference lib/model.py:106-118): out[b,a,t] = sum_c fa[b,c,a]*fb[b,c,t].
//
// A batched GEMM, M = nA, N = nB, K = C, both operands MN-major (the channel axis is the strided one).  A pack
// kernel converts each feature map to fp16 (optionally fusing the trunk's featureL2Norm, lib/model.py:85-86) and
// lays it out as  [b][C/64][n/8][64 c][8 positions]:  the 8 positions are the 16 bytes of a UMMA core-matrix row,
// 8 consecutive channels are its 8 K rows (LBO = 128 B), position groups are SBO = 1 KB apart, and the
// (128|256 positions x 64 channels) operand tile of a pipeline stage is ONE contiguous run -> one cp.async.bulk per
// operand per stage.  fp32 accumulation in TMEM (128 x 256 tile), epilogue: optional ReLU + singleton-dim L2 norm
// (normalization=True), rows streamed straight to the [B,nA,nB] volume.
//
// Fused relocalisation pooling (lib/model.py:269-272 + 177-191, k_size = 2): the pack kernel orders the positions of
// both feature maps window-major -- packed index = 4 * (2x2 window, row-major over the pooled grid) + (i%2)*2 + (j%2)
// -- so the four source cells of a pooling window are four consecutive accumulator rows (= four neighbouring lanes of
// an epilogue warp) and the four target cells four consecutive accumulator columns.  The epilogue takes the maximum
// of each 4x4 block (4 columns in registers, then a butterfly over the lane quad; ties -> lowest (i,j,k,l) code, the
// reference's scan order) and writes the pooled value and the four window offsets; the 16x larger volume never
// reaches memory.  Values are rounded to the storage type before they are compared, so the result is bit-identical
// to maxpool4d of the stored un-pooled volume.
#include "common.cuh"
#include "tc_common.cuh"

namespace ncnet {

using namespace tc;

constexpr int kCorrThreads = 320;      // warp 0: producer, warp 1: MMA issuer, warps 2-9: epilogue (two per TMEM lane quadrant)
constexpr int kCorrStages = 4;
constexpr int kCorrKB = 64;            // channels per stage
constexpr int kCorrBM = 128, kCorrBN = 256;

// [B, C, n] (T) -> fp16 [B][KB][npad/8][64][8]; positions >= n and channels >= C are zero.
// pool_w > 0: the map is [n / pool_w, pool_w] and packed position p holds the cell of 2x2 window p/4 (row-major over the
// pooled grid) at in-window offset ((p%4)/2, (p%4)%2).
template <typename T>
__global__ void __launch_bounds__(256)
corr_pack_kernel(const T *__restrict__ f, __half *__restrict__ out, int C, int n, int npad, int kblocks, int l2norm,
                 int pool_w) {
    const int b = blockIdx.z;
    const int mg = blockIdx.x * blockDim.x + threadIdx.x;       // group of 8 positions
    if (mg >= npad / 8) return;
    const T *fb = f + (size_t)b * C * n;
    int src[8];                                                 // source position of each packed position, -1 = padding
#pragma unroll
    for (int u = 0; u < 8; ++u) {
        const int a = mg * 8 + u;
        src[u] = a < n ? a : -1;
        if (pool_w > 0 && a < n) {
            const int win = a >> 2, q = a & 3, ww = pool_w >> 1;
            src[u] = ((win / ww) * 2 + (q >> 1)) * pool_w + (win % ww) * 2 + (q & 1);
        }
    }
    float inv[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) inv[u] = 1.f;
    if (l2norm) {
        float ss[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int c = 0; c < C; ++c)
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const float v = src[u] >= 0 ? ldf(fb + (size_t)c * n + src[u]) : 0.f;
                ss[u] += v * v;
            }
#pragma unroll
        for (int u = 0; u < 8; ++u) inv[u] = 1.f / sqrtf(ss[u] + 1e-6f);
    }
    for (int c = blockIdx.y; c < kblocks * kCorrKB; c += gridDim.y) {
        __align__(16) __half v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u)
            v[u] = __float2half_rn((c < C && src[u] >= 0) ? ldf(fb + (size_t)c * n + src[u]) * inv[u] : 0.f);
        const size_t off = ((((size_t)b * kblocks + c / kCorrKB) * (npad / 8) + mg) * kCorrKB + (c % kCorrKB)) * 8;
        *reinterpret_cast<uint4 *>(out + off) = *reinterpret_cast<const uint4 *>(v);
    }
}

struct CorrArgs {
    const __half *pa, *pb;
    void *out;
    int B, nA, nB, npadA, npadB, kblocks;
    int relu_l2norm, out_f16;
    uint32_t idesc;
    // fused 2x2x2x2 max-pool (pool != 0): out = pooled values [nA/4, nB/4]; idx[4] = max_i, max_j, max_k, max_l
    int pool, idx_bytes;
    void *idx[4];
};

template <typename T>
__device__ __forceinline__ float corr_post(float x, int relu_l2norm);
template <>
__device__ __forceinline__ float corr_post<float>(float x, int relu_l2norm) {
    if (relu_l2norm) {
        x = fmaxf(x, 0.f);
        x = x / sqrtf(x * x + 1e-6f);
    }
    return x;
}
template <>
__device__ __forceinline__ float corr_post<__half>(float x, int relu_l2norm) {
    if (relu_l2norm) {
        x = rnd<__half>(fmaxf(rnd<__half>(x), 0.f));
        x = x / rnd<__half>(sqrtf(rnd<__half>(rnd<__half>(x * x) + 1e-6f)));
    }
    return rnd<__half>(x);
}

// epilogue of the fused pooling: this thread's accumulator row r (a window-major source position) and 16 columns
template <typename T>
__device__ __forceinline__ void corr_pool_chunk(const CorrArgs &p, const float (&v)[16], int r, int t0, int lane) {
    const int aq = lane & 3;                       // in-window source offset code (i%2)*2 + (j%2)
    float bv[4];
    int bc[4];
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        float best = corr_post<T>(v[4 * g], p.relu_l2norm);
        int code = 0;
#pragma unroll
        for (int c = 1; c < 4; ++c) {
            const float x = corr_post<T>(v[4 * g + c], p.relu_l2norm);
            if (x > best) { best = x; code = c; }
        }
        code += aq * 4;                            // (i,j,k,l) row-major window scan order
#pragma unroll
        for (int o = 1; o <= 2; o <<= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int oc = __shfl_xor_sync(0xffffffffu, code, o);
            if (ov > best || (ov == best && oc < code)) { best = ov; code = oc; }
        }
        bv[g] = best; bc[g] = code;
    }
    // lane aq of the quad writes pooled column aq of this chunk
    const float mv = aq == 0 ? bv[0] : (aq == 1 ? bv[1] : (aq == 2 ? bv[2] : bv[3]));
    const int mc = aq == 0 ? bc[0] : (aq == 1 ? bc[1] : (aq == 2 ? bc[2] : bc[3]));
    const int prow = r >> 2, pcol = (t0 >> 2) + aq;
    const int pA = p.nA >> 2, pB = p.nB >> 2;
    if (prow >= pA || pcol >= pB) return;
    const size_t o = (size_t)prow * pB + pcol;
    stf((T *)p.out + o, mv);
    const int off[4] = {mc >> 3, (mc >> 2) & 1, (mc >> 1) & 1, mc & 1};
#pragma unroll
    for (int d = 0; d < 4; ++d) {
        if (p.idx_bytes == 1) ((unsigned char *)p.idx[d])[o] = (unsigned char)off[d];
        else ((long long *)p.idx[d])[o] = off[d];
    }
}

// fp16 outputs: the same 4x4 reduction on packed integer keys -- (order-preserving 16 bits of the rounded value) << 4 |
// (15 - window code) -- so one integer max per comparison returns the larger value and, on ties, the lower code.
__device__ __forceinline__ void corr_pool_chunk_half(const CorrArgs &p, const float (&v)[16], int r, int t0, int lane) {
    const int aq = lane & 3;
    uint32_t bk[4];
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        uint32_t best = 0u;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const float x = p.relu_l2norm ? corr_post<__half>(v[4 * g + c], 1) : v[4 * g + c];
            uint32_t h = __half_as_ushort(__float2half_rn(x));
            h = h == 0x8000u ? 0u : h;                                   // -0 compares equal to +0
            const uint32_t o = (h & 0x8000u) ? (~h & 0xffffu) : (h | 0x8000u);
            const uint32_t key = (o << 4) | (uint32_t)(15 - (aq * 4 + c));
            best = max(best, key);
        }
        best = max(best, __shfl_xor_sync(0xffffffffu, best, 1));
        best = max(best, __shfl_xor_sync(0xffffffffu, best, 2));
        bk[g] = best;
    }
    const uint32_t mk = aq == 0 ? bk[0] : (aq == 1 ? bk[1] : (aq == 2 ? bk[2] : bk[3]));
    const int prow = r >> 2, pcol = (t0 >> 2) + aq;
    const int pA = p.nA >> 2, pB = p.nB >> 2;
    if (prow >= pA || pcol >= pB) return;
    const size_t o = (size_t)prow * pB + pcol;
    const uint32_t ob = mk >> 4;
    const unsigned short hb = (unsigned short)((ob & 0x8000u) ? (ob & 0x7fffu) : (~ob & 0xffffu));
    ((__half *)p.out)[o] = __ushort_as_half(hb);
    const int mc = 15 - (int)(mk & 15u);
    const int off[4] = {mc >> 3, (mc >> 2) & 1, (mc >> 1) & 1, mc & 1};
#pragma unroll
    for (int d = 0; d < 4; ++d) {
        if (p.idx_bytes == 1) ((unsigned char *)p.idx[d])[o] = (unsigned char)off[d];
        else ((long long *)p.idx[d])[o] = off[d];
    }
}

__global__ void __launch_bounds__(kCorrThreads, 1)
corr4d_tc_kernel(const CorrArgs p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    // Persistent: CTA c works on tiles c, c + gridDim.x, ... (N tile fastest, then M tile, then batch).  The 512 TMEM
    // columns hold TWO 128 x 256 accumulators: the MMA warp fills one while the eight epilogue warps drain the other,
    // and the operand ring keeps streaming across tile boundaries.
    __shared__ __align__(8) uint64_t full_bar[kCorrStages], empty_bar[kCorrStages], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_base_s;
    constexpr uint32_t kABytes = kCorrBM / 8 * kCorrKB * 16, kBBytes = kCorrBN / 8 * kCorrKB * 16;   // 16 KB, 32 KB
    constexpr uint32_t kStage = kABytes + kBBytes;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_tiles_n = (p.npadB + kCorrBN - 1) / kCorrBN, n_tiles_m = p.npadA / kCorrBM;
    const int total_tiles = p.B * n_tiles_m * n_tiles_n;

    if (warp == 1 && lane == 0) {
        for (int s = 0; s < kCorrStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], 8); }      // 8 epilogue warps
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base0 = tmem_base_s;

    if (warp == 0) {
        int stage = 0; uint32_t phase = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int nt = tile % n_tiles_n, mt = (tile / n_tiles_n) % n_tiles_m, b = tile / (n_tiles_n * n_tiles_m);
        // the last N tile may be narrower than 256 positions (npadB is a multiple of 128)
        const int n_cols = min(kCorrBN, p.npadB - nt * kCorrBN);
        const uint32_t b_bytes = (uint32_t)(n_cols / 8) * kCorrKB * 16;
        for (int kb = 0; kb < p.kblocks; ++kb) {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            if (elect_one()) {
                uint8_t *sa = smem + (size_t)stage * kStage;
                mbar_expect_tx(&full_bar[stage], kABytes + b_bytes);
                const __half *ga = p.pa + ((((size_t)b * p.kblocks + kb) * (p.npadA / 8) + (size_t)mt * (kCorrBM / 8)) * kCorrKB) * 8;
                const __half *gb = p.pb + ((((size_t)b * p.kblocks + kb) * (p.npadB / 8) + (size_t)nt * (kCorrBN / 8)) * kCorrKB) * 8;
                bulk_g2s(sa, ga, kABytes, &full_bar[stage]);
                bulk_g2s(sa + kABytes, gb, b_bytes, &full_bar[stage]);
            }
            __syncwarp();
            if (++stage == kCorrStages) { stage = 0; phase ^= 1; }
        }
        }
    } else if (warp == 1) {
        int stage = 0; uint32_t phase = 0;
        // MN-major, no swizzle: LBO = 128 B between groups of 8 channels, SBO = 1 KB between groups of 8 positions
        const uint64_t hi = ((uint64_t)(((uint32_t)(kCorrKB * 16) >> 4) | (1u << 14))) << 32;
        const uint32_t lbo_field = (128u >> 4) << 16;
        const uint32_t smem0 = smem_u32_pinned(smem);
        int it = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
        const int nt = tile % n_tiles_n;
        const int n_cols = min(kCorrBN, p.npadB - nt * kCorrBN);
        const uint32_t idesc = (p.idesc & ~(0x3fu << 17)) | ((uint32_t)(n_cols >> 3) << 17);
        const int as = it & 1;                               // accumulator stage of this tile
        const uint32_t tmem_base = tmem_base0 + (uint32_t)(as * kCorrBN);
        if (it >= 2) {                                       // the epilogue must have drained this stage's previous tile
            mbar_wait(&acc_empty[as], (uint32_t)((it >> 1) - 1) & 1u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        }
        for (int kb = 0; kb < p.kblocks; ++kb) {
            mbar_wait(&full_bar[stage], phase);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (elect_one()) {
                const uint32_t sa4 = (smem0 + (uint32_t)stage * (uint32_t)kStage) >> 4;
                const uint32_t sb4 = sa4 + (kABytes >> 4);
#pragma unroll
                for (int kk = 0; kk < kCorrKB / 16; ++kk)        // 16 channels = 2 K groups = 256 B further
                    umma_f16(tmem_base, hi | (uint64_t)(lbo_field | (sa4 + kk * 16)), hi | (uint64_t)(lbo_field | (sb4 + kk * 16)),
                             idesc, (kb | kk) ? 1u : 0u);
                umma_commit(&empty_bar[stage]);
            }
            __syncwarp();
            if (++stage == kCorrStages) { stage = 0; phase ^= 1; }
        }
        if (elect_one()) umma_commit(&acc_full[as]);
        __syncwarp();
        }
    } else {
        const int q = warp & 3;                              // TMEM lane quadrant this warp may read
        const int chalf = (warp - 2) >> 2;                   // which half of the tile's columns it takes
        int it = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
        const int nt = tile % n_tiles_n, mt = (tile / n_tiles_n) % n_tiles_m, b = tile / (n_tiles_n * n_tiles_m);
        const int n_cols = min(kCorrBN, p.npadB - nt * kCorrBN);
        const int as = it & 1;
        const uint32_t tmem_base = tmem_base0 + (uint32_t)(as * kCorrBN);
        mbar_wait(&acc_full[as], (uint32_t)(it >> 1) & 1u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int a = mt * kCorrBM + q * 32 + lane;
        const size_t row = ((size_t)b * p.nA + a) * p.nB;
        // rows start 16-byte aligned when nB is a multiple of 8 (fp16) / 4 (fp32) and the base pointer is aligned
        const bool vec_ok = (p.nB % (p.out_f16 ? 8 : 4)) == 0 && (reinterpret_cast<uintptr_t>(p.out) & 15) == 0;
        const int c_end = min(n_cols, (chalf + 1) * (kCorrBN / 2));
        // tcgen05.ld has a few hundred cycles of latency: the next chunk's load is in flight while this one is processed
        const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
        const int c_begin = chalf * (kCorrBN / 2);
        uint32_t cur[16], nxt[16];
        if (c_begin < c_end) tmem_ld16_issue(lane_base + (uint32_t)c_begin, cur);
        for (int c0 = c_begin; c0 < c_end; c0 += 16) {
            float v[16];
            tmem_ld16_wait(cur);
#pragma unroll
            for (int u = 0; u < 16; ++u) v[u] = __uint_as_float(cur[u]);
            if (c0 + 16 < c_end) tmem_ld16_issue(lane_base + (uint32_t)(c0 + 16), nxt);
            // The per-lane work below diverges (rows / columns outside the matrix skip it); tcgen05.ld / wait::ld are
            // .sync.aligned, so the hand-over to the next chunk happens after the lanes have reconverged (end of the body).
            const bool has_next = c0 + 16 < c_end;
            [&] {
            if (p.pool) {        // whole warp stays converged: the quad butterfly needs all four lanes
                const int t0p = nt * kCorrBN + c0;
                if (p.out_f16) corr_pool_chunk_half(p, v, a, t0p, lane);
                else corr_pool_chunk<float>(p, v, a, t0p, lane);
                return;
            }
            if (a >= p.nA) return;
            const int t0 = nt * kCorrBN + c0;
            if (t0 >= p.nB) return;
#pragma unroll
            for (int u = 0; u < 16; ++u) {
                float x = v[u];
                if (p.relu_l2norm) {
                    if (p.out_f16) {
                        x = rnd<__half>(fmaxf(rnd<__half>(x), 0.f));
                        x = x / rnd<__half>(sqrtf(rnd<__half>(rnd<__half>(x * x) + 1e-6f)));
                    } else {
                        x = fmaxf(x, 0.f);
                        x = x / sqrtf(x * x + 1e-6f);
                    }
                }
                v[u] = x;
            }
            if (vec_ok && t0 + 16 <= p.nB) {            // whole 16-value chunk in range and 16-byte aligned
                if (p.out_f16) {
                    __align__(16) __half h[16];
#pragma unroll
                    for (int u = 0; u < 16; ++u) h[u] = __float2half_rn(v[u]);
                    uint4 *dst = reinterpret_cast<uint4 *>((__half *)p.out + row + t0);
                    dst[0] = reinterpret_cast<const uint4 *>(h)[0];
                    dst[1] = reinterpret_cast<const uint4 *>(h)[1];
                } else {
                    float4 *dst = reinterpret_cast<float4 *>((float *)p.out + row + t0);
#pragma unroll
                    for (int w4 = 0; w4 < 4; ++w4) dst[w4] = make_float4(v[4 * w4], v[4 * w4 + 1], v[4 * w4 + 2], v[4 * w4 + 3]);
                }
            } else {
#pragma unroll
                for (int u = 0; u < 16; ++u) {
                    if (t0 + u >= p.nB) break;
                    if (p.out_f16) ((__half *)p.out)[row + t0 + u] = __float2half_rn(v[u]);
                    else ((float *)p.out)[row + t0 + u] = v[u];
                }
            }
            }();
            __syncwarp();
            if (has_next) {
                tmem_ld16_wait(nxt);
#pragma unroll
                for (int u = 0; u < 16; ++u) cur[u] = nxt[u];
            }
        }
        // this warp has read its share of the accumulator: hand the stage back to the MMA warp
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&acc_empty[as])) : "memory");
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base0), "r"(512) : "memory");
}

static size_t corr_align(size_t v) { return (v + 255) & ~(size_t)255; }
static int pad128(int n) { return (n + 127) / 128 * 128; }

size_t corr4d_tc_scratch_bytes(int B, int C, int nA, int nB) {
    const size_t kblocks = (C + kCorrKB - 1) / kCorrKB;
    return corr_align((size_t)B * kblocks * pad128(nA) * kCorrKB * 2) + corr_align((size_t)B * kblocks * pad128(nB) * kCorrKB * 2) + 256;
}

struct CorrPool {
    int wA, wB, idx_bytes;       // widths of the two feature maps (0 = no pooling)
    void *idx[4];
};

template <typename T>
static int corr_tc_run(const void *fa, const void *fb, void *out, void *scratch, int B, int C, int nA, int nB, int relu_l2norm,
                       int l2norm_inputs, cudaStream_t st, CorrPool pool = CorrPool{}) {
    const int kblocks = (C + kCorrKB - 1) / kCorrKB;
    const int npadA = pad128(nA), npadB = pad128(nB);
    NC_CHECK(B <= 65535, "corr4d(tc): batch too large");
    __half *pa = (__half *)scratch;
    __half *pb = (__half *)((uint8_t *)scratch + corr_align((size_t)B * kblocks * npadA * kCorrKB * 2));
    corr_pack_kernel<T><<<dim3((npadA / 8 + 255) / 256, 64, B), 256, 0, st>>>((const T *)fa, pa, C, nA, npadA, kblocks, l2norm_inputs, pool.wA);
    NC_LAUNCH_OK("corr_pack_kernel(A)");
    corr_pack_kernel<T><<<dim3((npadB / 8 + 255) / 256, 64, B), 256, 0, st>>>((const T *)fb, pb, C, nB, npadB, kblocks, l2norm_inputs, pool.wB);
    NC_LAUNCH_OK("corr_pack_kernel(B)");
    CorrArgs a{};
    a.pa = pa; a.pb = pb; a.out = out; a.B = B; a.nA = nA; a.nB = nB; a.npadA = npadA; a.npadB = npadB; a.kblocks = kblocks;
    a.relu_l2norm = relu_l2norm; a.out_f16 = sizeof(T) == 2;
    a.pool = pool.wA > 0; a.idx_bytes = pool.idx_bytes;
    for (int d = 0; d < 4; ++d) a.idx[d] = pool.idx[d];
    // D = f32, A = B = f16, both MN-major, M = 128; N is patched per tile
    a.idesc = (1u << 4) | (1u << 15) | (1u << 16) | ((uint32_t)(kCorrBN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const size_t smem = (size_t)kCorrStages * (kCorrBM + kCorrBN) / 8 * kCorrKB * 16;
    cudaError_t e = cudaFuncSetAttribute(corr4d_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    NC_CHECK(e == cudaSuccess, "corr4d(tc): smem attribute: %s", cudaGetErrorString(e));
    const long long tiles = (long long)((npadB + kCorrBN - 1) / kCorrBN) * (npadA / kCorrBM) * B;
    NC_CHECK(tiles < (1ll << 31), "corr4d(tc): too many tiles");
    const unsigned grid = (unsigned)(tiles < kNumSMs ? tiles : kNumSMs);      // one persistent CTA per SM
    corr4d_tc_kernel<<<grid, kCorrThreads, smem, st>>>(a);
    NC_LAUNCH_OK("corr4d_tc_kernel");
    return 0;
}

int corr4d_tc_fwd(const void *fa, const void *fb, void *out, void *scratch, int B, int C, int nA, int nB, int dtype,
                  int relu_l2norm, int l2norm_inputs, cudaStream_t st) {
    if (dtype == NCNET_F32) return corr_tc_run<float>(fa, fb, out, scratch, B, C, nA, nB, relu_l2norm, l2norm_inputs, st);
    return corr_tc_run<__half>(fa, fb, out, scratch, B, C, nA, nB, relu_l2norm, l2norm_inputs, st);
}

int corr4d_pool_tc_fwd(const void *fa, const void *fb, void *value, void *const idx[4], void *scratch, int C, int hA, int wA,
                       int hB, int wB, int idx_bytes, int dtype, int relu_l2norm, int l2norm_inputs, cudaStream_t st) {
    CorrPool pool{};
    pool.wA = wA; pool.wB = wB; pool.idx_bytes = idx_bytes;
    for (int d = 0; d < 4; ++d) pool.idx[d] = idx[d];
    if (dtype == NCNET_F32)
        return corr_tc_run<float>(fa, fb, value, scratch, 1, C, hA * wA, hB * wB, relu_l2norm, l2norm_inputs, st, pool);
    return corr_tc_run<__half>(fa, fb, value, scratch, 1, C, hA * wA, hB * wB, relu_l2norm, l2norm_inputs, st, pool);
}

}  // namespace ncnet
