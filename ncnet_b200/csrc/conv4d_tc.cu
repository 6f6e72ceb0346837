// Conv4d on the 5th-gen tensor cores (mode NCNET_CONV_TC16): fp16 operands, fp32 accumulation in
// TMEM, one implicit-GEMM kernel per layer (reference semantics: lib/conv4d.py:11-51).
//
// GEMM view.  For one (batch, group, iA, jA) the CTA owns a chunk of the (iB,jB) plane, flattened
// with the padded pitch P = d4 + k - 1 so that a tap shift is an affine row offset:
//     A (M x K)  = activation rows: M = 128 consecutive flattened positions, K = 16 channels,
//                  K-major, no swizzle: smem [channel-half][position][8 x fp16] -> a tap is just a
//                  different descriptor start address (16 B per position);
//     B (N x K)  = weights of one (d1,d2[,d3]) tap for N = (tap-in-N, out-channel) columns;
//     D (M x N)  = fp32 accumulators in TMEM, NT tiles of N columns, live for the whole CTA.
// The taps that are folded into N are summed in the epilogue with the matching row shift
// (out[m] = sum_g D[m + shift_g][g]), which keeps N large enough (80 for a 16->16, k=5 layer)
// that one 4 KB A tile is amortised over 5 taps -- with N=16 the MMA would be bound by the
// 128 B/clk shared-memory read rate at 22 % of the tensor peak.
//   mode M1 (Cin<=16, Cout>1): channels = input channels, N = (d4, o), K-steps per plane = d3 taps
//   mode M2 (Cout == 1)      : channels = input channels, N = (d3, d4), one K-step per plane
//   mode M3 (Cin == 1)       : channels = the k taps along d4 (im2col in the pack kernel),
//                              N = (d3, o), one K-step per plane
// Pipeline: warp 0 streams (d1,d2) planes + their weight slice with cp.async.bulk (TMA engine,
// mbarrier complete_tx) through a 3-stage ring; one thread of warp 1 issues tcgen05.mma and
// commits to the stage's "empty" barrier; warps 2-5 drain TMEM (tcgen05.ld), apply the shifted
// sums, bias, ReLU / ReLU-mask, and store.
#include "conv4d_tc.cuh"
#include "conv4d_common.cuh"
#include "tc_common.cuh"

namespace ncnet {

constexpr int kTcThreads = 192;
constexpr int kStages = 3;
constexpr int kMaxGroups = 32;      // column groups (taps folded into N)
constexpr int kMaxKSteps = 8;
constexpr int kOutStride = 20;      // floats per outbuf row (16 channels + pad: conflict-free float4 RMW)

struct TcArgs {
    const __half *act;      // [B*Gx][2][d1*d2][d3][d4][8] fp16 (channel-blocked)
    const __half *wp;       // [G][k*k][KSTEPS][2][N][8] fp16
    const float *bias;      // [Cout] or null
    const void *mask;       // planar tensor shaped like the output, or null
    void *out;              // planar [B, G*Cout, vol] (out_f16 ? half : float)
    int B, G, Gx, Cout, CoutPad;
    int d1, d2, d3, d4, KS, P;
    int lpad;               // column offset of l=0 inside a smem row (pad, or 0 for M3)
    int R, nchunks;         // output k-rows per CTA, chunks per plane
    int NT, N, NG, KSTEPS;
    int npos_pad;           // positions per channel-half in a smem stage
    int relu, out_f16, mask_f16;
    uint32_t idesc;
    int aoff[kMaxKSteps];   // A row offset (positions) per K-step
    int gshift[kMaxGroups]; // epilogue row shift per column group
    uint32_t stage_bytes, data_bytes, w_bytes;
    long long *dbg;         // development only: per-CTA phase timestamps (8 per CTA) or null
};

using namespace tc;

__global__ void __launch_bounds__(kTcThreads, 1)
conv4d_tc_kernel(const TcArgs p, const __grid_constant__ CUtensorMap amap) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t full_bar[kStages], empty_bar[kStages], accum_bar;
    __shared__ uint32_t tmem_base_s;
    __shared__ int s_gshift[kMaxGroups];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int bx = blockIdx.x;
    const int chunk = bx % p.nchunks; bx /= p.nchunks;
    const int j = bx % p.d2; bx /= p.d2;
    const int i = bx % p.d1; bx /= p.d1;
    const int g = bx % p.G;
    const int b = bx / p.G;
    const int PAD = p.KS / 2;
    const int k0 = chunk * p.R;
    const int rows_out = min(p.R, p.d3 - k0);

    uint8_t *stage0 = smem;
    float *outbuf = reinterpret_cast<float *>(smem + (size_t)kStages * p.stage_bytes);
    const int out_rows = p.NT * 128;

    // ---- one-time setup -------------------------------------------------------------------
    if (threadIdx.x < kMaxGroups) s_gshift[threadIdx.x] = p.gshift[threadIdx.x];
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(&accum_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {    // TMEM: all 512 columns (one CTA per SM: the ring + outbuf fill shared memory)
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // zero the activation part of every stage once: halo columns/rows and the tail padding are
    // never written by the row copies, so they stay zero (= the conv's zero padding)
    {
        const uint4 z = make_uint4(0, 0, 0, 0);
        for (int s = 0; s < kStages; ++s) {
            uint4 *d = reinterpret_cast<uint4 *>(stage0 + (size_t)s * p.stage_bytes);
            for (uint32_t e = threadIdx.x; e < p.data_bytes / 16; e += kTcThreads) d[e] = z;
        }
        // outbuf starts at the bias
        if (p.CoutPad >= 16) {
            for (int e = threadIdx.x; e < out_rows * kOutStride; e += kTcThreads) {
                const int o = e % kOutStride;
                outbuf[e] = (p.bias && o < p.Cout) ? __ldg(p.bias + o) : 0.f;
            }
        } else {
            const float b0 = p.bias ? __ldg(p.bias) : 0.f;
            for (int e = threadIdx.x; e < 4 * out_rows; e += kTcThreads) outbuf[e] = e < out_rows ? b0 : 0.f;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy zeros -> visible to TMA/UMMA
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;
    const long long t_start = clock64();
#define TC_STAMP(slot) do { if (p.dbg && blockIdx.x < 4096) p.dbg[blockIdx.x * 8 + (slot)] = clock64() - t_start; } while (0)

    // list of valid (d1,d2) planes is implicit: both roles walk the same loop
    const size_t plane_pos = (size_t)p.d3 * p.d4;
    const size_t ij_count = (size_t)p.d1 * p.d2;

    if (warp == 0) {
        // ===== producer =====
        int stage = 0;
        uint32_t phase = 0;
        const int box_rows = p.R + p.KS - 1;
        const uint32_t half_bytes = (uint32_t)p.P * box_rows * 16u;       // one TMA tile: the halo'd plane chunk of one channel half
        const uint32_t tx_bytes = 2u * half_bytes + p.w_bytes;
        const int bg = b * p.Gx + (p.Gx == 1 ? 0 : g);
        if (elect_one()) prefetch_tmap(&amap);
        for (int t1 = 0; t1 < p.KS; ++t1) {
            const int ii = i + t1 - PAD;
            if (ii < 0 || ii >= p.d1) continue;
            for (int t2 = 0; t2 < p.KS; ++t2) {
                const int jj = j + t2 - PAD;
                if (jj < 0 || jj >= p.d2) continue;
                mbar_wait(&empty_bar[stage], phase ^ 1);
                if (elect_one()) {
                    uint8_t *sb = stage0 + (size_t)stage * p.stage_bytes;
                    mbar_expect_tx(&full_bar[stage], tx_bytes);
                    // rows/columns outside the volume arrive as zeros: that is the convolution's zero padding
                    tma_tile_5d(sb, &amap, 0, -p.lpad, k0 - PAD, ii * p.d2 + jj, bg * 2 + 0, &full_bar[stage]);
                    tma_tile_5d(sb + (size_t)p.npos_pad * 16, &amap, 0, -p.lpad, k0 - PAD, ii * p.d2 + jj, bg * 2 + 1,
                                &full_bar[stage]);
                    const __half *wsrc = p.wp + ((size_t)g * p.KS * p.KS + (size_t)t1 * p.KS + t2) * (p.w_bytes / 2);
                    bulk_g2s(sb + p.data_bytes, wsrc, p.w_bytes, &full_bar[stage]);
                }
                __syncwarp();
                if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (one thread) =====
        int stage = 0;
        uint32_t phase = 0;
        bool first = true;
        const uint32_t lbo_a = (uint32_t)p.npos_pad * 16u;
        const uint32_t lbo_b = (uint32_t)p.N * 16u;
        const uint32_t lbo_a_field = (lbo_a >> 4) << 16, lbo_b_field = (lbo_b >> 4) << 16;
        const uint64_t desc_hi = ((uint64_t)((128u >> 4) | (1u << 14))) << 32;      // SBO = 128 B, descriptor version 1
        for (int t1 = 0; t1 < p.KS; ++t1) {
            const int ii = i + t1 - PAD;
            if (ii < 0 || ii >= p.d1) continue;
            for (int t2 = 0; t2 < p.KS; ++t2) {
                const int jj = j + t2 - PAD;
                if (jj < 0 || jj >= p.d2) continue;
                mbar_wait(&full_bar[stage], phase);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                if (lane == 0 && first) TC_STAMP(0);
                // Descriptor arithmetic stays warp-uniform (no lane dependence) and only the low 32 bits
                // change, so ptxas keeps it on the uniform datapath; lane 0 alone executes the MMAs.
                const uint32_t sa4 = smem_u32(stage0 + (size_t)stage * p.stage_bytes) >> 4;
                const uint32_t sw4 = sa4 + (p.data_bytes >> 4);
                if (elect_one()) {
#pragma unroll
                for (int ks = 0; ks < kMaxKSteps; ++ks) {
                    if (ks < p.KSTEPS) {
                        const uint64_t bdesc = desc_hi | (uint64_t)(lbo_b_field | (sw4 + (uint32_t)ks * 2u * (lbo_b >> 4)));
                        const uint32_t a_lo0 = lbo_a_field | (sa4 + (uint32_t)p.aoff[ks]);
                        const uint32_t acc = (first && ks == 0) ? 0u : 1u;
                        for (int t0 = 0; t0 < p.NT; t0 += 8) {
#pragma unroll
                            for (int u = 0; u < 8; ++u) {
                                const int t = t0 + u;
                                if (t < p.NT)
                                    umma_f16(tmem_base + (uint32_t)(t * p.N), desc_hi | (uint64_t)(a_lo0 + (uint32_t)t * 128u),
                                             bdesc, p.idesc, acc);
                            }
                        }
                    }
                }
                umma_commit(&empty_bar[stage]);     // frees the smem slot when these MMAs have read it
                }
                __syncwarp();
                first = false;
                if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
        }
        if (elect_one()) { TC_STAMP(1); umma_commit(&accum_bar); }
        __syncwarp();
    } else {
        // ===== epilogue (warps 2..5 <-> TMEM lane quadrants warp%4) =====
        const int q = warp & 3;
        const int et = threadIdx.x - 64;                 // 0..127
        mbar_wait(&accum_bar, 0);
        if (et == 0) TC_STAMP(2);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (p.CoutPad >= 16) {
            // column group gi = 16 out-channels of one folded tap: out[m - shift_gi][o] += D[m][gi*16 + o].
            // Within a group every thread owns a distinct row; groups are separated by a barrier.
            for (int t = 0; t < p.NT; ++t) {
                const int m = t * 128 + q * 32 + lane;
                for (int gi = 0; gi < p.NG; ++gi) {
                    float v[16];
                    tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(t * p.N + gi * 16), v);
                    const int dest = m - s_gshift[gi];
                    if (dest >= 0) {
                        float4 *dst = reinterpret_cast<float4 *>(outbuf + (size_t)dest * kOutStride);
#pragma unroll
                        for (int w4 = 0; w4 < 4; ++w4) {
                            float4 a = dst[w4];
                            a.x += v[4 * w4 + 0]; a.y += v[4 * w4 + 1]; a.z += v[4 * w4 + 2]; a.w += v[4 * w4 + 3];
                            dst[w4] = a;
                        }
                    }
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                }
            }
        } else {
            // single out-channel: one column per folded (d3,d4) tap.  Each warp accumulates into its
            // own copy of the plane (stride 1), so no cross-warp ordering is needed; the store sums them.
            float *mine = outbuf + (size_t)q * out_rows;
            for (int t = 0; t < p.NT; ++t) {
                const int m = t * 128 + q * 32 + lane;
                for (int c0 = 0; c0 < p.NG; c0 += 16) {
                    float v[16];
                    tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(t * p.N + c0), v);
#pragma unroll
                    for (int u = 0; u < 16; ++u) {
                        if (c0 + u < p.NG) {
                            const int dest = m - s_gshift[c0 + u];
                            if (dest >= 0) mine[dest] += v[u];
                        }
                        __syncwarp();
                    }
                }
            }
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (et == 0) TC_STAMP(3);
        // ---- store: planar output, coalesced along l -----------------------------------------
        const size_t vol = ij_count * plane_pos;
        const int ew = et >> 5;                                   // epilogue warp 0..3
        const int npairs = p.Cout * rows_out;                     // (channel, output row) pairs, one warp each
        for (int pr = ew; pr < npairs; pr += 4) {
            const int o = pr / rows_out, rr = pr - o * rows_out;
            const size_t off0 = ((size_t)b * p.G * p.Cout + (size_t)g * p.Cout + o) * vol +
                                ((size_t)i * p.d2 + j) * plane_pos + (size_t)(k0 + rr) * p.d4;
            for (int l = lane; l < p.d4; l += 32) {
                const int row = rr * p.P + l;
                float v;
                if (p.CoutPad >= 16) v = outbuf[(size_t)row * kOutStride + o];
                else v = outbuf[row] + outbuf[out_rows + row] + outbuf[2 * out_rows + row] + outbuf[3 * out_rows + row];
                if (p.relu) v = fmaxf(v, 0.f);
                const size_t off = off0 + l;
                if (p.mask) {
                    const float mv = p.mask_f16 ? __half2float(((const __half *)p.mask)[off]) : ((const float *)p.mask)[off];
                    if (!(mv > 0.f)) v = 0.f;
                }
                if (p.out_f16) ((__half *)p.out)[off] = __float2half_rn(v);
                else ((float *)p.out)[off] = v;
            }
        }
    }

    if (threadIdx.x == 64) TC_STAMP(4);
    // ---- teardown ------------------------------------------------------------------------------
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
    }
}

// ---------------------------------------------------------------------------------------------
// operand packing
// ---------------------------------------------------------------------------------------------
// planar [B, Gx*Cin, vol] (T) -> channel-blocked fp16 [B*Gx][2][vol][8].
// taps_as_channels (mode M3, Cin == 1): channel c' of position (.., l) is x[.., l + c' - pad].
template <typename T>
__global__ void __launch_bounds__(256)
pack_act_kernel(const T *__restrict__ x, __half *__restrict__ act, int Cin, size_t vol, int d4, int taps, int pad) {
    const size_t bg = blockIdx.y;
    const int half = blockIdx.z;
    for (size_t pos = (size_t)blockIdx.x * blockDim.x + threadIdx.x; pos < vol; pos += (size_t)gridDim.x * blockDim.x) {
        __align__(16) __half v[8];
        if (taps == 0) {
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const int ch = half * 8 + c;
                v[c] = ch < Cin ? __float2half_rn(ldf(x + (bg * Cin + ch) * vol + pos)) : __float2half_rn(0.f);
            }
        } else {
            const int l = (int)(pos % d4);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const int tap = half * 8 + c;
                const int ll = l + tap - pad;
                v[c] = (tap < taps && ll >= 0 && ll < d4) ? __float2half_rn(ldf(x + bg * vol + pos + (tap - pad)))
                                                           : __float2half_rn(0.f);
            }
        }
        *reinterpret_cast<uint4 *>(act + ((bg * 2 + half) * vol + pos) * 8) = *reinterpret_cast<const uint4 *>(v);
    }
}

struct TcPackW {
    PackDesc d;
    int mode;       // 1,2,3
    int KSTEPS, N, CoutPad;
};

// -> [G][k*k][KSTEPS][2][N][8] fp16
__global__ void pack_w_tc_kernel(const float *__restrict__ w, __half *__restrict__ wp, TcPackW a) {
    const int k = a.d.k;
    const long long total = (long long)a.d.G_e * k * k * a.KSTEPS * 2 * a.N * 8;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        long long r = idx;
        const int c8 = (int)(r % 8); r /= 8;
        const int n = (int)(r % a.N); r /= a.N;
        const int half = (int)(r % 2); r /= 2;
        const int ks = (int)(r % a.KSTEPS); r /= a.KSTEPS;
        const int t2 = (int)(r % k); r /= k;
        const int t1 = (int)(r % k); r /= k;
        const int g = (int)r;
        const int kch = half * 8 + c8;
        int oe, ci, t3, t4;
        bool ok = true;
        if (a.mode == 1) {          // N = (t4, o), K-step = t3, channel = ci
            t3 = ks; t4 = n / a.CoutPad; oe = n % a.CoutPad; ci = kch;
        } else if (a.mode == 2) {   // N = (t3, t4), single out channel, channel = ci
            t3 = n / k; t4 = n % k; oe = 0; ci = kch;
            ok = n < k * k;
        } else {                    // N = (t3, o), channel = t4 tap, single in channel
            t3 = n / a.CoutPad; oe = n % a.CoutPad; t4 = kch; ci = 0;
            ok = kch < k;
        }
        float v = 0.f;
        if (ok && t3 < k && t4 < k && oe < a.d.Cout_e && ci < a.d.Cin_e)
            v = w[packed_source_index(a.d, g, oe, ci, t1, t2, t3, t4)];
        wp[idx] = __float2half_rn(v);
    }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
long long *g_tc_debug_buffer = nullptr;     // set through ncnet_debug_set_tc_timing (development only)

struct TcPlan {
    int mode, KSTEPS, N, NG, CoutPad, P, lpad, R, nchunks, NT, npos_pad;
    uint32_t data_bytes, w_bytes, stage_bytes;
    size_t smem_bytes;
    bool ok;
};

static TcPlan tc_plan(int Cin_e, int Cout_e, int k, int d3, int d4) {
    TcPlan pl{};
    pl.ok = false;
    if (Cin_e > 16 || Cout_e > 16 || (k != 3 && k != 5)) return pl;
    if (Cin_e == 1 && Cout_e > 1) { pl.mode = 3; pl.CoutPad = 16; pl.NG = k; pl.KSTEPS = 1; pl.lpad = 0; }
    else if (Cout_e == 1)          { pl.mode = 2; pl.CoutPad = 1; pl.NG = k * k; pl.KSTEPS = 1; pl.lpad = k / 2; }
    else                           { pl.mode = 1; pl.CoutPad = 16; pl.NG = k; pl.KSTEPS = k; pl.lpad = k / 2; }
    pl.N = pl.mode == 2 ? ((k * k + 15) / 16) * 16 : k * 16;
    pl.P = d4 + k - 1;
    if (pl.P * 16 > 65535 * 16) return pl;
    const int max_aoff = pl.mode == 1 ? (k - 1) * pl.P : 0;
    const int max_shift = pl.mode == 1 ? (k - 1) : (pl.mode == 2 ? (k - 1) * pl.P + (k - 1) : (k - 1) * pl.P);
    pl.w_bytes = (uint32_t)pl.KSTEPS * 2u * pl.N * 16u;
    auto fit = [&](int R, TcPlan &o) {
        const int mrows = (R - 1) * o.P + d4 + max_shift;
        const int NT = (mrows + 127) / 128;
        if (NT * o.N > 512) return false;
        int npos = NT * 128 + max_aoff;
        const int need = (R + k - 1) * o.P;
        if (need > npos) npos = need;
        npos = ((npos + 7) / 8) * 8;
        if (npos > 16383) return false;                      // 14-bit LBO field (16 B units)
        const uint32_t data = 2u * npos * 16u;
        const uint32_t stage = ((data + o.w_bytes + 127) / 128) * 128;
        const size_t outb = o.CoutPad >= 16 ? (size_t)NT * 128 * kOutStride * sizeof(float) : (size_t)NT * 128 * 4 * sizeof(float);
        const size_t smem = (size_t)kStages * stage + outb;
        if (smem > 220 * 1024) return false;
        o.R = R; o.NT = NT; o.npos_pad = npos; o.data_bytes = data; o.stage_bytes = stage; o.smem_bytes = smem;
        return true;
    };
    for (int R = d3; R >= 1; --R) {
        if (!fit(R, pl)) continue;
        pl.nchunks = (d3 + R - 1) / R;
        const int balanced = (d3 + pl.nchunks - 1) / pl.nchunks;    // equal chunks; smaller R always fits too
        if (!fit(balanced, pl)) return pl;
        pl.ok = true;
        break;
    }
    return pl;
}

static size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

size_t conv4d_tc_scratch_bytes(int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int symmetric) {
    const size_t vol = (size_t)d1 * d2 * d3 * d4;
    const int G = symmetric ? 2 : 1;
    (void)Cin; (void)Cout;
    // activations: up to G groups, 16 channels fp16 each; weights: largest N is 5*16=80 -> bound generously
    const size_t act = (size_t)B * G * 2 * vol * 16;
    const size_t wts = (size_t)G * k * k * k * 2 * 96 * 16;
    return align256(act) + align256(wts) + 512;
}

static bool tc_supported(int Cin_e, int Cout_e, int k, int d3, int d4) { return tc_plan(Cin_e, Cout_e, k, d3, d4).ok; }

// runs one effective convolution on the tensor cores; returns 2 when the shape is unsupported
static int run_tc(const void *x, const float *w, const float *bias, const void *mask, void *y, void *scratch,
                  PackDesc d, int B, int Gx, int d1, int d2, int d3, int d4, int relu, int dtype, cudaStream_t st) {
    const TcPlan pl = tc_plan(d.Cin_e, d.Cout_e, d.k, d3, d4);
    if (!pl.ok) return 2;
    const size_t vol = (size_t)d1 * d2 * d3 * d4;
    __half *act = reinterpret_cast<__half *>(scratch);
    __half *wp = reinterpret_cast<__half *>((uint8_t *)scratch + align256((size_t)B * Gx * 2 * vol * 16));

    // 1. activations -> channel-blocked fp16
    {
        unsigned gx = (unsigned)((vol + 255) / 256 < 4096 ? (vol + 255) / 256 : 4096);
        dim3 grid(gx, B * Gx, 2);
        NC_CHECK(B * Gx <= 65535, "conv4d(tc): batch too large");
        const int taps = pl.mode == 3 ? d.k : 0;
        if (dtype == NCNET_F32)
            pack_act_kernel<float><<<grid, 256, 0, st>>>((const float *)x, act, d.Cin_e, vol, d4, taps, d.k / 2);
        else
            pack_act_kernel<__half><<<grid, 256, 0, st>>>((const __half *)x, act, d.Cin_e, vol, d4, taps, d.k / 2);
        NC_LAUNCH_OK("pack_act_kernel");
    }
    // 2. weights
    {
        TcPackW a{};
        a.d = d; a.mode = pl.mode; a.KSTEPS = pl.KSTEPS; a.N = pl.N; a.CoutPad = pl.CoutPad;
        const long long total = (long long)d.G_e * d.k * d.k * pl.KSTEPS * 2 * pl.N * 8;
        const int blocks = (int)((total + 255) / 256 < 1184 ? (total + 255) / 256 : 1184);
        pack_w_tc_kernel<<<blocks, 256, 0, st>>>(w, wp, a);
        NC_LAUNCH_OK("pack_w_tc_kernel");
    }
    // 3. the implicit GEMM
    TcArgs a{};
    a.act = act; a.wp = wp; a.bias = bias; a.mask = mask; a.out = y;
    a.B = B; a.G = d.G_e; a.Gx = Gx; a.Cout = d.Cout_e; a.CoutPad = pl.CoutPad;
    a.d1 = d1; a.d2 = d2; a.d3 = d3; a.d4 = d4; a.KS = d.k; a.P = pl.P; a.lpad = pl.lpad;
    a.R = pl.R; a.nchunks = pl.nchunks; a.NT = pl.NT; a.N = pl.N; a.NG = pl.NG; a.KSTEPS = pl.KSTEPS;
    a.npos_pad = pl.npos_pad; a.relu = relu; a.out_f16 = dtype == NCNET_F16; a.mask_f16 = dtype == NCNET_F16;
    a.stage_bytes = pl.stage_bytes; a.data_bytes = pl.data_bytes; a.w_bytes = pl.w_bytes;
    // instruction descriptor: D=f32 (bits 4-5 = 1), A=B=f16 (0), both K-major, N>>3 at bit 17, M>>4 at bit 24
    a.dbg = g_tc_debug_buffer;
    a.idesc = (1u << 4) | ((uint32_t)(pl.N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    for (int ks = 0; ks < kMaxKSteps; ++ks) a.aoff[ks] = (pl.mode == 1 && ks < d.k) ? ks * pl.P : 0;
    for (int gi = 0; gi < kMaxGroups; ++gi) {
        int s = 0;
        if (gi < pl.NG) {
            if (pl.mode == 1) s = gi;                                   // group = d4 tap
            else if (pl.mode == 2) s = (gi / d.k) * pl.P + (gi % d.k);  // group = (d3, d4)
            else s = gi * pl.P;                                         // group = d3 tap
        }
        a.gshift[gi] = s;
    }
    cudaError_t e = cudaFuncSetAttribute(conv4d_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes);
    NC_CHECK(e == cudaSuccess, "conv4d(tc): smem attribute: %s", cudaGetErrorString(e));
    const long long ctas = (long long)B * d.G_e * d1 * d2 * pl.nchunks;
    NC_CHECK(ctas < (1ll << 31), "conv4d(tc): grid too large");
    CUtensorMap amap;
    if (make_act_tmap(&amap, act, (size_t)B * Gx * 2, d1, d2, d3, d4, pl.P, pl.R + d.k - 1)) return 1;
    conv4d_tc_kernel<<<(unsigned)ctas, kTcThreads, pl.smem_bytes, st>>>(a, amap);
    NC_LAUNCH_OK("conv4d_tc_kernel");
    return 0;
}

int conv4d_tc_fwd(const void *x, const float *w, const float *bias, void *y, void *scratch, int B, int Cin, int Cout,
                  int d1, int d2, int d3, int d4, int k, int w_layout, int symmetric, int shared_input, int relu,
                  int dtype, cudaStream_t st) {
    PackDesc d{};
    d.Cin = Cin; d.Cout = Cout; d.k = k; d.w_layout = w_layout; d.dgrad = 0; d.merge_groups = 0;
    d.G_e = symmetric ? 2 : 1; d.Cin_e = Cin; d.Cout_e = Cout;
    const int Gx = (symmetric && !shared_input) ? 2 : 1;
    return run_tc(x, w, bias, nullptr, y, scratch, d, B, Gx, d1, d2, d3, d4, relu, dtype, st);
}

int conv4d_tc_dgrad(const void *gy, const float *w, const void *mask, void *gx, void *scratch, int B, int Cin,
                    int Cout, int d1, int d2, int d3, int d4, int k, int w_layout, int symmetric, int shared_input,
                    int dtype, cudaStream_t st) {
    PackDesc d{};
    d.Cin = Cin; d.Cout = Cout; d.k = k; d.w_layout = w_layout; d.dgrad = 1;
    const int G = symmetric ? 2 : 1;
    int Gx;
    if (G == 2 && shared_input) { d.merge_groups = 1; d.G_e = 1; d.Cin_e = 2 * Cout; d.Cout_e = Cin; Gx = 1; }
    else { d.merge_groups = 0; d.G_e = G; d.Cin_e = Cout; d.Cout_e = Cin; Gx = G; }
    return run_tc(gy, w, nullptr, mask, gx, scratch, d, B, Gx, d1, d2, d3, d4, 0, dtype, st);
}

bool conv4d_tc_supports(int Cin_e, int Cout_e, int k, int d3, int d4) { return tc_supported(Cin_e, Cout_e, k, d3, d4); }

}  // namespace ncnet

// development hook (not part of include/ncnet_b200.h): device buffer of 4096*8 int64 phase timestamps
extern "C" void ncnet_debug_set_tc_timing(void *device_buffer) { ncnet::g_tc_debug_buffer = (long long *)device_buffer; }
