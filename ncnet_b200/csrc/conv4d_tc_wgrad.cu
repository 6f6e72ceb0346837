// Conv4d weight gradient on the tensor cores (mode NCNET_CONV_TC16).
//
//   dW[o,c,t1,t2,t3,t4] = sum_{b,i,j,k,l} G[b,o,i,j,k,l] * X[b,c,i+t1-p,j+t2-p,k+t3-p,l+t4-p]
//
// The contraction runs over POSITIONS, so both operands are "MN-major" for tcgen05: in the
// channel-blocked fp16 layout ([half][position][8 channels], 16 B per position) eight consecutive
// positions are the eight K rows of a core matrix and the 8 channels are its 16 B.  Channels alone
// would give M = N = 16; the MMA's M/N "8-row groups" are addressed with ONE stride (SBO), which is
// exactly what a tap shift is, so the groups of an operand are 8 shifted views of the same staged
// plane (5 of them are taps, 3 are ignored):
//   V2 (Cin,Cout<=16): A = X half hc, groups = t4 shifts (+16 B);  B = G half ho, groups = t3 row
//                      shifts (+P*16 B, stored reversed).  D[hc][ho] (64x64) = dW[o,c,t3,t4] of one
//                      (t1,t2), accumulated in TMEM over every plane of the CTA's slice.
//   V1 (Cin == 1):     A = X' whose 8 "channels" are the t4 taps (im2col by the pack kernel),
//                      groups = t3 row shifts;  B = G, both halves as the two N groups (N = 16).
//   V3 (Cout == 1):    A as V2;  B = G8 whose 8 "channels" are the t2 taps (pack kernel, shifts along
//                      the second volume axis), groups = t3 row shifts: one CTA covers all t2.
// CTA = (group, t1[, t2], slice of the (b,i,j) plane list); 3-stage cp.async.bulk ring, one thread
// issues the MMAs, the accumulators are flushed once at the end as per-slice partial sums which
// unpack_wgrad_kernel reduces deterministically (shared with the fp32 path).
#include "conv4d_tc.cuh"
#include "conv4d_common.cuh"
#include "tc_common.cuh"

namespace ncnet {

constexpr int kWgThreads = 192;
constexpr int kWgStages = 3;

struct WgArgs {
    const __half *xa;   // A side, CB16 [B*GxA][2][d1*d2][d3+2p][P][8]
    const __half *gb;   // B side, CB16 [B*G][2][d1*d2][d3+2p][P][8]
    float *part;        // [S][G][Cin][k][k][k][k][cop]
    int B, G, GxA;
    int d1, d2, d3, d4, KS, P;
    int variant, NAv, NBv, N, nb_copy;      // nb_copy: B halves copied per stage
    uint32_t a_sbo, b_sbo;
    int a_r0, b_r0;                         // buffer row at which the haloed X / G plane is placed
    int npos_a, npos_b, nkb, kb0;           // K blocks [kb0, nkb): blocks that only see halo zeros are skipped
    int S, Cin, Cout, cop;
    uint32_t idesc, tmem_cols;
    uint32_t stage_bytes, a_bytes, b_start;
    int stages;                             // ring depth
};

using namespace tc;

// Work item `pl` of a CTA: which grad plane it stages, which X plane(s), and which of the NAv "A variants" are live.
//   V2: pl = output plane (i,j); X plane (i+t1-p, j+t2-p); A variants = the channel halves of X (all live)
//   V1: pl = output plane (i,j); A variant av = X' plane (i+t1-p, j+av-p): one CTA covers every t2, a variant whose
//       plane falls outside the volume is skipped
//   V3: pl = X plane (i',j'); grad plane (i'-t1+p, j') whose channels hold the t2 taps; A variants = halves of X
struct WgItem { int b; size_t ij_g; int xi, xj; uint32_t amask; };

__device__ __forceinline__ bool wg_item(const WgArgs &p, int pl, int t1, int t2, WgItem &it) {
    const int PAD = p.KS / 2;
    it.b = pl / (p.d1 * p.d2);
    const int ij = pl - it.b * (p.d1 * p.d2);
    const int i = ij / p.d2, j = ij - i * p.d2;
    if (p.variant == 3) {
        const int ig = i - t1 + PAD;
        if (ig < 0 || ig >= p.d1) return false;
        it.xi = i; it.xj = j; it.ij_g = (size_t)ig * p.d2 + j; it.amask = (1u << p.NAv) - 1u;
        return true;
    }
    const int ii = i + t1 - PAD;
    if (ii < 0 || ii >= p.d1) return false;
    it.xi = ii; it.ij_g = (size_t)i * p.d2 + j;
    if (p.variant == 1) {
        it.xj = j - PAD;                      // A variant av reads plane column xj + av
        uint32_t m = 0;
        for (int av = 0; av < p.NAv; ++av) if (it.xj + av >= 0 && it.xj + av < p.d2) m |= 1u << av;
        it.amask = m;
        return m != 0;
    }
    const int jj = j + t2 - PAD;
    if (jj < 0 || jj >= p.d2) return false;
    it.xj = jj; it.amask = (1u << p.NAv) - 1u;
    return true;
}

__global__ void __launch_bounds__(kWgThreads, 1)
conv4d_tc_wgrad_kernel(const WgArgs p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t full_bar[kWgStages], empty_bar[kWgStages], accum_bar;
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int by = blockIdx.y;
    int t2 = 0;
    if (p.variant == 2) { t2 = by % p.KS; by /= p.KS; }
    const int t1 = by % p.KS; by /= p.KS;
    const int g = by;
    const int s = blockIdx.x;
    const int nplanes = p.B * p.d1 * p.d2;
    const int per = (nplanes + p.S - 1) / p.S;
    const int pbeg = s * per, pend = min(nplanes, pbeg + per);

    if (warp == 1 && lane == 0) {
        for (int st = 0; st < p.stages; ++st) { mbar_init(&full_bar[st], 1); mbar_init(&empty_bar[st], 1); }
        mbar_init(&accum_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(p.tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    {   // zero the whole ring once: halo rows/columns and tails are never written by the row copies
        const uint4 z = make_uint4(0, 0, 0, 0);
        uint4 *d = reinterpret_cast<uint4 *>(smem);
        const uint32_t n16 = (uint32_t)p.stages * p.stage_bytes / 16;
        for (uint32_t e = threadIdx.x; e < n16; e += kWgThreads) d[e] = z;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;

    if (warp == 0) {
        // ===== producer =====
        int stage = 0;
        uint32_t phase = 0;
        // a haloed plane is one contiguous run in CB16: one bulk copy per staged channel half
        const size_t plane_cb = (size_t)(p.d3 + 2 * (p.KS / 2)) * p.P;
        const size_t ijn = (size_t)p.d1 * p.d2;
        const uint32_t tile = (uint32_t)(plane_cb * 16);
        for (int pl = pbeg; pl < pend; ++pl) {
            WgItem it;
            if (!wg_item(p, pl, t1, t2, it)) continue;
            mbar_wait(&empty_bar[stage], phase ^ 1);
            if (elect_one()) {
                uint8_t *sa = smem + (size_t)stage * p.stage_bytes;
                uint8_t *sb = sa + p.a_bytes;
                mbar_expect_tx(&full_bar[stage], (uint32_t)(__popc(it.amask) + p.nb_copy) * tile);
                const size_t bga = (size_t)it.b * p.GxA + (p.GxA == 1 ? 0 : g);
                const size_t bgb = (size_t)it.b * p.G + g;
                for (int av = 0; av < p.NAv; ++av) {
                    if (!((it.amask >> av) & 1u)) continue;
                    // V1: variant = plane column, always channel half 0;  V2/V3: variant = channel half of one plane
                    const size_t ijx = (size_t)it.xi * p.d2 + (p.variant == 1 ? it.xj + av : it.xj);
                    const int half = p.variant == 1 ? 0 : av;
                    bulk_g2s(sa + ((size_t)av * p.npos_a + (size_t)p.a_r0 * p.P) * 16,
                             p.xa + (((bga * 2 + half) * ijn + ijx) * plane_cb) * 8, tile, &full_bar[stage]);
                }
                for (int h = 0; h < p.nb_copy; ++h)
                    bulk_g2s(sb + ((size_t)h * p.npos_b + (size_t)p.b_r0 * p.P) * 16,
                             p.gb + (((bgb * 2 + h) * ijn + it.ij_g) * plane_cb) * 8, tile, &full_bar[stage]);
            }
            __syncwarp();
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        int stage = 0;
        uint32_t phase = 0;
        uint32_t inited = 0;                                      // accumulator sets already written once
        const uint32_t lbo_field = (128u >> 4) << 16;                                   // LBO = 128 B (next 8 positions)
        const uint64_t a_hi = ((uint64_t)((p.a_sbo >> 4) | (1u << 14))) << 32;           // SBO + descriptor version 1
        const uint64_t b_hi = ((uint64_t)((p.b_sbo >> 4) | (1u << 14))) << 32;
        const uint32_t n = (uint32_t)p.N, nbv = (uint32_t)p.NBv;
        const uint32_t smem0 = smem_u32_pinned(smem);
        for (int pl = pbeg; pl < pend; ++pl) {
            WgItem it;
            if (!wg_item(p, pl, t1, t2, it)) continue;
            mbar_wait(&full_bar[stage], phase);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            // warp-uniform descriptor arithmetic (only the low word changes); one elected lane issues
            const uint32_t sa4 = (smem0 + (uint32_t)stage * p.stage_bytes) >> 4;
            const uint32_t sb4 = sa4 + ((p.a_bytes + p.b_start) >> 4);
            if (elect_one()) {
                if (p.NAv == 2 && p.NBv == 2) {
                    // 2x2 accumulators: K-block-major in snake order, so consecutive MMAs share one operand descriptor
                    const uint32_t acc0 = (inited & 1u) ? 1u : 0u;
                    uint32_t a0 = lbo_field | (sa4 + (uint32_t)p.kb0 * 16u), a1 = a0 + (uint32_t)p.npos_a;
                    uint32_t b0 = lbo_field | (sb4 + (uint32_t)p.kb0 * 16u), b1 = b0 + (uint32_t)p.npos_b;
                    umma_f16(tmem_base, a_hi | (uint64_t)a0, b_hi | (uint64_t)b0, p.idesc, acc0);
                    umma_f16(tmem_base + n, a_hi | (uint64_t)a0, b_hi | (uint64_t)b1, p.idesc, acc0);
                    umma_f16(tmem_base + 3 * n, a_hi | (uint64_t)a1, b_hi | (uint64_t)b1, p.idesc, acc0);
                    umma_f16(tmem_base + 2 * n, a_hi | (uint64_t)a1, b_hi | (uint64_t)b0, p.idesc, acc0);
#pragma unroll 4
                    for (int kb = p.kb0 + 1; kb < p.nkb; ++kb) {
                        a0 += 16u; a1 += 16u; b0 += 16u; b1 += 16u;
                        umma_f16(tmem_base, a_hi | (uint64_t)a0, b_hi | (uint64_t)b0, p.idesc, 1u);
                        umma_f16(tmem_base + n, a_hi | (uint64_t)a0, b_hi | (uint64_t)b1, p.idesc, 1u);
                        umma_f16(tmem_base + 3 * n, a_hi | (uint64_t)a1, b_hi | (uint64_t)b1, p.idesc, 1u);
                        umma_f16(tmem_base + 2 * n, a_hi | (uint64_t)a1, b_hi | (uint64_t)b0, p.idesc, 1u);
                    }
                } else
                // one accumulator set at a time, its K blocks back to back: the inner loop is two adds and one MMA
                for (int av = 0; av < p.NAv; ++av) {
                    if (!((it.amask >> av) & 1u)) continue;
                    const uint32_t a_lo0 = lbo_field | (sa4 + (uint32_t)av * p.npos_a + (uint32_t)p.kb0 * 16u);
                    const uint32_t acc0 = ((inited >> av) & 1u) ? 1u : 0u;
                    for (int bv = 0; bv < p.NBv; ++bv) {
                        const uint32_t d = tmem_base + ((uint32_t)av * nbv + bv) * n;
                        uint32_t a_lo = a_lo0;
                        uint32_t b_lo = lbo_field | (sb4 + (uint32_t)bv * p.npos_b + (uint32_t)p.kb0 * 16u);
                        umma_f16(d, a_hi | (uint64_t)a_lo, b_hi | (uint64_t)b_lo, p.idesc, acc0);
#pragma unroll 8
                        for (int kb = p.kb0 + 1; kb < p.nkb; ++kb) {
                            a_lo += 16u; b_lo += 16u;                    // next 16 positions (16-byte units)
                            umma_f16(d, a_hi | (uint64_t)a_lo, b_hi | (uint64_t)b_lo, p.idesc, 1u);
                        }
                    }
                }
                umma_commit(&empty_bar[stage]);
            }
            __syncwarp();
            inited |= it.amask;
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
        if (elect_one()) umma_commit(&accum_bar);
        __syncwarp();
    } else {
        // ===== epilogue: flush the accumulators as this slice's partial sums =====
        const int q = warp & 3;
        mbar_wait(&accum_bar, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        uint32_t live = 0;      // accumulator sets the issuer touched (same walk as the other roles)
        for (int pl = pbeg; pl < pend; ++pl) {
            WgItem it;
            if (wg_item(p, pl, t1, t2, it)) live |= it.amask;
            if (live == (1u << p.NAv) - 1u) break;
        }
        if (live) {
            const int KS = p.KS;
            const size_t per_group = (size_t)p.Cin * KS * KS * KS * KS * p.cop;
            float *dst = p.part + ((size_t)s * p.G + g) * per_group;
            const int m = q * 16 + lane;           // M = 64: rows 16q..16q+15 sit in lanes 0..15 of quadrant q
            for (int av = 0; av < p.NAv; ++av) {
                if (!((live >> av) & 1u)) continue;
                for (int bv = 0; bv < p.NBv; ++bv)
                    for (int n0 = 0; n0 < p.N; n0 += 8) {
                        float v[8];
                        tmem_ld8(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((av * p.NBv + bv) * p.N + n0), v);
                        if (lane >= 16) continue;
                        const int sA = m >> 3, c8 = m & 7;
#pragma unroll
                        for (int u = 0; u < 8; ++u) {
                            const int n = n0 + u;
                            int o, c, t3, t4, tt2 = t2;
                            if (p.variant == 2) { t4 = sA; c = av * 8 + c8; t3 = KS - 1 - (n >> 3); o = bv * 8 + (n & 7); }
                            else if (p.variant == 1) { t3 = sA; t4 = c8; c = 0; o = n; tt2 = av; }
                            else { t4 = sA; c = av * 8 + c8; t3 = KS - 1 - (n >> 3); tt2 = n & 7; o = 0; }
                            if (t3 < 0 || t3 >= KS || t4 >= KS || tt2 >= KS || c >= p.Cin || o >= p.Cout) continue;
                            dst[(((((size_t)c * KS + t1) * KS + tt2) * KS + t3) * KS + t4) * p.cop + o] = v[u];
                        }
                    }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
    }
}

// conv4d_fp32.cu: reduces the per-slice partial sums into the reference layout and sums the bias gradient
int wgrad_finish(const float *part, float *gw, float *gb, float *bias_part, const float *out_scale, const void *gy, int S, int G, int B, int Cin,
                 int Cout, int k, int cop, int w_layout, size_t vol, int dtype, cudaStream_t st);

// bias gradient from a CB16 grad_y, stage 1: db[(chunk*nblk + blk)*16 + o] = partial sum over the chunk's positions
// (halo is zero); stage 2 is bias_reduce (conv4d_fp32.cu)
__global__ void __launch_bounds__(256)
cb16_bias_grad_kernel(const __half *__restrict__ gy, float *__restrict__ db, int Cout, size_t npad, int chunks) {
    const size_t blkhalf = blockIdx.y;            // (block, half)
    const int half = (int)(blkhalf & 1);
    const __half *src = gy + blkhalf * npad * 8;
    const size_t per = (npad + chunks - 1) / chunks;
    const size_t beg = (size_t)blockIdx.x * per, end = beg + per < npad ? beg + per : npad;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    size_t e = beg + threadIdx.x;
    for (; e + blockDim.x < end; e += 2 * blockDim.x) {       // two 16-byte loads in flight per thread
        const uint4 r0 = *reinterpret_cast<const uint4 *>(src + e * 8);
        const uint4 r1 = *reinterpret_cast<const uint4 *>(src + (e + blockDim.x) * 8);
        const __half *h0 = reinterpret_cast<const __half *>(&r0), *h1 = reinterpret_cast<const __half *>(&r1);
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[c] += __half2float(h0[c]) + __half2float(h1[c]);
    }
    for (; e < end; e += blockDim.x) {
        const uint4 raw = *reinterpret_cast<const uint4 *>(src + e * 8);
        const __half *h = reinterpret_cast<const __half *>(&raw);
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[c] += __half2float(h[c]);
    }
    __shared__ float sm[8][8];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const float v = warp_sum(acc[c]);
        if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5][c] = v;
    }
    __syncthreads();
    if (threadIdx.x < 8) {
        float t = 0.f;
        for (int w8 = 0; w8 < 8; ++w8) t += sm[w8][threadIdx.x];
        const int o = half * 8 + threadIdx.x;
        db[((size_t)blockIdx.x * (gridDim.y >> 1) + (blkhalf >> 1)) * 16 + o] = t;
    }
}

// conv4d_tc_wgrad_pm.cu: full-size tiles for layers with more than 8 channels on both sides
bool conv4d_tc_wgrad_pm_supported(int d3, int d4, int k);
int conv4d_tc_wgrad_pm_run(const __half *x_cb, const __half *g_cb, const __half *x_pm_in, const __half *g_pm_in, float *part,
                           void *pm_scratch, int B, int G, int GxA, int Cin, int Cout, int d1, int d2, int d3, int d4, int k,
                           int S, int cop, cudaStream_t st);

// conv4d_tc_wgrad_thin.cu: M128 tiles for the single-channel layers (Cin == 1 or Cout == 1)
bool conv4d_tc_wgrad_thin_supported(int d3, int d4, int k);
int conv4d_tc_wgrad_thin_run(const __half *w16_cb, const __half *w16_pm_in, const void *one, int one_dtype, float *part, void *scratch, int cin1,
                             int B, int G, int Gt, int C16, int d1, int d2, int d3, int d4, int k, int S, int cop,
                             cudaStream_t st);

struct WgPlan {
    bool ok;
    int variant, NAv, NBv, N, nb_copy, P, a_r0, b_r0, npos_a, npos_b, nkb, kb0, S, cop, stages;
    uint32_t a_sbo, b_sbo, a_bytes, stage_bytes, tmem_cols, b_start;
    size_t smem, part_bytes;
};

static size_t wg_align(size_t v) { return (v + 255) & ~(size_t)255; }

static WgPlan wg_plan(int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int G) {
    WgPlan w{};
    w.ok = false;
    if (Cin > 16 || Cout > 16 || (k != 3 && k != 5)) return w;
    const int pad = k / 2;
    w.P = d4 + 2 * pad;
    const int plane_cb = (d3 + 2 * pad) * w.P;          // the K range: every position of a haloed plane
    w.nkb = (plane_cb + 15) / 16;
    if (Cin == 1 && Cout > 1) {
        // A = X' (d4 taps in the channels of half 0), groups = d3 row shifts; B = G, the two halves are the N groups
        w.variant = 1; w.NAv = k; w.NBv = 1; w.N = 16; w.nb_copy = 2;       // A variants = the k planes j+t2-p
        w.a_sbo = (uint32_t)w.P * 16u;
        w.a_r0 = pad; w.b_r0 = 0; w.b_start = 0;
        // the three unused row-shift groups read up to 7 rows past a plane's buffer: into the next variant's buffer or
        // the grad buffer behind it (finite data, discarded), so no per-buffer margin is reserved for them
        w.npos_a = w.nkb * 16 + 8;
        if (w.npos_a < (pad + d3 + 2 * pad) * w.P) w.npos_a = (pad + d3 + 2 * pad) * w.P;
        w.npos_b = w.nkb * 16 + 8;
        w.cop = ((Cout + 3) / 4) * 4;
    } else {
        // A = X half, groups = d4 shifts; B = G half (V2) or G8 with the d2 taps in its channels (V3),
        // only the k useful d3 row shifts, in reversed order (group r' <-> tap k-1-r'): N = 8k
        w.variant = Cout == 1 ? 3 : 2;
        w.NAv = (Cin + 7) / 8; w.NBv = w.variant == 3 ? 1 : (Cout + 7) / 8; w.N = 8 * k; w.nb_copy = w.NBv;
        w.a_sbo = 16u;
        w.b_sbo = (uint32_t)w.P * 16u;
        w.a_r0 = 0; w.b_r0 = 8 - pad; w.b_start = (uint32_t)((9 - k) * w.P + pad) * 16u;
        w.npos_a = w.nkb * 16 + 8;
        w.npos_b = w.nkb * 16 + 8 * w.P + pad + 8;
        if (w.npos_b < (w.b_r0 + d3 + 2 * pad) * w.P) w.npos_b = (w.b_r0 + d3 + 2 * pad) * w.P;
        w.cop = w.variant == 3 ? 4 : ((Cout + 3) / 4) * 4;
    }
    w.npos_a = ((w.npos_a + 7) / 8) * 8;
    w.npos_b = ((w.npos_b + 7) / 8) * 8;
    if (w.variant == 1) w.b_sbo = (uint32_t)w.npos_b * 16u;
    if (w.npos_a > 16383 || w.npos_b > 16383 || w.P > 1023) return w;
    w.a_bytes = (uint32_t)w.NAv * w.npos_a * 16u;
    const uint32_t b_bytes = (uint32_t)w.nb_copy * w.npos_b * 16u;
    w.stage_bytes = ((w.a_bytes + b_bytes + 127) / 128) * 128;
    w.stages = kWgStages;
    if ((size_t)w.stages * w.stage_bytes > 220 * 1024) w.stages = 2;
    w.smem = (size_t)w.stages * w.stage_bytes;
    if (w.smem > 220 * 1024) return w;
    const int cols = w.NAv * w.NBv * w.N;
    w.tmem_cols = cols <= 32 ? 32 : (cols <= 64 ? 64 : (cols <= 128 ? 128 : (cols <= 256 ? 256 : 512)));
    // slices: 4-12 waves of CTAs, picking the count whose last wave is fullest (600 CTAs on 148 SMs would idle
    // 140 SMs for a fifth of the run)
    const long long base = (long long)G * k * (w.variant == 2 ? k : 1);
    const long long planes = (long long)B * d1 * d2;
    long long smin = (4ll * kNumSMs + base - 1) / base, smax = (12ll * kNumSMs + base - 1) / base;
    if (smax > planes) smax = planes;
    if (smin > smax) smin = smax;
    if (smin < 1) smin = smax = 1;
    long long S = smin;
    double best = 0.0;
    for (long long c = smin; c <= smax; ++c) {
        const long long ctas = base * c, waves = (ctas + kNumSMs - 1) / kNumSMs;
        const double eff = (double)ctas / (double)(waves * kNumSMs);
        if (eff > best + 0.004) { best = eff; S = c; }
    }
    w.S = (int)S;
    w.part_bytes = wg_align((size_t)w.S * G * Cin * k * k * k * k * w.cop * sizeof(float));
    w.ok = true;
    return w;
}

size_t conv4d_tc_wgrad_scratch_bytes(int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int symmetric) {
    const int G = symmetric ? 2 : 1;
    const WgPlan w = wg_plan(B, Cin, Cout, d1, d2, d3, d4, k, G);
    if (!w.ok) return 0;
    // CB16 copies of planar operands (worst case both) + the per-slice partial sums (+ the position-major copies
    // the full-tile kernel reads)
    const size_t cb = wg_align(cb16_bytes((size_t)B * G, d1, d2, d3, d4, k));
    const bool pm = w.variant == 2 && w.NAv == 2 && w.NBv == 2 && conv4d_tc_wgrad_pm_supported(d3, d4, k);
    const bool thin = w.variant != 2 && conv4d_tc_wgrad_thin_supported(d3, d4, k);
    return (pm ? 4 : (thin ? 5 : 2)) * cb + w.part_bytes + 1024;
}

// returns 2 when the shape/layout is outside the tensor-core kernel's envelope
static int wg_run(const void *x, int x_dtype, const void *gy, int gy_dtype, float *gw, float *gb, void *scratch,
                  float *bias_part, const float *out_scale, int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int w_layout,
                  int symmetric, int shared_input, cudaStream_t st) {
    const int G = symmetric ? 2 : 1;
    const int GxA = (G == 2 && !shared_input) ? 2 : 1;
    const WgPlan w = wg_plan(B, Cin, Cout, d1, d2, d3, d4, k, G);
    if (!w.ok) return 2;
    // NCNET_CB16PM: the position-major twin sits behind the CB16 tensor
    const __half *x_pm = nullptr, *g_pm = nullptr;
    if (x_dtype == NCNET_CB16PM) {
        x_dtype = NCNET_CB16;
        x_pm = (const __half *)((const uint8_t *)x + wg_align(cb16_bytes((size_t)B * GxA, d1, d2, d3, d4, k)));
    }
    if (gy_dtype == NCNET_CB16PM) {
        gy_dtype = NCNET_CB16;
        g_pm = (const __half *)((const uint8_t *)gy + wg_align(cb16_bytes((size_t)B * G, d1, d2, d3, d4, k)));
    }
    if (w.variant == 1 && x_dtype == NCNET_CB16) return 2;      // the tap im2col needs the planar single-channel tensors
    if (w.variant == 3 && gy_dtype == NCNET_CB16) return 2;
    const size_t vol = (size_t)d1 * d2 * d3 * d4;
    uint8_t *sc = (uint8_t *)scratch;
    size_t used = 0;
    const __half *xa, *gbk;
    // single-channel layers: the single-channel tensor is tap-packed by the thin kernel's own pack
    const bool thin = w.variant != 2 && (w.variant == 1 || GxA == G) && conv4d_tc_wgrad_thin_supported(d3, d4, k);
    xa = gbk = nullptr;
    if (thin && w.variant == 1) {
    } else if (x_dtype == NCNET_CB16) xa = (const __half *)x;
    else {
        int rc = cb16_pack(x, x_dtype, sc + used, B * GxA, Cin, d1, d2, d3, d4, k, w.variant == 1 ? 1 : 0, st);
        if (rc) return rc;
        xa = (const __half *)(sc + used);
        used += wg_align(cb16_bytes((size_t)B * GxA, d1, d2, d3, d4, k));
    }
    if (thin && w.variant == 3) {
    } else if (gy_dtype == NCNET_CB16) gbk = (const __half *)gy;
    else {
        int rc = cb16_pack(gy, gy_dtype, sc + used, B * G, Cout, d1, d2, d3, d4, k, w.variant == 3 ? 2 : 0, st);
        if (rc) return rc;
        gbk = (const __half *)(sc + used);
        used += wg_align(cb16_bytes((size_t)B * G, d1, d2, d3, d4, k));
    }
    float *part = (float *)(sc + used);
    cudaError_t e = cudaMemsetAsync(part, 0, w.part_bytes, st);
    NC_CHECK(e == cudaSuccess, "conv4d_wgrad(tc): memset: %s", cudaGetErrorString(e));

    bool launched = false;
    if (thin) {
        const size_t cb = wg_align(cb16_bytes((size_t)B * G, d1, d2, d3, d4, k));
        void *pm_scratch = sc + 2 * cb + w.part_bytes;
        int rc = w.variant == 1
            ? conv4d_tc_wgrad_thin_run(gbk, g_pm, x, x_dtype, part, pm_scratch, 1, B, G, GxA, Cout, d1, d2, d3, d4, k, w.S, w.cop, st)
            : conv4d_tc_wgrad_thin_run(xa, x_pm, gy, gy_dtype, part, pm_scratch, 0, B, G, G, Cin, d1, d2, d3, d4, k, w.S, w.cop, st);
        if (rc) return rc == 2 ? fail("conv4d_wgrad(tc,thin): plan failed after the capability check") : rc;
        launched = true;
    }
    if (w.variant == 2 && w.NAv == 2 && w.NBv == 2 && conv4d_tc_wgrad_pm_supported(d3, d4, k)) {
        // both sides have more than 8 channels: position-major operands, one M128 MMA per 16 positions
        const size_t cb = wg_align(cb16_bytes((size_t)B * G, d1, d2, d3, d4, k));
        void *pm_scratch = sc + 2 * cb + w.part_bytes;
        // the partial sums sit right behind the CB16 copies actually made; the PM16 copies go behind the worst case
        int rc = conv4d_tc_wgrad_pm_run(xa, gbk, x_pm, g_pm, part, pm_scratch, B, G, GxA, Cin, Cout, d1, d2, d3, d4, k, w.S, w.cop, st);
        if (rc == 1) return rc;
        launched = rc == 0;
    }
    WgArgs a{};
    a.xa = xa; a.gb = gbk; a.part = part;
    a.B = B; a.G = G; a.GxA = GxA;
    a.d1 = d1; a.d2 = d2; a.d3 = d3; a.d4 = d4; a.KS = k; a.P = w.P;
    a.variant = w.variant; a.NAv = w.NAv; a.NBv = w.NBv; a.N = w.N; a.nb_copy = w.nb_copy;
    a.a_sbo = w.a_sbo; a.b_sbo = w.b_sbo; a.a_r0 = w.a_r0; a.b_r0 = w.b_r0;
    // K blocks outside [kb0, kb1) multiply nothing but halo zeros (every shifted view of the X side is zero there)
    {
        const int pad = k / 2;
        int lo = pad * w.P + pad - 7;
        if (lo < 0) lo = 0;
        const int hi = (pad + d3 - 1) * w.P + pad + d4;
        a.kb0 = lo / 16;
        a.nkb = (hi + 15) / 16 < w.nkb ? (hi + 15) / 16 : w.nkb;
    }
    a.npos_a = w.npos_a; a.npos_b = w.npos_b; a.S = w.S; a.Cin = Cin; a.Cout = Cout; a.cop = w.cop;
    a.tmem_cols = w.tmem_cols; a.stage_bytes = w.stage_bytes; a.a_bytes = w.a_bytes; a.b_start = w.b_start;
    a.stages = w.stages;
    // D = f32, A = B = f16, both MN-major (bits 15/16), N>>3 at bit 17, M=64 -> 4 at bit 24
    a.idesc = (1u << 4) | (1u << 15) | (1u << 16) | ((uint32_t)(w.N >> 3) << 17) | ((uint32_t)(64 >> 4) << 24);
    e = cudaFuncSetAttribute(conv4d_tc_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)w.smem);
    NC_CHECK(e == cudaSuccess, "conv4d_wgrad(tc): smem attribute: %s", cudaGetErrorString(e));
    if (!launched) {
        dim3 grid(w.S, G * k * (w.variant == 2 ? k : 1));
        conv4d_tc_wgrad_kernel<<<grid, kWgThreads, w.smem, st>>>(a);
        NC_LAUNCH_OK("conv4d_tc_wgrad_kernel");
    }
    if (gy_dtype != NCNET_CB16)
        return wgrad_finish(part, gw, gb, bias_part, out_scale, gy, w.S, G, B, Cin, Cout, k, w.cop, w_layout, vol, gy_dtype, st);
    int rc = wgrad_finish(part, gw, nullptr, nullptr, out_scale, nullptr, w.S, G, B, Cin, Cout, k, w.cop, w_layout, vol, NCNET_F32, st);
    if (rc || !gb) return rc;
    const int pad = k / 2;
    const size_t npad = (size_t)d1 * d2 * (d3 + 2 * pad) * (d4 + 2 * pad);
    int chunks = (int)((kBiasTargetCtas + (size_t)B * G * 2 - 1) / ((size_t)B * G * 2));     // fill the GPU at small batches
    const int max_chunks = (int)((npad + 2047) / 2048);
    if (chunks > max_chunks) chunks = max_chunks;
    if (chunks < 1) chunks = 1;
    NC_CHECK(B * G * 2 <= 65535, "conv4d_wgrad(tc): too many blocks for the bias grid");
    cb16_bias_grad_kernel<<<dim3(chunks, B * G * 2), 256, 0, st>>>(gbk, bias_part, Cout, npad, chunks);
    NC_LAUNCH_OK("cb16_bias_grad_kernel");
    return bias_reduce(bias_part, gb, Cout, chunks * B * G, 16, out_scale, st);
}

int conv4d_tc_wgrad(const void *x, int x_dtype, const void *gy, int gy_dtype, float *gw, float *gb, void *scratch,
                    float *bias_part, const float *out_scale, int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int w_layout,
                    int symmetric, int shared_input, cudaStream_t st) {
    return wg_run(x, x_dtype, gy, gy_dtype, gw, gb, scratch, bias_part, out_scale, B, Cin, Cout, d1, d2, d3, d4, k, w_layout, symmetric,
                  shared_input, st);
}

}  // namespace ncnet
