// Conv4d weight gradient, 16 -> 16 channel layers, on full-size tensor-core tiles (mode NCNET_CONV_TC16).
//
//   dW[o,c,t1,t2,t3,t4] = sum_{b,i,j,k,l} G[b,o,i,j,k,l] * X[b,c,i+t1-p,j+t2-p,k+t3-p,l+t4-p]     (lib/conv4d.py:11-51)
//
// conv4d_tc_wgrad.cu reads the channel-blocked layout ([half][position][8 ch]), where one MMA group is 8 channels, so
// its tiles are M64 x N40 -- four half-rate MMAs (26 cycles each) per 16 positions.  Here both operands are first
// re-laid-out POSITION-MAJOR ("PM16": [plane][position][16 ch] fp16, 32 B per position, the two 16-byte halves swapped
// where bit 7 of the byte offset is set = the SWIZZLE_32B pattern).  With SWIZZLE_32B an MN-major "atom" is 16
// channels x 8 positions, atoms are addressed with one stride (LBO), and a tap shift is exactly such a stride:
//     A (M = 128): atom a = X shifted by a positions (the d4 taps; 5 of 8 used),        LBO = 32 B
//     B (N = 16k): atom r = G shifted by r rows      (the d3 taps, reversed order),     LBO = P * 32 B
//     K = 16 consecutive haloed positions per MMA (two 8-row K atoms, SBO = 256 B)
// so ONE M128 x N80 x K16 MMA (52 cycles) per 16 positions replaces the four M64 x N40 ones (105 cycles).  The
// hardware applies the swizzle to absolute shared-memory address bits, so shifted descriptor start addresses read
// the verbatim-copied image correctly (scripts/umma_sw32_check.cu verifies this on the device).
// CTA = (group, t1, t2, slice of the (b,i,j) plane list); D[(t4,c)][(r,o)] accumulates in TMEM over the slice and is
// flushed as per-slice partial sums in the layout unpack_wgrad_kernel reduces.
#include "conv4d_tc.cuh"
#include "tc_common.cuh"
#include <cstdlib>

namespace ncnet {

using namespace tc;

constexpr int kPmThreads = 192;
constexpr int kPmStages = 4;

// CB16 [blk][2][ij][npos][8] -> PM16 [blk][ij][npos][16] with the SWIZZLE_32B half swap (npos = haloed plane positions)
__global__ void __launch_bounds__(256)
pm16_from_cb16_kernel(const uint4 *__restrict__ cb, uint4 *__restrict__ pm, size_t nblk, size_t ij, int npos) {
    const size_t per_half = ij * (size_t)npos;
    const size_t total = nblk * per_half;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const size_t blk = e / per_half, r = e - blk * per_half;
        const int pos = (int)(r % npos);
        const uint4 h0 = cb[(blk * 2 + 0) * per_half + r];
        const uint4 h1 = cb[(blk * 2 + 1) * per_half + r];
        const bool swap = (pos >> 2) & 1;
        pm[e * 2 + 0] = swap ? h1 : h0;
        pm[e * 2 + 1] = swap ? h0 : h1;
    }
}

// shared with conv4d_tc_wgrad_thin.cu
void pm16_from_cb16(const __half *cb, __half *pm, size_t nblk, size_t ij, int npos, cudaStream_t st) {
    const size_t n = nblk * ij * (size_t)npos;
    const int blocks = (int)((n + 255) / 256 < 148 * 16 ? (n + 255) / 256 : 148 * 16);
    pm16_from_cb16_kernel<<<blocks, 256, 0, st>>>((const uint4 *)cb, (uint4 *)pm, nblk, ij, npos);
}

struct WgPmArgs {
    const __half *x;    // PM16 [B*GxA][d1*d2][npos][16]
    const __half *g;    // PM16 [B*G][d1*d2][npos][16]
    float *part;        // [S][G][Cin][k][k][k][k][cop]
    int B, G, GxA, d1, d2, KS, N;
    int nkb, kb0;                   // K blocks [kb0, nkb) of 16 positions
    int S, Cin, Cout, cop, stages;
    uint32_t plane_bytes;           // bytes of one haloed PM16 plane
    int nav, ncls;                  // d2 taps (X plane variants) per CTA, CTA classes per (group, t1) = ceil(k / nav)
    uint32_t a1_bytes;              // one X plane buffer
    uint32_t a_bytes, stage_bytes;  // all X buffers of a stage / whole stage
    uint32_t g_off;                 // byte offset of the G plane inside the B buffer (multiple of 256)
    uint32_t b_rel;                 // byte offset, from the B buffer, of B's K position 0 (atom 0)
    uint32_t lbo_b;                 // P * 32
    uint32_t idesc, tmem_cols;
};

// Walk over the CTA's output planes (b,i,j) without divisions in the loop (a lone lane runs ~10 cycles per instruction;
// two integer divisions per item cost as much as ten MMAs).
struct PmWalk {
    int b, i, j;
    __device__ __forceinline__ void start(const WgPmArgs &p, int pl) {
        b = pl / (p.d1 * p.d2);
        const int ij = pl - b * (p.d1 * p.d2);
        i = ij / p.d2; j = ij - i * p.d2;
    }
    __device__ __forceinline__ void next(const WgPmArgs &p) {
        if (++j == p.d2) { j = 0; if (++i == p.d1) { i = 0; ++b; } }
    }
    // mask of the CTA's d2 taps t2a+v whose X plane (i+t1-p, j+t2a+v-p) exists (0: nothing to do)
    __device__ __forceinline__ uint32_t mask(const WgPmArgs &p, int t1, int t2a, int nv) const {
        const int PAD = p.KS / 2;
        const int ii = i + t1 - PAD, jj = j + t2a - PAD;
        if (ii < 0 || ii >= p.d1) return 0u;
        uint32_t m = 0;
        for (int v = 0; v < nv; ++v) if (jj + v >= 0 && jj + v < p.d2) m |= 1u << v;
        return m;
    }
};

__global__ void __launch_bounds__(kPmThreads, 1)
conv4d_tc_wgrad_pm_kernel(const WgPmArgs p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t full_bar[kPmStages], empty_bar[kPmStages], accum_bar;
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // grid.x = (group, t1, tap class), grid.y = slice: the CTAs that run together walk the SAME planes with different
    // taps, so every plane is fetched from HBM once and then served by L2 (slice-major order re-read it per tap: 20 GB)
    int by = blockIdx.x;
    const int cls = by % p.ncls; by /= p.ncls;
    const int t1 = by % p.KS; by /= p.KS;
    const int t2a = cls * p.nav, nv = min(p.nav, p.KS - t2a);
    const int g = by;
    const int s = blockIdx.y;
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
    {   // zero the ring once: the margins around the planes are read by the shifted views and must stay zero
        const uint4 z = make_uint4(0, 0, 0, 0);
        uint4 *d = reinterpret_cast<uint4 *>(smem);
        const uint32_t n16 = (uint32_t)p.stages * p.stage_bytes / 16;
        for (uint32_t e = threadIdx.x; e < n16; e += kPmThreads) d[e] = z;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;
    const uint32_t sm0 = smem_u32_pinned(smem), full0 = smem_u32_pinned(full_bar), empty0 = smem_u32_pinned(empty_bar);

    if (warp == 0) {
        // ===== producer: one lane streams (X plane, G plane) pairs, one bulk copy each =====
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            const size_t ijn = (size_t)p.d1 * p.d2;
            const size_t plane_h = p.plane_bytes / 2;
            PmWalk w;
            w.start(p, pbeg);
            const int PAD = p.KS / 2;
            for (int pl = pbeg; pl < pend; ++pl, w.next(p)) {
                const uint32_t mask = w.mask(p, t1, t2a, nv);
                if (!mask) continue;
                const int b = w.b;
                const size_t ij_g = (size_t)w.i * p.d2 + w.j;
                const long long ij_x = (long long)(w.i + t1 - PAD) * p.d2 + (w.j + t2a - PAD);
                mbar_wait_u32(empty0 + (uint32_t)stage * 8u, phase ^ 1);
                const uint32_t sa = sm0 + (uint32_t)stage * p.stage_bytes, fb = full0 + (uint32_t)stage * 8u;
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fb), "r"((uint32_t)(__popc(mask) + 1) * p.plane_bytes) : "memory");
                const size_t bga = (size_t)b * p.GxA + (p.GxA == 1 ? 0 : g), bgb = (size_t)b * p.G + g;
                for (int v = 0; v < nv; ++v)
                    if ((mask >> v) & 1u)
                        bulk_g2s_u32(sa + (uint32_t)v * p.a1_bytes, p.x + ((long long)(bga * ijn) + ij_x + v) * (long long)plane_h, p.plane_bytes, fb);
                bulk_g2s_u32(sa + p.a_bytes + p.g_off, p.g + (bgb * ijn + ij_g) * plane_h, p.plane_bytes, fb);
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===== MMA issuer: one lane, one M128 x N x K16 MMA per 16 positions =====
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0, inited = 0;
            // descriptor high word: SBO = 256 B (K atoms), version 1, SWIZZLE_32B (layout type 6 at bits 61-63)
            const uint64_t hi = (uint64_t)((256u >> 4) | (1u << 14) | (6u << 29)) << 32;
            const uint32_t lbo_a = (32u >> 4) << 16, lbo_b = (p.lbo_b >> 4) << 16;
            const int nk = p.nkb - p.kb0;
            const uint32_t a1 = p.a1_bytes >> 4, n = (uint32_t)p.N;
            PmWalk w;
            w.start(p, pbeg);
            for (int pl = pbeg; pl < pend; ++pl, w.next(p)) {
                const uint32_t mask = w.mask(p, t1, t2a, nv);
                if (!mask) continue;
                mbar_wait_u32(full0 + (uint32_t)stage * 8u, phase);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t sa4 = (sm0 + (uint32_t)stage * p.stage_bytes) >> 4;
                uint32_t a_lo = lbo_a | (sa4 + (uint32_t)p.kb0 * 32u);                                  // 16 positions = 512 B
                uint32_t b_lo = lbo_b | (sa4 + ((p.a_bytes + p.b_rel) >> 4) + (uint32_t)p.kb0 * 32u);
                if (mask == 3u) {
                    // both d2 taps: K-block-major, the two MMAs of a block share the grad descriptor
                    const uint32_t acc0 = inited & 1u, acc1 = (inited >> 1) & 1u;
                    umma_f16(tmem_base, hi | (uint64_t)a_lo, hi | (uint64_t)b_lo, p.idesc, acc0);
                    umma_f16(tmem_base + n, hi | (uint64_t)(a_lo + a1), hi | (uint64_t)b_lo, p.idesc, acc1);
#pragma unroll 4
                    for (int kb = 1; kb < nk; ++kb) {
                        a_lo += 32u; b_lo += 32u;
                        umma_f16(tmem_base, hi | (uint64_t)a_lo, hi | (uint64_t)b_lo, p.idesc, 1u);
                        umma_f16(tmem_base + n, hi | (uint64_t)(a_lo + a1), hi | (uint64_t)b_lo, p.idesc, 1u);
                    }
                } else {
                    const uint32_t v = mask >> 1;            // the single live variant (mask is 1 or 2)
                    const uint32_t d = tmem_base + v * n;
                    a_lo += v * a1;
                    umma_f16(d, hi | (uint64_t)a_lo, hi | (uint64_t)b_lo, p.idesc, (inited >> v) & 1u);
#pragma unroll 8
                    for (int kb = 1; kb < nk; ++kb) {
                        a_lo += 32u; b_lo += 32u;
                        umma_f16(d, hi | (uint64_t)a_lo, hi | (uint64_t)b_lo, p.idesc, 1u);
                    }
                }
                inited |= mask;
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                             ::"r"(empty0 + (uint32_t)stage * 8u) : "memory");
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
            umma_commit(&accum_bar);
        }
        __syncwarp();
    } else {
        // ===== epilogue: flush D[(t4,c)][(r,o)] as this slice's partial sums =====
        const int q = warp & 3;
        mbar_wait(&accum_bar, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        uint32_t live = 0;      // accumulators the issuer wrote (same walk); untouched ones hold garbage, their partials stay zero
        {
            PmWalk w;
            w.start(p, pbeg);
            for (int pl = pbeg; pl < pend && live != (1u << nv) - 1u; ++pl, w.next(p)) live |= w.mask(p, t1, t2a, nv);
        }
        const int KS = p.KS;
        const int m = q * 32 + lane, t4 = m >> 4, c = m & 15;
        if (q * 2 < KS) {          // warp q holds taps 2q and 2q+1
            const size_t per_group = (size_t)p.Cin * KS * KS * KS * KS * p.cop;
            float *dst = p.part + ((size_t)s * p.G + g) * per_group;
            for (int v = 0; v < nv; ++v) {
                if (!((live >> v) & 1u)) continue;
                for (int n0 = 0; n0 < p.N; n0 += 16) {
                    float val[16];
                    tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(v * p.N + n0), val);
                    const int t3 = KS - 1 - (n0 >> 4);
                    if (t4 < KS && c < p.Cin) {
                        float *row = dst + (((((size_t)c * KS + t1) * KS + (t2a + v)) * KS + t3) * KS + t4) * p.cop;
#pragma unroll
                        for (int o = 0; o < 16; ++o) if (o < p.Cout) row[o] = val[o];
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

// ---------------------------------------------------------------------------------------------
// The same MMA as a WALK along jA (used when the planes are small enough for k + 2 of them to sit in shared memory).
//
// In the kernel above a CTA owns two d2 taps and streams (X, X, G) plane triples: 1.6 planes from L2 per tap-plane of
// 47 MMAs, which holds it to the L2 -> shared-memory rate (39 GB per launch of 32 PF-Pascal volumes, tensor pipe 69 %
// busy), and consecutive MMAs never share an operand.  Here a CTA owns (group, t1, slice of the (b, iA) rows) and walks
// jj along the row: the X plane (iA+t1-p, jj) meets the k grad planes (iA, jj+p-t2), t2 = 0..k-1, which live in a ring
// of k shared-memory buffers -- every step loads ONE X plane and ONE grad plane for k tap-planes of MMAs (0.4 planes
// per tap-plane), and the k MMAs of a 16-position K block share their A tile (X), which the tensor core keeps in its
// collector: 41 instead of 52 cycles per M128 x N80 x K16 MMA (scripts/umma_rate.cu).  The oldest grad plane's tap is
// issued first, alone, so that its buffer is free for the producer while the other k-1 taps run.
// Accumulators: tap t2 -> TMEM columns [t2*N, (t2+1)*N), live for the CTA's whole life, flushed as per-slice partials.
// ---------------------------------------------------------------------------------------------
constexpr int kPwMaxClasses = 10;
constexpr int kPwFlushRows = 27;

struct WgPwArgs {
    const __half *x;    // PM16 [B*GxA][d1*d2][npos][16]
    const __half *g;    // PM16 [B*G][d1*d2][npos][16]
    float *part;        // [S][G][Cin][k][k][k][k][cop]
    int B, G, GxA, d1, d2, N;
    int nkb, kb0;
    int Cin, Cout, cop;
    uint32_t plane_bytes, a1_bytes, gbuf_bytes, g_off, b_rel, lbo_b, idesc;
    int cls_start[kPwMaxClasses + 1];       // CTAs of class c = group * k + t1: [cls_start[c], cls_start[c+1])
};

template <int K>
__global__ void __launch_bounds__(kPmThreads, 1)
conv4d_tc_wgrad_pmw_kernel(const WgPwArgs p) {
    constexpr int PAD = K / 2;
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t xfull[2], xempty[2], gfull[K], gempty[K], accum_bar, drained_bar;
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int c = 0;
    while ((int)blockIdx.x >= p.cls_start[c + 1]) ++c;
    const int s = (int)blockIdx.x - p.cls_start[c], ncta = p.cls_start[c + 1] - p.cls_start[c];
    const int g = c / K, t1 = c % K;
    const int i_lo = max(0, PAD - t1), i_hi = min(p.d1 - 1, p.d1 - 1 + PAD - t1), nvi = i_hi - i_lo + 1;
    const int rows = p.B * nvi, per = (rows + ncta - 1) / ncta;
    const int rbeg = s * per, rend = min(rows, rbeg + per);
    // The tensor core adds into its fp32 accumulators with truncation: over a chain of n MMAs a coherent sum drifts low by
    // ~n * 2^-24 of its value.  The accumulators are therefore flushed into the slice's partial sums (proper fp32 adds)
    // every kPwFlushRows rows (~32 000 MMAs per accumulator at the PF-Pascal shape, about the chain length of the per-plane
    // kernel above; a flush stalls the issuer for the drain) and restarted from zero.
    const int nflush = rend > rbeg ? (rend - rbeg + kPwFlushRows - 1) / kPwFlushRows : 0;

    if (warp == 1 && lane == 0) {
        for (int b = 0; b < 2; ++b) { mbar_init(&xfull[b], 1); mbar_init(&xempty[b], 1); }
        for (int b = 0; b < K; ++b) { mbar_init(&gfull[b], 1); mbar_init(&gempty[b], 1); }
        mbar_init(&accum_bar, 1);
        mbar_init(&drained_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    const uint32_t smem_total = 2u * p.a1_bytes + (uint32_t)K * p.gbuf_bytes;
    {   // zero everything once: the margins around the planes are read by the shifted views and must stay zero
        const uint4 z = make_uint4(0, 0, 0, 0);
        uint4 *d = reinterpret_cast<uint4 *>(smem);
        for (uint32_t e = threadIdx.x; e < smem_total / 16; e += kPmThreads) d[e] = z;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;
    const uint32_t sm0 = smem_u32_pinned(smem), smg = sm0 + 2u * p.a1_bytes;
    const uint32_t xf0 = smem_u32_pinned(xfull), xe0 = smem_u32_pinned(xempty), gf0 = smem_u32_pinned(gfull), ge0 = smem_u32_pinned(gempty);
    const size_t ijn = (size_t)p.d1 * p.d2, plane_h = p.plane_bytes / 2;

    if (warp == 0) {
        // ===== producer: per step of the walk one grad plane (the one that becomes live) and one X plane =====
        if (elect_one()) {
            int xl = 0, gl = 0;                     // X / grad planes loaded so far (buffer = count mod ring size)
            for (int r = rbeg; r < rend; ++r) {
                const int b = r / nvi, i = i_lo + (r - b * nvi);
                const __half *xrow = p.x + (((size_t)b * p.GxA + (p.GxA == 1 ? 0 : g)) * ijn + (size_t)(i + t1 - PAD) * p.d2) * plane_h;
                const __half *grow = p.g + (((size_t)b * p.G + g) * ijn + (size_t)i * p.d2) * plane_h;
                for (int jj = 0; jj < p.d2; ++jj) {
                    const int j_first = jj == 0 ? 0 : jj + PAD, j_last = min(p.d2 - 1, jj + PAD);
                    for (int j = j_first; j <= j_last; ++j, ++gl) {
                        const int buf = gl % K;
                        if (gl >= K) mbar_wait_u32(ge0 + (uint32_t)buf * 8u, (uint32_t)((gl / K) - 1) & 1u);
                        const uint32_t fb = gf0 + (uint32_t)buf * 8u;
                        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fb), "r"(p.plane_bytes) : "memory");
                        bulk_g2s_u32(smg + (uint32_t)buf * p.gbuf_bytes + p.g_off, grow + (size_t)j * plane_h, p.plane_bytes, fb);
                    }
                    const int xb = xl & 1;
                    if (xl >= 2) mbar_wait_u32(xe0 + (uint32_t)xb * 8u, (uint32_t)((xl >> 1) - 1) & 1u);
                    const uint32_t fb = xf0 + (uint32_t)xb * 8u;
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fb), "r"(p.plane_bytes) : "memory");
                    bulk_g2s_u32(sm0 + (uint32_t)xb * p.a1_bytes, xrow + (size_t)jj * plane_h, p.plane_bytes, fb);
                    ++xl;
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (elect_one()) {
            // descriptor high word: SBO = 256 B (K atoms), version 1, SWIZZLE_32B (layout type 6 at bits 61-63)
            const uint64_t hi = (uint64_t)((256u >> 4) | (1u << 14) | (6u << 29)) << 32;
            const uint32_t lbo_a = (32u >> 4) << 16, lbo_b = (p.lbo_b >> 4) << 16;
            const int nk = p.nkb - p.kb0;
            const uint32_t n = (uint32_t)p.N;
            uint32_t inited = 0;
            int xl = 0, gbase = 0;
            for (int r = rbeg; r < rend; ++r) {
                for (int jj = 0; jj < p.d2; ++jj, ++xl) {
                    const int xb = xl & 1;
                    mbar_wait_u32(xf0 + (uint32_t)xb * 8u, (uint32_t)(xl >> 1) & 1u);
                    const int j_first = jj == 0 ? 0 : jj + PAD, j_last = min(p.d2 - 1, jj + PAD);
                    for (int j = j_first; j <= j_last; ++j) {
                        const int gi = gbase + j;
                        mbar_wait_u32(gf0 + (uint32_t)(gi % K) * 8u, (uint32_t)(gi / K) & 1u);
                    }
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const int t2_lo = max(0, jj + PAD - (p.d2 - 1)), t2_hi = min(K - 1, jj + PAD);       // taps whose grad plane j = jj+PAD-t2 exists
                    const uint32_t a0 = lbo_a | (((sm0 + (uint32_t)xb * p.a1_bytes) >> 4) + (uint32_t)p.kb0 * 32u);
                    uint32_t b0[K];
#pragma unroll
                    for (int t2 = 0; t2 < K; ++t2) {
                        const int gi = gbase + jj + PAD - t2;
                        b0[t2] = lbo_b | (((smg + (uint32_t)((gi + K) % K) * p.gbuf_bytes + p.b_rel) >> 4) + (uint32_t)p.kb0 * 32u);
                    }
                    if (t2_lo == 0 && t2_hi == K - 1) {
                        // interior step.  Oldest grad plane first (tap k-1, its last use): its buffer goes back to the producer.
                        {
                            uint32_t a = a0, b = b0[K - 1];
                            const uint32_t d = tmem_base + (uint32_t)(K - 1) * n;
                            umma_f16(d, hi | (uint64_t)a, hi | (uint64_t)b, p.idesc, (inited >> (K - 1)) & 1u);
#pragma unroll 4
                            for (int kb = 1; kb < nk; ++kb) {
                                a += 32u; b += 32u;
                                umma_f16_coll<0, true>(d, hi | (uint64_t)a, hi | (uint64_t)b, p.idesc, 1u);
                            }
                            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                                         ::"r"(ge0 + (uint32_t)((gbase + jj - PAD) % K) * 8u) : "memory");
                        }
                        // the other k-1 taps: K block outer, taps inner, the X tile held in the collector
                        uint32_t a = a0;
                        uint32_t b[K - 1];
#pragma unroll
                        for (int t2 = 0; t2 < K - 1; ++t2) b[t2] = b0[t2];
#pragma unroll
                        for (int t2 = 0; t2 < K - 1; ++t2) {
                            const uint32_t acc = (inited >> t2) & 1u, d = tmem_base + (uint32_t)t2 * n;
                            if (t2 == 0) umma_f16_coll<1, false>(d, hi | (uint64_t)a, hi | (uint64_t)b[t2], p.idesc, acc);
                            else if (t2 == K - 2) umma_f16_coll<3, false>(d, hi | (uint64_t)a, hi | (uint64_t)b[t2], p.idesc, acc);
                            else umma_f16_coll<2, false>(d, hi | (uint64_t)a, hi | (uint64_t)b[t2], p.idesc, acc);
                        }
#pragma unroll 2
                        for (int kb = 1; kb < nk; ++kb) {
                            a += 32u;
#pragma unroll
                            for (int t2 = 0; t2 < K - 1; ++t2) {
                                b[t2] += 32u;
                                const uint32_t d = tmem_base + (uint32_t)t2 * n;
                                if (t2 == 0) umma_f16_coll<1, true>(d, hi | (uint64_t)a, hi | (uint64_t)b[t2], p.idesc, 1u);
                                else if (t2 == K - 2) umma_f16_coll<3, true>(d, hi | (uint64_t)a, hi | (uint64_t)b[t2], p.idesc, 1u);
                                else umma_f16_coll<2, true>(d, hi | (uint64_t)a, hi | (uint64_t)b[t2], p.idesc, 1u);
                            }
                        }
                        inited |= (1u << K) - 1u;
                    } else {
                        // first / last columns of the row: the live taps one after the other, oldest grad plane first
                        for (int t2 = t2_hi; t2 >= t2_lo; --t2) {
                            uint32_t a = a0, b = b0[0];
#pragma unroll
                            for (int u = 1; u < K; ++u) if (u == t2) b = b0[u];
                            const uint32_t d = tmem_base + (uint32_t)t2 * n;
                            umma_f16(d, hi | (uint64_t)a, hi | (uint64_t)b, p.idesc, (inited >> t2) & 1u);
#pragma unroll 4
                            for (int kb = 1; kb < nk; ++kb) {
                                a += 32u; b += 32u;
                                umma_f16_coll<0, true>(d, hi | (uint64_t)a, hi | (uint64_t)b, p.idesc, 1u);
                            }
                            inited |= 1u << t2;
                            const int j = jj + PAD - t2;
                            if (jj == min(j + PAD, p.d2 - 1))         // this grad plane's last use
                                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                                             ::"r"(ge0 + (uint32_t)((gbase + j) % K) * 8u) : "memory");
                        }
                    }
                    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                                 ::"r"(xe0 + (uint32_t)xb * 8u) : "memory");
                }
                gbase += p.d2;
                const int done = r - rbeg + 1;
                if (done % kPwFlushRows == 0 || r + 1 == rend) {
                    umma_commit(&accum_bar);                    // the epilogue warps add the accumulators into the partial sums ...
                    if (r + 1 < rend) {                         // ... and hand them back
                        mbar_wait(&drained_bar, (uint32_t)((done / kPwFlushRows) - 1) & 1u);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        inited = 0;
                    }
                }
            }
        }
        __syncwarp();
    } else {
        // ===== epilogue: add D[t2][(t4,c)][(r,o)] into this slice's partial sums at every flush =====
        const int q = warp & 3;
        const int m = q * 32 + lane, t4 = m >> 4, ch = m & 15;
        const size_t per_group = (size_t)p.Cin * K * K * K * K * p.cop;
        float *dst = p.part + ((size_t)s * p.G + g) * per_group;
        for (int f = 0; f < nflush; ++f) {
            mbar_wait_sleep(&accum_bar, (uint32_t)f & 1u, 1000);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (q * 2 < K) {          // warp q holds the d4 taps 2q and 2q+1
                for (int t2 = 0; t2 < K; ++t2) {
                    for (int n0 = 0; n0 < p.N; n0 += 16) {
                        float val[16];
                        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(t2 * p.N + n0), val);
                        const int t3 = K - 1 - (n0 >> 4);
                        if (t4 < K && ch < p.Cin) {
                            float *row = dst + (((((size_t)ch * K + t1) * K + t2) * K + t3) * K + t4) * p.cop;
                            if (p.cop == 16) {          // 64-byte rows: four 16-byte read-modify-writes
                                float4 *r4 = reinterpret_cast<float4 *>(row);
                                float4 acc4[4];
#pragma unroll
                                for (int v = 0; v < 4; ++v) acc4[v] = f ? r4[v] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                                for (int v = 0; v < 4; ++v)
                                    r4[v] = make_float4(acc4[v].x + val[4 * v], acc4[v].y + val[4 * v + 1], acc4[v].z + val[4 * v + 2], acc4[v].w + val[4 * v + 3]);
                            } else {
#pragma unroll
                                for (int o = 0; o < 16; ++o) if (o < p.Cout) row[o] = (f ? row[o] : 0.f) + val[o];
                            }
                        }
                    }
                }
            }
            if (f + 1 < nflush) {
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                asm volatile("bar.sync 1, 128;" ::: "memory");
                if (threadIdx.x == 64) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&drained_bar)) : "memory");
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
}

struct WgPmPlan {
    bool ok;
    int N, nkb, kb0, stages, nav, ncls;
    uint32_t plane_bytes, a1_bytes, a_bytes, stage_bytes, g_off, b_rel, lbo_b, tmem_cols;
    uint32_t stage_bytes_b;         // one grad buffer with its margins (a multiple of 256 B)
    size_t smem;
};

static WgPmPlan pm_plan(int d3, int d4, int k) {
    WgPmPlan w{};
    w.ok = false;
    if (k != 3 && k != 5) return w;
    const int pad = k / 2, P = d4 + 2 * pad;
    const int npos = (d3 + 2 * pad) * P;
    const int nkb_full = (npos + 15) / 16;
    // K blocks outside [kb0, nkb) multiply nothing but halo zeros (every used shifted view of X is zero there)
    int lo = pad * P + pad - 7;
    if (lo < 0) lo = 0;
    const int hi = (pad + d3 - 1) * P + pad + d4;
    w.kb0 = lo / 16;
    w.nkb = (hi + 15) / 16 < nkb_full ? (hi + 15) / 16 : nkb_full;
    auto up8 = [](int v) { return (v + 7) / 8 * 8; };
    const int npos_a = up8(npos > w.nkb * 16 + 8 ? npos : w.nkb * 16 + 8);
    const int g_off_pos = up8(pad * P - pad);
    const int reach = w.nkb * 16 + pad * P + pad;
    const int npos_b = up8(g_off_pos + (npos > reach ? npos : reach));
    w.plane_bytes = (uint32_t)npos * 32u;
    w.a1_bytes = (uint32_t)npos_a * 32u;
    w.g_off = (uint32_t)g_off_pos * 32u;
    w.b_rel = (uint32_t)(g_off_pos - pad * P + pad) * 32u;
    w.lbo_b = (uint32_t)P * 32u;
    w.stage_bytes_b = (uint32_t)npos_b * 32u;
    if ((w.lbo_b >> 4) > 16383u) return w;
    // two d2 taps per CTA (one grad plane staged per two X planes: the kernel is close to the L2->shared ceiling) when a
    // 2-deep ring of such stages fits; else one tap and a deeper ring
    w.N = 16 * k;
    for (w.nav = 2; w.nav >= 1; --w.nav) {
        w.a_bytes = (uint32_t)w.nav * w.a1_bytes;
        w.stage_bytes = ((w.a_bytes + (uint32_t)npos_b * 32u + 1023u) / 1024u) * 1024u;
        int stages = (int)((220u * 1024u) / w.stage_bytes);
        if (stages > kPmStages) stages = kPmStages;
        w.stages = stages;
        if (stages >= 2 && w.nav * w.N <= 512) break;
    }
    if (w.nav < 1) return w;
    w.ncls = (k + w.nav - 1) / w.nav;
    w.smem = (size_t)w.stages * w.stage_bytes;
    const int cols = w.nav * w.N;
    w.tmem_cols = cols <= 32 ? 32 : (cols <= 64 ? 64 : (cols <= 128 ? 128 : (cols <= 256 ? 256 : 512)));
    w.ok = true;
    return w;
}

bool conv4d_tc_wgrad_pm_supported(int d3, int d4, int k) { return pm_plan(d3, d4, k).ok; }

// x_cb / g_cb: CB16 operands; pm_scratch: room for both PM16 copies (same byte counts as the CB16 tensors);
// part: [S][G][Cin][k][k][k][k][cop], already zeroed.  Returns 2 when the shape is outside the envelope.
int conv4d_tc_wgrad_pm_run(const __half *x_cb, const __half *g_cb, const __half *x_pm_in, const __half *g_pm_in, float *part,
                           void *pm_scratch, int B, int G, int GxA, int Cin, int Cout, int d1, int d2, int d3, int d4, int k,
                           int S, int cop, cudaStream_t st) {
    const WgPmPlan w = pm_plan(d3, d4, k);
    if (!w.ok || Cin > 16 || Cout > 16) return 2;
    const int pad = k / 2;
    const int npos = (d3 + 2 * pad) * (d4 + 2 * pad);
    const size_t ij = (size_t)d1 * d2;
    const size_t xa_bytes = (size_t)B * GxA * ij * npos * 32, gb_bytes = (size_t)B * G * ij * npos * 32;
    // operands that already come with their position-major twin (NCNET_CB16PM) skip the conversion pass
    const __half *x_pm = x_pm_in, *g_pm = g_pm_in;
    if (!x_pm) {
        __half *dst = (__half *)pm_scratch;
        pm16_from_cb16(x_cb, dst, (size_t)B * GxA, ij, npos, st);
        NC_LAUNCH_OK("pm16_from_cb16_kernel");
        x_pm = dst;
    }
    if (!g_pm) {
        __half *dst = (__half *)((uint8_t *)pm_scratch + ((xa_bytes + 255) & ~(size_t)255));
        pm16_from_cb16(g_cb, dst, (size_t)B * G, ij, npos, st);
        NC_LAUNCH_OK("pm16_from_cb16_kernel");
        g_pm = dst;
    }
    (void)gb_bytes;
    // planes small enough for k + 2 of them in shared memory: the walk along jA (one X + one grad plane per k tap-planes)
    {
        const uint32_t gbuf_bytes = w.stage_bytes_b;
        const size_t smem_w = 2 * (size_t)w.a1_bytes + (size_t)k * gbuf_bytes;
        const int ncls = G * k;
        if (smem_w <= 225u * 1024u && k * w.N <= 512 && ncls <= kPwMaxClasses && d1 >= k && d2 > k && !getenv("NCNET_NO_PMW")) {
            WgPwArgs q{};
            q.x = x_pm; q.g = g_pm; q.part = part;
            q.B = B; q.G = G; q.GxA = GxA; q.d1 = d1; q.d2 = d2; q.N = w.N;
            q.nkb = w.nkb; q.kb0 = w.kb0; q.Cin = Cin; q.Cout = Cout; q.cop = cop;
            q.plane_bytes = w.plane_bytes; q.a1_bytes = w.a1_bytes; q.gbuf_bytes = gbuf_bytes;
            q.g_off = w.g_off; q.b_rel = w.b_rel; q.lbo_b = w.lbo_b;
            q.idesc = (1u << 4) | (1u << 15) | (1u << 16) | ((uint32_t)(w.N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
            // CTAs per class (group, t1) in proportion to its valid output rows; at most S slices of partial sums each
            const int pad_ = k / 2;
            int wsum = 0, wc[kPwMaxClasses], nc[kPwMaxClasses];
            for (int c = 0; c < ncls; ++c) { const int t1 = c % k; wc[c] = d1 - (t1 > pad_ ? t1 - pad_ : pad_ - t1); wsum += wc[c]; }
            int used_ctas = 0;
            for (int c = 0; c < ncls; ++c) {
                int v = (int)((long long)kNumSMs * wc[c] / wsum);
                const long long rows = (long long)B * wc[c];
                if (v > S) v = S;
                if (v > rows) v = (int)rows;
                if (v < 1) v = 1;
                nc[c] = v; used_ctas += v;
            }
            for (bool more = true; more && used_ctas < kNumSMs;) {      // hand out the remainder to the most loaded classes
                more = false;
                int best = -1;
                double load = 0.0;
                for (int c = 0; c < ncls; ++c) {
                    const double l = (double)wc[c] / nc[c];
                    if (nc[c] < S && nc[c] < (long long)B * wc[c] && l > load) { load = l; best = c; }
                }
                if (best >= 0) { ++nc[best]; ++used_ctas; more = true; }
            }
            q.cls_start[0] = 0;
            for (int c = 0; c < ncls; ++c) q.cls_start[c + 1] = q.cls_start[c] + nc[c];
            for (int c = ncls + 1; c <= kPwMaxClasses; ++c) q.cls_start[c] = q.cls_start[ncls];
            cudaError_t e = k == 5 ? cudaFuncSetAttribute(conv4d_tc_wgrad_pmw_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_w)
                                   : cudaFuncSetAttribute(conv4d_tc_wgrad_pmw_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_w);
            NC_CHECK(e == cudaSuccess, "conv4d_wgrad(tc,pm walk): smem attribute: %s", cudaGetErrorString(e));
            if (k == 5) conv4d_tc_wgrad_pmw_kernel<5><<<q.cls_start[ncls], kPmThreads, smem_w, st>>>(q);
            else conv4d_tc_wgrad_pmw_kernel<3><<<q.cls_start[ncls], kPmThreads, smem_w, st>>>(q);
            NC_LAUNCH_OK("conv4d_tc_wgrad_pmw_kernel");
            return 0;
        }
    }
    WgPmArgs a{};
    a.x = x_pm; a.g = g_pm; a.part = part;
    a.B = B; a.G = G; a.GxA = GxA; a.d1 = d1; a.d2 = d2; a.KS = k; a.N = w.N;
    a.nkb = w.nkb; a.kb0 = w.kb0; a.S = S; a.Cin = Cin; a.Cout = Cout; a.cop = cop; a.stages = w.stages;
    a.plane_bytes = w.plane_bytes; a.a1_bytes = w.a1_bytes; a.a_bytes = w.a_bytes; a.stage_bytes = w.stage_bytes;
    a.nav = w.nav; a.ncls = w.ncls;
    a.g_off = w.g_off; a.b_rel = w.b_rel; a.lbo_b = w.lbo_b; a.tmem_cols = w.tmem_cols;
    // D = f32, A = B = f16, both MN-major (bits 15/16), N>>3 at bit 17, M = 128 -> 8 at bit 24
    a.idesc = (1u << 4) | (1u << 15) | (1u << 16) | ((uint32_t)(w.N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    cudaError_t e = cudaFuncSetAttribute(conv4d_tc_wgrad_pm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)w.smem);
    NC_CHECK(e == cudaSuccess, "conv4d_wgrad(tc,pm): smem attribute: %s", cudaGetErrorString(e));
    dim3 grid(G * k * w.ncls, S);
    conv4d_tc_wgrad_pm_kernel<<<grid, kPmThreads, w.smem, st>>>(a);
    NC_LAUNCH_OK("conv4d_tc_wgrad_pm_kernel");
    return 0;
}

}  // namespace ncnet
