// Conv4d weight gradient of the single-channel layers (Cin == 1 or Cout == 1) on M128 tiles (mode NCNET_CONV_TC16).
//
//   Cin == 1 :  dW[o,t1,t2,t3,t4] = sum G[o; i,j,k,l] * x[i+t1-p, j+t2-p, k+t3-p, l+t4-p]
//   Cout == 1:  dW[c,t1,t2,t3,t4] = sum g[i,j,k,l]    * X[c; i+t1-p, j+t2-p, k+t3-p, l+t4-p]     (lib/conv4d.py:11-51)
//
// The 16-channel tensor ("W16": G, or X) is the A operand in the position-major SWIZZLE_32B layout of
// conv4d_tc_wgrad_pm.cu, its atoms are the d3 taps (row-shifted views, 5 of 8 used, M = 128).  The single-channel
// tensor is im2col'ed over its (d2,d4) taps into 16-"channel" blocks ("T2": k=5 -> 25 taps in 2 blocks, k=3 -> 9 in 1)
// by a pack kernel that writes the same position-major layout; the blocks of one plane are the N atoms (N = 32 / 16).
// One M128 x N32 x K16 MMA per 16 positions then covers all k^3 (d2,d3,d4) taps of one d1 tap -- the channel-blocked
// kernel needs five half-rate M64 x N16 MMAs for that.  With anchor plane (i,j) of W16 the T2 plane is (i +/- (t1-p), j):
//   Cin == 1 :  T2[(k,l)][(t2,t4)] = x[i+t1-p, j+t2-p, k, l+t4-p],  W16 row shift  r = k-1-t3  (reversed)
//   Cout == 1:  T2[(k,l)][(t2,t4)] = g[i-t1+p, j-t2+p, k, l-t4+p],  W16 row shift  r = t3
// and in both cases  W16 index = T2 index + (r - p) * P + p  in haloed flattened positions.
// CTA = (group, t1, slice of the plane list); partial sums per slice in the layout unpack_wgrad_kernel reduces.
#include "conv4d_tc.cuh"
#include "tc_common.cuh"

namespace ncnet {

using namespace tc;

constexpr int kThinThreads = 192;
constexpr int kThinStages = 3;

// Single-channel planar [nblk, vol] -> PM16 tap blocks [nblk*ngrp][ij][npos][16] (SWIZZLE_32B half swap), haloed rows,
// column index = the un-haloed l (columns >= d4 are zero).  mirror = 0: x[.., j+t2-p, k, l+t4-p];  1: minus signs.
// Block = one (i,j) plane of one (output block, channel half); compile-time tap octet as in cb16_pack_taps2d_narrow_kernel.
template <typename T, int k, int BASE>
__device__ __forceinline__ void pm_taps_rows(const T *__restrict__ xp, __half *__restrict__ dst, int jx, int d2, int d3, int d4,
                                             size_t plane, int half, int mirror, int warp, int lane) {
    constexpr int pad = k / 2, T2A = BASE / k;
    const int P = d4 + 2 * pad, rows = d3 + 2 * pad;
    const int sgn = mirror ? -1 : 1;
    // lane holds source column lane - pad; channel tap t4 of output column l needs source column l + sgn*(t4 - pad)
    const int col = lane - pad;
    const bool col_ok = col >= 0 && col < d4;
    for (int kk = warp; kk < rows; kk += 8) {
        const int kq = kk - pad;
        const bool row_ok = kq >= 0 && kq < d3;
        float e[3];
#pragma unroll
        for (int u = 0; u < 3; ++u) {
            const int dj = sgn * (T2A + u - pad), jt = jx + dj;
            e[u] = (row_ok && col_ok && T2A + u < k && jt >= 0 && jt < d2)
                       ? ldf(xp + (ptrdiff_t)dj * (ptrdiff_t)plane + (size_t)kq * d4 + col) : 0.f;
        }
        __align__(16) __half v[8];
#pragma unroll
        for (int c8 = 0; c8 < 8; ++c8) {
            const int c = BASE + c8;                     // compile-time after unrolling
            const int t2 = c / k, t4 = c - t2 * k;
            float o = 0.f;
            if (c < k * k) {
                const int srcl = lane + pad + sgn * (t4 - pad);      // lane that holds column (lane) + sgn*(t4-pad)
                const float got = __shfl_sync(0xffffffffu, e[t2 - T2A], srcl & 31);
                o = (srcl >= 0 && srcl < 32 && lane < d4) ? got : 0.f;
            }
            v[c8] = __float2half_rn(o);
        }
        if (lane < P) {
            const int pos = kk * P + lane;
            const int hs = half ^ ((pos >> 2) & 1);
            *reinterpret_cast<uint4 *>(dst + ((size_t)pos * 2 + hs) * 8) = *reinterpret_cast<const uint4 *>(v);
        }
    }
}

template <typename T, int k>
__global__ void __launch_bounds__(256)
pm16_pack_taps2d_kernel(const T *__restrict__ x, __half *__restrict__ pm, int d1, int d2, int d3, int d4, int mirror) {
    constexpr int pad = k / 2, ngrp = (k * k + 15) / 16;
    const int jx = blockIdx.x, ix = blockIdx.y;
    const unsigned z = blockIdx.z, blk_out = z >> 1, half = z & 1u;
    const unsigned blk = blk_out / ngrp, gr = blk_out - blk * ngrp;
    const int P = d4 + 2 * pad, rows = d3 + 2 * pad;
    const size_t ij = (size_t)d1 * d2, plane = (size_t)d3 * d4;
    const size_t ijx = (size_t)ix * d2 + jx;
    const T *xp = x + ((size_t)blk * ij + ijx) * plane;
    __half *dst = pm + (((size_t)blk_out * ij + ijx) * rows * P) * 16;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int base = (int)(gr * 16 + half * 8);
    if (base == 0) pm_taps_rows<T, k, 0>(xp, dst, jx, d2, d3, d4, plane, (int)half, mirror, warp, lane);
    else if (base == 8) pm_taps_rows<T, k, 8>(xp, dst, jx, d2, d3, d4, plane, (int)half, mirror, warp, lane);
    else if (base == 16) pm_taps_rows<T, k, 16>(xp, dst, jx, d2, d3, d4, plane, (int)half, mirror, warp, lane);
    else pm_taps_rows<T, k, 24>(xp, dst, jx, d2, d3, d4, plane, (int)half, mirror, warp, lane);
}

struct WgThinArgs {
    const __half *w16;  // PM16 [B*G][d1*d2][npos][16]      the 16-channel tensor (A operand)
    const __half *t2;   // PM16 [B*Gt*ngrp][d1*d2][npos][16] tap blocks of the single-channel tensor (B operand)
    float *part;        // [S][G][Cin][k][k][k][k][cop]
    int B, G, Gt, d1, d2, KS, N, ngrp;
    int sgn;            // T2 plane = (i + sgn*(t1-p), j): +1 for Cin == 1, -1 for Cout == 1
    int cin1;           // 1: Cin == 1 (rows of D are out-channels, d3 atoms reversed); 0: Cout == 1 (rows are in-channels)
    int nkb, kb0, S, C16, cop, stages;
    uint32_t plane_bytes, a_bytes, b1_bytes, stage_bytes;
    uint32_t a_off;     // byte offset of the W16 plane inside the A buffer (multiple of 256)
    uint32_t a_rel;     // byte offset, from the A buffer, of atom 0 at K position 0
    uint32_t lbo_a;     // P * 32
    uint32_t idesc, tmem_cols;
};

struct ThinWalk {
    int b, i, j;
    __device__ __forceinline__ void start(const WgThinArgs &p, int pl) {
        b = pl / (p.d1 * p.d2);
        const int ij = pl - b * (p.d1 * p.d2);
        i = ij / p.d2; j = ij - i * p.d2;
    }
    __device__ __forceinline__ void next(const WgThinArgs &p) {
        if (++j == p.d2) { j = 0; if (++i == p.d1) { i = 0; ++b; } }
    }
    // The walk runs over the T2 planes (the tap-packed single-channel tensor: the larger stream, ngrp planes per
    // step), so the k tap classes that run together read the SAME T2 plane and L2 serves k-1 of them; the 16-channel
    // plane each class pairs it with is row  i - sgn*(t1 - p)  (returned here, -1 when it falls outside the volume)
    // and is the one re-read from DRAM a few rows later.  (Anchoring the walk on the 16-channel planes re-read the
    // T2 planes instead: 5.5 GB of DRAM reads per launch for 0.43 GB of input at the PF-Pascal shape.)
    __device__ __forceinline__ int w16_row(const WgThinArgs &p, int t1) const {
        const int ii = i - p.sgn * (t1 - p.KS / 2);
        return (ii < 0 || ii >= p.d1) ? -1 : ii;
    }
};

__global__ void __launch_bounds__(kThinThreads, 1)
conv4d_tc_wgrad_thin_kernel(const WgThinArgs p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t full_bar[kThinStages], empty_bar[kThinStages], accum_bar;
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int t1 = blockIdx.x % p.KS, g = blockIdx.x / p.KS;       // tap-major grid: co-resident CTAs walk the same planes
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
        for (uint32_t e = threadIdx.x; e < n16; e += kThinThreads) d[e] = z;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;
    const uint32_t sm0 = smem_u32_pinned(smem), full0 = smem_u32_pinned(full_bar), empty0 = smem_u32_pinned(empty_bar);

    if (warp == 0) {
        // ===== producer: W16 plane + the ngrp tap-block planes of the T2 plane, one bulk copy each =====
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            const size_t ijn = (size_t)p.d1 * p.d2;
            const size_t plane_h = p.plane_bytes / 2;
            ThinWalk w;
            w.start(p, pbeg);
            for (int pl = pbeg; pl < pend; ++pl, w.next(p)) {
                const int ii = w.w16_row(p, t1);
                if (ii < 0) continue;
                mbar_wait_u32(empty0 + (uint32_t)stage * 8u, phase ^ 1);
                const uint32_t sa = sm0 + (uint32_t)stage * p.stage_bytes, fb = full0 + (uint32_t)stage * 8u;
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fb), "r"((uint32_t)(1 + p.ngrp) * p.plane_bytes) : "memory");
                const size_t bga = (size_t)w.b * p.G + g, bgt = (size_t)w.b * p.Gt + (p.Gt == 1 ? 0 : g);
                bulk_g2s_u32(sa + p.a_off, p.w16 + (bga * ijn + (size_t)ii * p.d2 + w.j) * plane_h, p.plane_bytes, fb);
                for (int gr = 0; gr < p.ngrp; ++gr)
                    bulk_g2s_u32(sa + p.a_bytes + (uint32_t)gr * p.b1_bytes,
                                 p.t2 + ((bgt * p.ngrp + gr) * ijn + (size_t)w.i * p.d2 + w.j) * plane_h, p.plane_bytes, fb);
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===== MMA issuer: one lane, one M128 x N x K16 MMA per 16 positions =====
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0, acc = 0;
            // descriptor high word: SBO = 256 B (K atoms), version 1, SWIZZLE_32B (layout type 6 at bits 61-63)
            const uint64_t hi = (uint64_t)((256u >> 4) | (1u << 14) | (6u << 29)) << 32;
            const uint32_t lbo_a = (p.lbo_a >> 4) << 16, lbo_b = (p.b1_bytes >> 4) << 16;
            const int nk = p.nkb - p.kb0;
            ThinWalk w;
            w.start(p, pbeg);
            for (int pl = pbeg; pl < pend; ++pl, w.next(p)) {
                if (w.w16_row(p, t1) < 0) continue;
                mbar_wait_u32(full0 + (uint32_t)stage * 8u, phase);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t sa4 = (sm0 + (uint32_t)stage * p.stage_bytes) >> 4;
                uint32_t a_lo = lbo_a | (sa4 + (p.a_rel >> 4) + (uint32_t)p.kb0 * 32u);         // 16 positions = 512 B
                uint32_t b_lo = lbo_b | (sa4 + (p.a_bytes >> 4) + (uint32_t)p.kb0 * 32u);
                umma_f16(tmem_base, hi | (uint64_t)a_lo, hi | (uint64_t)b_lo, p.idesc, acc);
                acc = 1u;
#pragma unroll 8
                for (int kb = 1; kb < nk; ++kb) {
                    a_lo += 32u; b_lo += 32u;
                    umma_f16(tmem_base, hi | (uint64_t)a_lo, hi | (uint64_t)b_lo, p.idesc, 1u);
                }
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                             ::"r"(empty0 + (uint32_t)stage * 8u) : "memory");
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
            umma_commit(&accum_bar);
        }
        __syncwarp();
    } else {
        // ===== epilogue: D[(r, ch16)][(block, tap channel)] -> this slice's partial sums =====
        const int q = warp & 3;
        mbar_wait(&accum_bar, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        bool any = false;           // no MMA issued: the accumulator is garbage, the (zeroed) partials stay
        {
            ThinWalk w;
            w.start(p, pbeg);
            for (int pl = pbeg; pl < pend && !any; ++pl, w.next(p)) any = w.w16_row(p, t1) >= 0;
        }
        const int KS = p.KS;
        const int m = q * 32 + lane, r = m >> 4, ch = m & 15;
        if (any && q * 2 < KS) {
            const int t3 = p.cin1 ? KS - 1 - r : r;
            const size_t per_group = (size_t)(p.cin1 ? 1 : p.C16) * KS * KS * KS * KS * p.cop;
            float *dst = p.part + ((size_t)s * p.G + g) * per_group;
            for (int n0 = 0; n0 < p.N; n0 += 16) {
                float v[16];
                tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)n0, v);
                if (r < KS && ch < p.C16) {
#pragma unroll
                    for (int c = 0; c < 16; ++c) {
                        const int tap = n0 + c;
                        if (tap < KS * KS) {
                            const int t2 = tap / KS, t4 = tap - t2 * KS;
                            const int cin = p.cin1 ? 0 : ch, o = p.cin1 ? ch : 0;
                            dst[(((((size_t)cin * KS + t1) * KS + t2) * KS + t3) * KS + t4) * p.cop + o] = v[c];
                        }
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

struct WgThinPlan {
    bool ok;
    int N, ngrp, nkb, kb0, stages;
    uint32_t plane_bytes, a_bytes, b1_bytes, stage_bytes, a_off, a_rel, lbo_a, tmem_cols;
    size_t smem;
};

static WgThinPlan thin_plan(int d3, int d4, int k) {
    WgThinPlan w{};
    w.ok = false;
    if ((k != 3 && k != 5) || d4 + 2 * (k / 2) > 32) return w;         // the tap pack kernel holds a row in one warp
    const int pad = k / 2, P = d4 + 2 * pad;
    const int npos = (d3 + 2 * pad) * P;
    // K runs over the T2 positions; outside rows [p, p+d3) T2 is zero
    const int lo = pad * P, hi = (pad + d3 - 1) * P + d4;
    w.kb0 = lo / 16;
    w.nkb = (hi + 15) / 16;
    auto up8 = [](int v) { return (v + 7) / 8 * 8; };
    const int a_off_pos = up8(pad * P - pad);
    const int reach = w.nkb * 16 + (7 - pad) * P + pad + 1;             // atoms 5..7 are read too (discarded)
    const int npos_a = up8(a_off_pos + (npos > reach ? npos : reach));
    const int npos_b = up8(npos > w.nkb * 16 + 8 ? npos : w.nkb * 16 + 8);
    w.ngrp = (k * k + 15) / 16;
    w.N = 16 * w.ngrp;
    w.plane_bytes = (uint32_t)npos * 32u;
    w.a_bytes = (uint32_t)npos_a * 32u;
    w.b1_bytes = (uint32_t)npos_b * 32u;
    w.a_off = (uint32_t)a_off_pos * 32u;
    w.a_rel = (uint32_t)(a_off_pos - pad * P + pad) * 32u;
    w.lbo_a = (uint32_t)P * 32u;
    if ((w.lbo_a >> 4) > 16383u || (w.b1_bytes >> 4) > 16383u) return w;
    w.stage_bytes = ((w.a_bytes + (uint32_t)w.ngrp * w.b1_bytes + 1023u) / 1024u) * 1024u;
    int stages = (int)((220u * 1024u) / w.stage_bytes);
    if (stages > kThinStages) stages = kThinStages;
    if (stages < 2) return w;
    w.stages = stages;
    w.smem = (size_t)stages * w.stage_bytes;
    w.tmem_cols = 32;
    w.ok = true;
    return w;
}

bool conv4d_tc_wgrad_thin_supported(int d3, int d4, int k) { return thin_plan(d3, d4, k).ok; }

// pm16_from_cb16_kernel lives in conv4d_tc_wgrad_pm.cu
void pm16_from_cb16(const __half *cb, __half *pm, size_t nblk, size_t ij, int npos, cudaStream_t st);

// w16_cb: the 16-channel tensor in CB16 ([B*G] blocks);  one: the single-channel tensor, planar [B*Gt, vol] (dtype);
// cin1 = 1: Cin == 1 (w16 = grad_y, one = x);  0: Cout == 1 (w16 = x, one = grad_y).  scratch: room for the PM16 copy
// of w16 (same bytes as its CB16 form) + the tap blocks (2x that for k = 5).  part: zeroed partial sums.
int conv4d_tc_wgrad_thin_run(const __half *w16_cb, const __half *w16_pm_in, const void *one, int one_dtype, float *part, void *scratch, int cin1,
                             int B, int G, int Gt, int C16, int d1, int d2, int d3, int d4, int k, int S, int cop,
                             cudaStream_t st) {
    const WgThinPlan w = thin_plan(d3, d4, k);
    if (!w.ok || C16 > 16) return 2;
    NC_CHECK(one_dtype == NCNET_F32 || one_dtype == NCNET_F16, "conv4d_wgrad(tc,thin): bad dtype %d", one_dtype);
    const int pad = k / 2;
    const int npos = (d3 + 2 * pad) * (d4 + 2 * pad);
    const size_t ij = (size_t)d1 * d2;
    const size_t w16_bytes = ((size_t)B * G * ij * npos * 32 + 255) & ~(size_t)255;
    const __half *w16_pm = w16_pm_in;         // NCNET_CB16PM operands bring their position-major twin
    __half *t2_pm = (__half *)((uint8_t *)scratch + w16_bytes);
    if (!w16_pm) {
        pm16_from_cb16(w16_cb, (__half *)scratch, (size_t)B * G, ij, npos, st);
        NC_LAUNCH_OK("pm16_from_cb16_kernel");
        w16_pm = (const __half *)scratch;
    }
    {
        const long long zb = (long long)B * Gt * w.ngrp * 2;
        NC_CHECK(zb <= 65535 && d1 <= 65535, "conv4d_wgrad(tc,thin): grid too large");
        dim3 grid(d2, d1, (unsigned)zb);
        const int mirror = cin1 ? 0 : 1;
        if (one_dtype == NCNET_F32) {
            if (k == 5) pm16_pack_taps2d_kernel<float, 5><<<grid, 256, 0, st>>>((const float *)one, t2_pm, d1, d2, d3, d4, mirror);
            else pm16_pack_taps2d_kernel<float, 3><<<grid, 256, 0, st>>>((const float *)one, t2_pm, d1, d2, d3, d4, mirror);
        } else {
            if (k == 5) pm16_pack_taps2d_kernel<__half, 5><<<grid, 256, 0, st>>>((const __half *)one, t2_pm, d1, d2, d3, d4, mirror);
            else pm16_pack_taps2d_kernel<__half, 3><<<grid, 256, 0, st>>>((const __half *)one, t2_pm, d1, d2, d3, d4, mirror);
        }
        NC_LAUNCH_OK("pm16_pack_taps2d_kernel");
    }
    WgThinArgs a{};
    a.w16 = w16_pm; a.t2 = t2_pm; a.part = part;
    a.B = B; a.G = G; a.Gt = Gt; a.d1 = d1; a.d2 = d2; a.KS = k; a.N = w.N; a.ngrp = w.ngrp;
    a.sgn = cin1 ? 1 : -1; a.cin1 = cin1;
    a.nkb = w.nkb; a.kb0 = w.kb0; a.S = S; a.C16 = C16; a.cop = cop; a.stages = w.stages;
    a.plane_bytes = w.plane_bytes; a.a_bytes = w.a_bytes; a.b1_bytes = w.b1_bytes; a.stage_bytes = w.stage_bytes;
    a.a_off = w.a_off; a.a_rel = w.a_rel; a.lbo_a = w.lbo_a; a.tmem_cols = w.tmem_cols;
    // D = f32, A = B = f16, both MN-major (bits 15/16), N>>3 at bit 17, M = 128 -> 8 at bit 24
    a.idesc = (1u << 4) | (1u << 15) | (1u << 16) | ((uint32_t)(w.N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    cudaError_t e = cudaFuncSetAttribute(conv4d_tc_wgrad_thin_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)w.smem);
    NC_CHECK(e == cudaSuccess, "conv4d_wgrad(tc,thin): smem attribute: %s", cudaGetErrorString(e));
    dim3 grid(G * k, S);
    conv4d_tc_wgrad_thin_kernel<<<grid, kThinThreads, w.smem, st>>>(a);
    NC_LAUNCH_OK("conv4d_tc_wgrad_thin_kernel");
    return 0;
}

}  // namespace ncnet
