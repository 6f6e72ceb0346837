// Single-input-channel Conv4d layers (mode M4: Cin == 1, Cout <= 16 -- the first layer of a NeighConsensus stack and
// the input-gradient convolution of its last layer) as a persistent WALK along iA with CB16 output
// (reference semantics: lib/conv4d.py:11-51).
//
// Mode M4 turns the (d2,d4) taps of the single channel into the 16 "channels" of a CB16 block (im2col by the pack
// kernel: k=5 -> 25 taps in 2 blocks, k=3 -> 9 taps in 1) and the d3 taps into row-shifted K-steps, which leaves the
// d1 taps.  conv4d_tc_kernel loops over them per OUTPUT plane: N = 16 (or 32 with both branches side by side) columns
// per MMA -- an M128xN32xK16 MMA occupies the tensor pipe for 40 cycles to do 16 cycles of arithmetic -- and every
// tap-packed plane is staged k times.  Here a CTA owns (batch item, jA, chunk of 128 positions) and steps through the
// INPUT planes ii: a staged plane contributes to the k outputs i = ii-p .. ii+p at once, so the d1 taps are folded
// into N (k*16 = 80 columns, or 160 with both branches: 80 cycles for 10x the columns), every plane is staged once,
// and the weights of all taps (26 / 51 KB) stay resident in shared memory.
//
// Accumulators: output i of work item t lives in slot (n mod R) of a ring over all 512 TMEM columns, n the running
// output count, R = 512 / (16*OG) slots.  One MMA writes the k consecutive slots of its live outputs (two MMAs when the
// run wraps around the ring).  Every MMA accumulates: a slot is zeroed (tcgen05.st) by the epilogue warps right after
// they have read it, and all of TMEM once at start.  Output i is complete after plane ii = i+p: the issuer commits to
// the slot's barrier and goes on; the epilogue warps turn the slot into CB16 (+ position-major twin) rows straight
// from registers: bias, ReLU / ReLU mask, fp16, 16-byte stores with lanes along the contiguous position axis.
#include "conv4d_tc.cuh"
#include "conv4d_common.cuh"
#include "tc_common.cuh"

namespace ncnet {
using namespace tc;

constexpr int kW1Threads = 320;      // TMA producer, MMA issuer, 8 epilogue warps
constexpr int kW1StagesMax = 8;
constexpr int kW1SlotsMax = 32;

struct Walk1Args {
    const __half *act;      // tap-packed CB16 input: NB channel blocks per (batch item, input group)
    const __half *wp;       // [Gp][NB][k t3][2][NTOT][8] fp16, NTOT rows ordered (u = k-1-t1, branch in merge, out channel)
    const float *bias;      // [Cout] or null
    const __half *mask;     // CB16 tensor shaped like the output or null
    __half *out;            // CB16 output, G groups per batch item
    __half *out_pm;         // position-major twin or null
    int B, G, Gp, OG, Gx, Cout;         // G = Gp * OG output groups; OG branches share one input (side by side in N)
    int d1, d2, d3, d4, P;
    int NB;                             // channel blocks of the (d2,d4) taps
    int nchunks, M_out;
    int items, cpg;                     // work items per input group, CTAs per input group
    int npos_pad, stages;
    int relu;
    uint32_t idesc_n[6];                // instruction descriptor for N = cnt * 16 * OG, cnt = 1..k
    uint32_t stage_bytes, w_bytes, w_off;
};

struct W1Item { int b, j, chunk; };
__device__ __forceinline__ W1Item w1_item(const Walk1Args &p, int t) {
    W1Item it;
    it.chunk = t % p.nchunks; t /= p.nchunks;
    it.j = t % p.d2;
    it.b = t / p.d2;
    return it;
}

__device__ __forceinline__ void tmem_st16_zero(uint32_t taddr) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1};"
        ::"r"(taddr), "r"(0u) : "memory");
}

template <int K, int OG>
__global__ void __launch_bounds__(kW1Threads, 1)
conv4d_tc_walk1_kernel(const Walk1Args p) {
    constexpr int PAD = K / 2, NCOL = 16 * OG, NTOT = K * NCOL, R = 512 / NCOL;
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t full_bar[kW1StagesMax], empty_bar[kW1StagesMax], w_bar;
    __shared__ __align__(8) uint64_t acc_full[kW1SlotsMax], acc_empty[kW1SlotsMax];
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int gp = blockIdx.x / p.cpg, cl = blockIdx.x % p.cpg;
    const int rows_pad = p.d3 + 2 * PAD;
    const int plane_cb = rows_pad * p.P;
    const size_t ij_count = (size_t)p.d1 * p.d2;
    const int npos_need = 128 + (K - 1) * p.P;

    if (warp == 1 && lane == 0) {
        for (int s = 0; s < p.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int s = 0; s < R; ++s) { mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], OG); }      // OG == 2: both warp sets read the slot
        mbar_init(&w_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    {   // stage rows behind a clipped copy feed discarded accumulator rows only; start them out finite
        const uint4 z = make_uint4(0, 0, 0, 0);
        const int n16 = (int)(p.w_off >> 4);
        for (int e = threadIdx.x; e < n16; e += kW1Threads) reinterpret_cast<uint4 *>(smem)[e] = z;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;
    if (warp >= 2 && warp < 6) {        // every accumulator column starts at zero: all MMAs accumulate
        const uint32_t lane_base = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
        for (int c = 0; c < 512; c += 16) tmem_st16_zero(lane_base + (uint32_t)c);
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t sm0 = smem_u32_pinned(smem);
    const uint32_t full0 = smem_u32_pinned(full_bar), empty0 = smem_u32_pinned(empty_bar);
    const uint32_t accf0 = smem_u32_pinned(acc_full), acce0 = smem_u32_pinned(acc_empty);

    if (warp == 0) {
        // ===== producer: the group's weights once, then this chunk of every tap-packed plane (ii, j), block by block =====
        if (elect_one()) {
            const uint32_t wb = smem_u32(&w_bar);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(wb), "r"(p.w_bytes) : "memory");
            bulk_g2s_u32(sm0 + p.w_off, (const uint8_t *)p.wp + (size_t)gp * p.w_bytes, p.w_bytes, wb);
            const size_t plane_halves = (size_t)plane_cb * 8;
            const size_t half_off = ij_count * plane_halves;            // between the two channel halves of a block
            const uint32_t half1_off = (uint32_t)p.npos_pad * 16u;
            int stage = 0;
            uint32_t phase = 0;
            for (int t = cl; t < p.items; t += p.cpg) {
                const W1Item it = w1_item(p, t);
                const int m0 = it.chunk * 128;
                const uint32_t bytes = (uint32_t)min(npos_need, plane_cb - m0) * 16u;
                const size_t bg = (size_t)it.b * p.Gx + (p.Gx == 1 ? 0 : gp);
                const __half *src_base = p.act + bg * p.NB * 2 * half_off + (size_t)it.j * plane_halves + (size_t)m0 * 8;
                for (int ii = 0; ii < p.d1; ++ii)
                    for (int blk = 0; blk < p.NB; ++blk) {
                        mbar_wait_u32(empty0 + (uint32_t)stage * 8u, phase ^ 1);
                        const uint32_t sb = sm0 + (uint32_t)stage * p.stage_bytes, fb = full0 + (uint32_t)stage * 8u;
                        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fb), "r"(2u * bytes) : "memory");
                        const __half *src0 = src_base + (size_t)blk * 2 * half_off + (size_t)ii * p.d2 * plane_halves;
                        bulk_g2s_u32(sb, src0, bytes, fb);
                        bulk_g2s_u32(sb + half1_off, src0 + half_off, bytes, fb);
                        if (++stage == p.stages) { stage = 0; phase ^= 1; }
                    }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (elect_one()) {
            mbar_wait(&w_bar, 0);
            const uint32_t lbo_a = (uint32_t)p.npos_pad << 16, lbo_b = (uint32_t)NTOT << 16;
            const uint64_t hi = (uint64_t)((128u >> 4) | (1u << 14)) << 32;             // SBO = 128 B, descriptor version 1
            const uint32_t sm4 = sm0 >> 4, stage4 = p.stage_bytes >> 4, w4 = (sm0 + p.w_off) >> 4;
            uint32_t a_rel[K];
#pragma unroll
            for (int t3 = 0; t3 < K; ++t3) a_rel[t3] = lbo_a | (uint32_t)(t3 * p.P);
            int stage = 0;
            uint32_t phase = 0;
            int n_base = 0;
            for (int t = cl; t < p.items; t += p.cpg) {
                for (int ii = 0; ii < p.d1; ++ii) {
                    const int u_lo = max(0, PAD - ii), u_hi = min(K - 1, p.d1 - 1 - ii + PAD);
                    const int n_lo = n_base + ii - PAD + u_lo;
                    // outputs touched for the first time by this plane: their slot's previous output must have been drained
                    for (int n = (ii == 0 ? n_lo : n_lo + (u_hi - u_lo)); n <= n_lo + (u_hi - u_lo); ++n) {
                        if ((ii == 0 || u_hi == K - 1) && n >= R)
                            mbar_wait_u32(acce0 + (uint32_t)(n % R) * 8u, (uint32_t)((n / R) - 1) & 1u);
                    }
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const int cnt = u_hi - u_lo + 1, s_lo = n_lo % R;
                    const int c1 = min(cnt, R - s_lo), c2 = cnt - c1;       // the run of live slots may wrap around the ring
                    const uint32_t d1col = tmem_base + (uint32_t)(s_lo * NCOL);
                    const uint32_t id1 = p.idesc_n[c1], id2 = p.idesc_n[c2];
                    for (int blk = 0; blk < p.NB; ++blk) {
                        mbar_wait_u32(full0 + (uint32_t)stage * 8u, phase);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        const uint32_t s4 = sm4 + (uint32_t)stage * stage4;
                        const uint32_t b0 = lbo_b | (w4 + (uint32_t)(blk * K * 2 * NTOT + u_lo * NCOL));
#pragma unroll
                        for (int t3 = 0; t3 < K; ++t3) {
                            const uint64_t adesc = hi | (uint64_t)(a_rel[t3] + s4);
                            const uint32_t b3 = b0 + (uint32_t)(t3 * 2 * NTOT);
                            umma_f16(d1col, adesc, hi | (uint64_t)b3, id1, 1u);
                            if (c2) umma_f16(tmem_base, adesc, hi | (uint64_t)(b3 + (uint32_t)(c1 * NCOL)), id2, 1u);
                        }
                        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                                     ::"r"(empty0 + (uint32_t)stage * 8u) : "memory");
                        if (++stage == p.stages) { stage = 0; phase ^= 1; }
                    }
                    // outputs whose last contributing plane this was
                    const int i_first = ii == p.d1 - 1 ? max(0, ii - PAD) : ii - PAD, i_last = ii == p.d1 - 1 ? p.d1 - 1 : ii - PAD;
                    for (int i = max(0, i_first); i <= i_last; ++i)
                        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                                     ::"r"(accf0 + (uint32_t)((n_base + i) % R) * 8u) : "memory");
                }
                n_base += p.d1;
            }
        }
        __syncwarp();
    } else {
        // ===== epilogue: 8 warps (two per TMEM lane quadrant warp%4), one output chunk per completed plane =====
        // OG == 2: warp set s (warps 2..5 / 6..9) owns branch s of every output; OG == 1: set s owns the outputs with n % 2 == s.
        // Every address is a per-item base plus a constant stride per output plane: the loop body is TMEM load, 16 x
        // (bias, ReLU / mask, fp16), two to four 16-byte stores.
        const int ew = warp - 2, set = ew >> 2, q = warp & 3;
        const int et = (ew & 3) * 32 + lane;             // 0..127 within the set
        const int og = OG == 2 ? set : 0;
        const int off_h = PAD * p.P + PAD;               // haloed position of flattened output m: m + off_h
        float bv[16];
#pragma unroll
        for (int o = 0; o < 16; ++o) bv[o] = (p.bias && o < p.Cout) ? __ldg(p.bias + o) : 0.f;     // the branches share the layer's bias
        const uint4 z4 = make_uint4(0, 0, 0, 0);
        const size_t half_halves = ij_count * (size_t)plane_cb * 8;          // between the channel halves of a CB16 block
        const size_t step_cb = (size_t)p.d2 * plane_cb * 8, step_pm = (size_t)p.d2 * plane_cb * 16;     // per output plane i
        const bool full_ch = p.Cout == 16;
        int n = 0;
        for (int t = cl; t < p.items; t += p.cpg) {
            const W1Item it = w1_item(p, t);
            const int m = it.chunk * 128 + q * 32 + lane;            // this lane's flattened output position
            const int hp = m + off_h;
            const bool in_plane = m < p.M_out;
            const bool valid = in_plane && (m % p.P) < p.d4;
            const size_t blk_idx = (size_t)it.b * p.G + gp * OG + og;
            const size_t cb0 = ((blk_idx * 2) * ij_count + (size_t)it.j) * (size_t)plane_cb * 8;       // plane (i = 0, j), half 0
            const size_t pm0 = (blk_idx * ij_count + (size_t)it.j) * (size_t)plane_cb * 16;
            const int swap8 = ((hp >> 2) & 1) * 8;
            for (int i = 0; i < p.d1; ++i, ++n) {
                if (OG == 1 && (n & 1) != set) continue;             // warp-uniform
                const int slot = n % R;
                __half *ob = p.out + cb0 + (size_t)i * step_cb;
                __half *opm = p.out_pm ? p.out_pm + pm0 + (size_t)i * step_pm : nullptr;
                // the ReLU mask of this position (dgrad) is fetched while the plane's last MMAs are still running
                uint4 mk0 = z4, mk1 = z4;
                if (p.mask && valid) {
                    const __half *mp = p.mask + cb0 + (size_t)i * step_cb + (size_t)hp * 8;
                    mk0 = *reinterpret_cast<const uint4 *>(mp);
                    mk1 = *reinterpret_cast<const uint4 *>(mp + half_halves);
                }
                mbar_wait_sleep(&acc_full[slot], (uint32_t)(n / R) & 1u, 50);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(slot * NCOL + og * 16);
                uint32_t r[16];
                tmem_ld16_issue(taddr, r);
                tmem_ld16_wait(r);
                tmem_st16_zero(taddr);                               // ready for the slot's next output
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                asm volatile("bar.sync %0, 128;" ::"r"(1 + set) : "memory");
                if (et == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(acce0 + (uint32_t)slot * 8u) : "memory");
                if (in_plane) {
                    uint4 h0 = z4, h1 = z4;
                    if (valid) {
                        float x[16];
#pragma unroll
                        for (int o = 0; o < 16; ++o) {
                            x[o] = __uint_as_float(r[o]) + bv[o];
                            if (p.relu) x[o] = fmaxf(x[o], 0.f);
                            if (!full_ch && o >= p.Cout) x[o] = 0.f;
                        }
                        if (p.mask) {
                            const __half *mh0 = reinterpret_cast<const __half *>(&mk0), *mh1 = reinterpret_cast<const __half *>(&mk1);
#pragma unroll
                            for (int c8 = 0; c8 < 8; ++c8) {
                                if (!(__half2float(mh0[c8]) > 0.f)) x[c8] = 0.f;
                                if (!(__half2float(mh1[c8]) > 0.f)) x[8 + c8] = 0.f;
                            }
                        }
                        __half2 *p0 = reinterpret_cast<__half2 *>(&h0), *p1 = reinterpret_cast<__half2 *>(&h1);
#pragma unroll
                        for (int c2 = 0; c2 < 4; ++c2) {
                            p0[c2] = __floats2half2_rn(x[2 * c2], x[2 * c2 + 1]);
                            p1[c2] = __floats2half2_rn(x[8 + 2 * c2], x[8 + 2 * c2 + 1]);
                        }
                    }
                    *reinterpret_cast<uint4 *>(ob + (size_t)hp * 8) = h0;
                    *reinterpret_cast<uint4 *>(ob + half_halves + (size_t)hp * 8) = h1;
                    if (opm) {       // position-major twin: 32 B per position, halves swapped by bit 7 of the in-plane byte offset
                        *reinterpret_cast<uint4 *>(opm + (size_t)hp * 16 + swap8) = h0;
                        *reinterpret_cast<uint4 *>(opm + (size_t)hp * 16 + (8 - swap8)) = h1;
                    }
                }
                // the halo rows above the first / below the last output position
                if (it.chunk == 0) {
                    for (int e = et; e < 2 * off_h; e += 128) {
                        const int half = e >= off_h ? 1 : 0, pos = e - half * off_h;
                        *reinterpret_cast<uint4 *>(ob + (size_t)half * half_halves + (size_t)pos * 8) = z4;
                        if (opm) *reinterpret_cast<uint4 *>(opm + (size_t)pos * 16 + (size_t)half * 8) = z4;
                    }
                }
                if (it.chunk == p.nchunks - 1) {
                    const int lo = p.M_out + off_h, cnt = plane_cb - lo;
                    for (int e = et; e < 2 * cnt; e += 128) {
                        const int half = e >= cnt ? 1 : 0, pos = lo + (e - half * cnt);
                        *reinterpret_cast<uint4 *>(ob + (size_t)half * half_halves + (size_t)pos * 8) = z4;
                        if (opm) *reinterpret_cast<uint4 *>(opm + (size_t)pos * 16 + (size_t)half * 8) = z4;
                    }
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
}

// -> [Gp][NB][k t3][2 halves][NTOT = k*OG*16 rows (u, branch, o)][8 channels]; channel c = blk*16 + half*8 + c8 is the
// (d2,d4) tap (c / k, c % k) as the tap-packed activations order them; row u holds the d1 tap k-1-u
__global__ void pack_w_walk1_kernel(const float *__restrict__ w, __half *__restrict__ wp, PackDesc d, int OG, int NB, int Gp) {
    const int k = d.k, NTOT = k * OG * 16;
    const long long total = (long long)Gp * NB * k * 2 * NTOT * 8;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        long long r = idx;
        const int c8 = (int)(r % 8); r /= 8;
        const int row = (int)(r % NTOT); r /= NTOT;
        const int half = (int)(r % 2); r /= 2;
        const int t3 = (int)(r % k); r /= k;
        const int blk = (int)(r % NB); r /= NB;
        const int gp = (int)r;
        const int u = row / (OG * 16), og = (row / 16) % OG, o = row % 16;
        const int c = blk * 16 + half * 8 + c8;
        float v = 0.f;
        if (c < k * k && o < d.Cout_e) v = w[packed_source_index(d, gp * OG + og, o, 0, k - 1 - u, c / k, t3, c % k)];
        wp[idx] = __float2half_rn(v);
    }
}

struct Walk1Plan {
    bool ok;
    int NB, nchunks, M_out, npos_pad, stages;
    uint32_t stage_bytes, w_bytes, w_off;
    size_t smem;
};

static Walk1Plan walk1_plan(int k, int og, int d3, int d4) {
    Walk1Plan w{};
    w.ok = false;
    if ((k != 3 && k != 5) || (og != 1 && og != 2)) return w;
    const int P = d4 + k - 1;
    w.NB = (k * k + 15) / 16;
    w.M_out = (d3 - 1) * P + d4;
    w.nchunks = (w.M_out + 127) / 128;
    int npos = 128 + (k - 1) * P;
    npos = (npos + 7) / 8 * 8;
    if (npos > 16383) return w;
    w.npos_pad = npos;
    w.stage_bytes = (uint32_t)(2 * npos * 16);
    w.w_bytes = (uint32_t)(w.NB * k * 2 * (k * og * 16) * 16);
    const uint32_t budget = 200u * 1024u;
    if (w.w_bytes + 3u * w.stage_bytes > budget) return w;
    const int stages = (int)((budget - w.w_bytes) / w.stage_bytes);
    w.stages = stages > kW1StagesMax ? kW1StagesMax : stages;
    w.w_off = (uint32_t)w.stages * w.stage_bytes;
    w.smem = (size_t)w.w_off + w.w_bytes;
    if (w.smem < 120u * 1024u) w.smem = 120u * 1024u;          // one CTA per SM: each takes all 512 TMEM columns
    w.ok = true;
    return w;
}

bool conv4d_tc_walk1_ok(int k, int og, int d3, int d4) { return walk1_plan(k, og, d3, d4).ok; }

template <int K, int OG>
static int walk1_launch(const Walk1Args &a, const Walk1Plan &w, cudaStream_t st) {
    cudaError_t e = cudaFuncSetAttribute(conv4d_tc_walk1_kernel<K, OG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)w.smem);
    NC_CHECK(e == cudaSuccess, "conv4d(tc,walk1): smem attribute: %s", cudaGetErrorString(e));
    conv4d_tc_walk1_kernel<K, OG><<<(unsigned)(a.Gp * a.cpg), kW1Threads, w.smem, st>>>(a);
    NC_LAUNCH_OK("conv4d_tc_walk1_kernel");
    return 0;
}

// act: tap-packed CB16 input (cb16_pack tap_mode 4); w: the layer's weights as the reference stores them; wp: scratch for
// the re-laid-out weights (G * NB * k * 2 * k*16 * 16 bytes); out (+ mask): CB16 with G groups per batch item
int conv4d_tc_walk1_run(const __half *act, const float *w, __half *wp, const PackDesc &d, const float *bias, const __half *mask,
                        __half *out, __half *out_pm, int B, int og, int Gx, int d1, int d2, int d3, int d4, int relu,
                        cudaStream_t st) {
    const int k = d.k;
    const Walk1Plan pl = walk1_plan(k, og, d3, d4);
    if (!pl.ok) return 2;
    Walk1Args a{};
    a.act = act; a.wp = wp; a.bias = bias; a.mask = mask; a.out = out; a.out_pm = out_pm;
    a.B = B; a.G = d.G_e; a.OG = og; a.Gp = d.G_e / og; a.Gx = Gx; a.Cout = d.Cout_e;
    a.d1 = d1; a.d2 = d2; a.d3 = d3; a.d4 = d4; a.P = d4 + k - 1;
    a.NB = pl.NB; a.nchunks = pl.nchunks; a.M_out = pl.M_out;
    a.npos_pad = pl.npos_pad; a.stages = pl.stages; a.relu = relu;
    a.stage_bytes = pl.stage_bytes; a.w_bytes = pl.w_bytes; a.w_off = pl.w_off;
    for (int c = 1; c <= k; ++c)      // D=f32, A=B=f16, both K-major, N>>3 at bit 17, M>>4 at bit 24
        a.idesc_n[c] = (1u << 4) | ((uint32_t)((c * 16 * og) >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const long long items = (long long)B * d2 * pl.nchunks;
    NC_CHECK(items < (1ll << 31), "conv4d(tc,walk1): too many work items");
    a.items = (int)items;
    a.cpg = kNumSMs / a.Gp > 0 ? kNumSMs / a.Gp : 1;
    if (a.cpg > a.items) a.cpg = a.items;
    {
        const long long total = (long long)a.Gp * pl.NB * k * 2 * (k * og * 16) * 8;
        const int blocks = (int)((total + 255) / 256 < 1184 ? (total + 255) / 256 : 1184);
        pack_w_walk1_kernel<<<blocks, 256, 0, st>>>(w, wp, d, og, pl.NB, a.Gp);
        NC_LAUNCH_OK("pack_w_walk1_kernel");
    }
    if (k == 5) return og == 2 ? walk1_launch<5, 2>(a, pl, st) : walk1_launch<5, 1>(a, pl, st);
    return og == 2 ? walk1_launch<3, 2>(a, pl, st) : walk1_launch<3, 1>(a, pl, st);
}

}  // namespace ncnet
