// Many-channel Conv4d layers (mode M1: Cin <= 16, Cout > 1) as a persistent WALK along jA on CTA pairs, with the
// layer's weights resident in shared memory (reference semantics: lib/conv4d.py:11-51).
//
// Why: in conv4d_tc_kernel a CTA owns one output plane and streams its k*k input planes AND their k*k weight slabs
// (12.8 KB per (d1,d2) tap at 16->16, k=5) from L2; ncu shows that kernel bound by the L2->SM crossbar (39 GB per launch
// of 32 PF-Pascal volumes, one N=80 MMA per ~69 cycles where the tensor pipe could retire one per 52).  Here
//   * a CTA PAIR (cluster of 2, tcgen05 cta_group::2, one M256 MMA stream) holds ALL k*k tap slabs of one branch group
//     for its whole life -- 320 KB of fp16 at 16->16, k=5: half of the N columns (160 KB) in each CTA's shared memory;
//   * the pair is persistent: it walks a list of work items (batch item, iA, pair of position chunks, jA segment);
//   * inside an item it steps through the input columns jj; a staged chunk of plane (iA+t1-p, jj) is used at once for
//     the k outputs j = jj-p .. jj+p it contributes to, whose accumulators are live slots of TMEM (k+1 slots of N
//     columns: the extra slot lets an output drain for a whole column step before its slot is needed again);
// so per output chunk the CTA stages k*k small activation chunks (8 KB) and nothing else: the tensor pipe is the bound.
//
// GEMM view (as in conv4d_tc.cu): positions of the haloed (iB,jB) plane flattened with pitch P; output m = r*P + c
// (valid when c < d4); A = 128 consecutive positions x 16 channels, K-major, a d3 tap is a row offset of P positions;
// N = (d4 tap, out-channel) = k*16 columns; D[m'][t4*16+o] and out[m] = sum_t4 D[m - m0 + t4][t4*16+o], so a chunk of
// 128 D rows yields OUT = 128-(k-1) outputs.  Chunks tile the flattened output index, not rows.
#include "conv4d_tc.cuh"
#include "conv4d_common.cuh"
#include "tc_common.cuh"

namespace ncnet {
using namespace tc;

constexpr int kW16Threads = 192;
constexpr int kW16StagesMax = 8;
constexpr int kW16SlotsMax = 6;
constexpr int kW16OutStride = 20;      // floats per row of the shifted-sum buffer (16 channels + pad)

struct Walk16Args {
    const __half *act;      // CB16 input
    const __half *wp;       // [G][pair rank][k*k][k K-steps][2][N/2][8] fp16
    const float *bias;
    const __half *mask;     // CB16 tensor shaped like the output (ReLU mask of a dgrad) or null
    __half *out;            // CB16 output
    __half *out_pm;         // its position-major twin (NCNET_CB16PM) or null
    int B, G, Gx, Cout;
    int d1, d2, d3, d4, P;
    int OUT, M_out, npairs;             // outputs per chunk, flattened outputs per plane, chunk pairs per plane
    int nseg, seg_len;                  // jA segments per (b, iA, chunk pair)
    int items, cpg;                     // work items per branch group, clusters per branch group
    int npos_pad, stages;
    int relu;
    uint32_t idesc, stage_bytes, w_bytes, w_off, out_off;
};

struct W16Item { int b, i, cp, j0, j1; };
__device__ __forceinline__ W16Item w16_item(const Walk16Args &p, int t) {
    W16Item it;
    it.i = t % p.d1; t /= p.d1;         // iA fastest: clusters that run together share (t1-shifted) input planes in L2
    it.cp = t % p.npairs; t /= p.npairs;
    const int seg = t % p.nseg;
    it.b = t / p.nseg;
    it.j0 = seg * p.seg_len;
    it.j1 = min(p.d2, it.j0 + p.seg_len);
    return it;
}

__device__ __forceinline__ void umma_f16_pair(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
// MODE: A-collector hint (0 none, 1 fill, 2 use, 3 lastuse) for MMAs issued back to back on the same A tile; ACC1: the
// accumulate flag is the constant 1 (no predicate set-up in the single issuing lane)
template <int MODE, bool ACC1>
__device__ __forceinline__ void pair_mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
#define NC_PAIR_MMA(Q)                                                                                                                      \
    do {                                                                                                                                    \
        if (ACC1)                                                                                                                           \
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::2.kind::f16" Q " [%0], %1, %2, %3, p;\n\t}"  \
                         ::"r"(d), "l"(a), "l"(b), "r"(idesc) : "memory");                                                                  \
        else                                                                                                                                \
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16" Q " [%0], %1, %2, %3, p;\n\t}" \
                         ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");                                                        \
    } while (0)
    if (MODE == 1) NC_PAIR_MMA(".collector::a::fill");
    else if (MODE == 2) NC_PAIR_MMA(".collector::a::use");
    else if (MODE == 3) NC_PAIR_MMA(".collector::a::lastuse");
    else NC_PAIR_MMA("");
#undef NC_PAIR_MMA
}
// arrives on the barrier at this shared-memory offset in BOTH CTAs of the pair once the MMAs issued so far have completed
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ uint32_t leader_addr(const void *p) {      // the same variable in the pair's rank-0 CTA
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_u32(p)), "r"(0));
    return r;
}
__device__ __forceinline__ void arrive_cluster(uint32_t addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(addr) : "memory");
}

template <int K>
__global__ void __launch_bounds__(kW16Threads, 1)
conv4d_tc_walk16_kernel(const Walk16Args p) {
    constexpr int PAD = K / 2, SL = K + 1, N = K * 16, NLOC = N / 2;
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t full_bar[kW16StagesMax], empty_bar[kW16StagesMax], peer_full[kW16StagesMax];
    __shared__ __align__(8) uint64_t acc_full[kW16SlotsMax], acc_empty[kW16SlotsMax], w_bar, peer_w;
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t cta_rank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta_rank));
    const int cluster = blockIdx.x >> 1;
    const int g = cluster / p.cpg, cl = cluster % p.cpg;
    const int rows_pad = p.d3 + 2 * PAD;
    const int plane_cb = rows_pad * p.P;                  // positions of a haloed plane
    const size_t ij_count = (size_t)p.d1 * p.d2;
    const int npos_need = 128 + (K - 1) * p.P;            // positions a stage's MMAs read

    if (warp == 1 && lane == 0) {
        for (int s = 0; s < p.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); mbar_init(&peer_full[s], 1); }
        for (int s = 0; s < SL; ++s) { mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], 2); }    // both CTAs' epilogues
        mbar_init(&w_bar, 1);
        mbar_init(&peer_w, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    {   // stage rows behind a clipped copy are read by discarded accumulator rows only; start them out finite
        const uint4 z = make_uint4(0, 0, 0, 0);
        const int n16 = (int)(p.w_off >> 4);
        for (int e = threadIdx.x; e < n16; e += kW16Threads) reinterpret_cast<uint4 *>(smem)[e] = z;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");       // the peer's barriers are initialised
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;
    const uint32_t sm0 = smem_u32_pinned(smem);
    const uint32_t full0 = smem_u32_pinned(full_bar), empty0 = smem_u32_pinned(empty_bar);
    const uint32_t accf0 = smem_u32_pinned(acc_full);

    if (warp == 0) {
        // ===== producer: this CTA's half of every weight slab once, then its activation chunk per (jj, t1) =====
        if (elect_one()) {
            const uint32_t wb = smem_u32(&w_bar);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(wb), "r"(p.w_bytes) : "memory");
            const uint8_t *wsrc = (const uint8_t *)p.wp + ((size_t)g * 2 + cta_rank) * p.w_bytes;
            const uint32_t part = p.w_bytes / K;
            for (int t1 = 0; t1 < K; ++t1) bulk_g2s_u32(sm0 + p.w_off + (uint32_t)t1 * part, wsrc + (size_t)t1 * part, part, wb);
            const size_t plane_halves = (size_t)plane_cb * 8;
            const size_t half_off = ij_count * plane_halves;
            const uint32_t half1_off = (uint32_t)p.npos_pad * 16u;
            int stage = 0;
            uint32_t phase = 0;
            for (int t = cl; t < p.items; t += p.cpg) {
                const W16Item it = w16_item(p, t);
                const int m0 = (it.cp * 2 + (int)cta_rank) * p.OUT;
                const int nvalid = max(0, min(npos_need, plane_cb - m0));
                const uint32_t bytes = (uint32_t)nvalid * 16u;
                const size_t bg = (size_t)it.b * p.Gx + (p.Gx == 1 ? 0 : g);
                const __half *src_base = p.act + bg * 2 * ij_count * plane_halves + (size_t)m0 * 8;
                const int t1_lo = max(0, PAD - it.i), t1_hi = min(K - 1, p.d1 - 1 - it.i + PAD);
                const int jj_lo = max(0, it.j0 - PAD), jj_hi = min(p.d2 - 1, it.j1 - 1 + PAD);
                for (int jj = jj_lo; jj <= jj_hi; ++jj)
                    for (int t1 = t1_lo; t1 <= t1_hi; ++t1) {
                        mbar_wait_u32(empty0 + (uint32_t)stage * 8u, phase ^ 1);
                        const uint32_t sb = sm0 + (uint32_t)stage * p.stage_bytes, fb = full0 + (uint32_t)stage * 8u;
                        if (bytes) {
                            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fb), "r"(2u * bytes) : "memory");
                            const __half *src0 = src_base + (size_t)((it.i + t1 - PAD) * p.d2 + jj) * plane_halves;
                            bulk_g2s_u32(sb, src0, bytes, fb);
                            bulk_g2s_u32(sb + half1_off, src0 + half_off, bytes, fb);
                        } else {      // a chunk past the end of the plane (odd chunk count padded to a pair): nothing to load
                            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(fb) : "memory");
                        }
                        if (++stage == p.stages) { stage = 0; phase ^= 1; }
                    }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (elect_one()) {
            if (cta_rank == 0) {
                // ===== MMA issuer (leader CTA): one M256 x N x K16 stream for the pair =====
                mbar_wait(&w_bar, 0);
                mbar_wait(&peer_w, 0);
                const uint32_t peer0 = smem_u32_pinned(peer_full), acce0 = smem_u32_pinned(acc_empty);
                const uint32_t lbo_a = (uint32_t)p.npos_pad << 16, lbo_b = (uint32_t)NLOC << 16;
                const uint64_t hi = (uint64_t)((128u >> 4) | (1u << 14)) << 32;         // SBO = 128 B, descriptor version 1
                const uint32_t sm4 = sm0 >> 4, stage4 = p.stage_bytes >> 4, w4 = (sm0 + p.w_off) >> 4;
                constexpr uint32_t SLAB4 = (uint32_t)(K * 2 * NLOC), BSTEP = (uint32_t)(2 * NLOC);   // 16-byte units
                uint32_t a_rel[K];
#pragma unroll
                for (int ks = 0; ks < K; ++ks) a_rel[ks] = lbo_a | (uint32_t)(ks * p.P);
                int stage = 0;
                uint32_t phase = 0;
                int n_base = 0;                     // outputs of earlier items (slot = n mod SL)
                bool prewaited = false;             // the current stage's barriers were already waited for during the previous stage
                for (int t = cl; t < p.items; t += p.cpg) {
                    const W16Item it = w16_item(p, t);
                    const int t1_lo = max(0, PAD - it.i), t1_hi = min(K - 1, p.d1 - 1 - it.i + PAD);
                    const int jj_lo = max(0, it.j0 - PAD), jj_hi = min(p.d2 - 1, it.j1 - 1 + PAD);
                    const bool last_item = t + p.cpg >= p.items;
                    for (int jj = jj_lo; jj <= jj_hi; ++jj) {
                        // interior column: all k outputs j = jj-PAD+dd are live and only the last can be new
                        const bool interior = jj - PAD >= it.j0 && jj + PAD < it.j1;
                        const int n0 = n_base + (jj - PAD - it.j0);
                        uint32_t dcol[K];
                        {
                            int slot = interior ? n0 % SL : 0;
#pragma unroll
                            for (int dd = 0; dd < K; ++dd) { dcol[dd] = tmem_base + (uint32_t)(slot * N); if (++slot == SL) slot = 0; }
                        }
                        for (int t1 = t1_lo; t1 <= t1_hi; ++t1) {
                            if (!prewaited) {
                                mbar_wait_u32(full0 + (uint32_t)stage * 8u, phase);
                                mbar_wait_u32(peer0 + (uint32_t)stage * 8u, phase);
                            }
                            prewaited = false;
                            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                            const uint32_t s4 = sm4 + (uint32_t)stage * stage4;
                            const uint32_t wrow = lbo_b | (w4 + (uint32_t)(t1 * K + (K - 1)) * SLAB4);   // tap t2 = K-1-dd
                            if (interior) {
                                // The stage's K-step tile A is shared by the k outputs: K-step outer, outputs inner, A held in the
                                // collector (fill, use ..., lastuse: 41 instead of 46 cycles per M256 x N80 MMA); straight-line issue.
                                const bool nw = t1 == t1_lo;
                                if (nw && n0 + K - 1 >= SL) {      // the new output's slot: its previous owner has been drained by both CTAs
                                    const int n = n0 + K - 1;
                                    mbar_wait_u32(acce0 + (uint32_t)(n % SL) * 8u, (uint32_t)((n / SL) - 1) & 1u);
                                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                                }
#pragma unroll
                                for (int ks = 0; ks < K; ++ks) {
                                    if (ks == K - 1 && !(last_item && jj == jj_hi && t1 == t1_hi)) {
                                        // the next stage's barriers are polled while the MMAs issued so far keep the tensor pipe busy
                                        const int ns = stage + 1 == p.stages ? 0 : stage + 1;
                                        const uint32_t np = stage + 1 == p.stages ? phase ^ 1u : phase;
                                        mbar_wait_u32(full0 + (uint32_t)ns * 8u, np);
                                        mbar_wait_u32(peer0 + (uint32_t)ns * 8u, np);
                                        prewaited = true;
                                    }
                                    const uint64_t adesc = hi | (uint64_t)(a_rel[ks] + s4);
                                    const uint32_t b0 = wrow + (uint32_t)ks * BSTEP;
                                    pair_mma<1, true>(dcol[0], adesc, hi | (uint64_t)b0, p.idesc, 1u);
#pragma unroll
                                    for (int dd = 1; dd < K - 1; ++dd)
                                        pair_mma<2, true>(dcol[dd], adesc, hi | (uint64_t)(b0 - (uint32_t)dd * SLAB4), p.idesc, 1u);
                                    if (ks == 0) pair_mma<3, false>(dcol[K - 1], adesc, hi | (uint64_t)(b0 - (uint32_t)(K - 1) * SLAB4), p.idesc, nw ? 0u : 1u);
                                    else pair_mma<3, true>(dcol[K - 1], adesc, hi | (uint64_t)(b0 - (uint32_t)(K - 1) * SLAB4), p.idesc, 1u);
                                }
                            } else {
#pragma unroll
                                for (int dd = 0; dd < K; ++dd) {        // outputs j = jj - PAD + dd, ascending: a new one comes last
                                    const int j = jj - PAD + dd;
                                    if (j >= it.j0 && j < it.j1) {
                                        const int n = n_base + (j - it.j0);
                                        const int slot = n % SL;
                                        const bool is_new = t1 == t1_lo && jj == max(j - PAD, 0);
                                        if (is_new && n >= SL) {        // the slot's previous output (n - SL) has been drained by both CTAs
                                            mbar_wait_u32(acce0 + (uint32_t)slot * 8u, (uint32_t)((n / SL) - 1) & 1u);
                                            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                                        }
                                        const uint32_t b0 = wrow - (uint32_t)dd * SLAB4;
                                        const uint32_t dcol = tmem_base + (uint32_t)(slot * N);
#pragma unroll
                                        for (int ks = 0; ks < K; ++ks)
                                            umma_f16_pair(dcol, hi | (uint64_t)(a_rel[ks] + s4), hi | (uint64_t)(b0 + (uint32_t)ks * BSTEP), p.idesc,
                                                          (is_new && ks == 0) ? 0u : 1u);
                                    }
                                }
                            }
                            umma_commit_pair(empty0 + (uint32_t)stage * 8u);       // frees the stage in both CTAs
                            if (++stage == p.stages) { stage = 0; phase ^= 1; }
                        }
                        // outputs whose last contributing column this was
#pragma unroll
                        for (int dd = 0; dd < K; ++dd) {
                            const int j = jj - PAD + dd;
                            if (j >= it.j0 && j < it.j1 && min(p.d2 - 1, j + PAD) == jj)
                                umma_commit_pair(accf0 + (uint32_t)((n_base + (j - it.j0)) % SL) * 8u);
                        }
                    }
                    n_base += it.j1 - it.j0;
                }
            } else {
                // ===== peer CTA: no MMAs of its own; tells the leader when its weights and each of its stages have landed =====
                const uint32_t r_peer_w = leader_addr(&peer_w), r_peer0 = leader_addr(peer_full);
                mbar_wait(&w_bar, 0);
                arrive_cluster(r_peer_w);
                int stage = 0;
                uint32_t phase = 0;
                for (int t = cl; t < p.items; t += p.cpg) {
                    const W16Item it = w16_item(p, t);
                    const int nt1 = min(K - 1, p.d1 - 1 - it.i + PAD) - max(0, PAD - it.i) + 1;
                    const int njj = min(p.d2 - 1, it.j1 - 1 + PAD) - max(0, it.j0 - PAD) + 1;
                    for (int s = 0; s < nt1 * njj; ++s) {
                        mbar_wait_u32(full0 + (uint32_t)stage * 8u, phase);
                        arrive_cluster(r_peer0 + (uint32_t)stage * 8u);
                        if (++stage == p.stages) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
        __syncwarp();
    } else {
        // ===== epilogue (warps 2..5 <-> TMEM lane quadrants warp%4): one output chunk per completed column =====
        const int q = warp & 3;
        const int et = threadIdx.x - 64;
        const int row = q * 32 + lane;
        const uint32_t osm = sm0 + p.out_off;
        const uint32_t r_acce0 = leader_addr(acc_empty);
        const int off_h = PAD * p.P + PAD;               // haloed position of output m: m + off_h
        float bv[16];
#pragma unroll
        for (int o = 0; o < 16; ++o) bv[o] = (p.bias && o < p.Cout) ? __ldg(p.bias + o) : 0.f;
        int n = 0;
        for (int t = cl; t < p.items; t += p.cpg) {
            const W16Item it = w16_item(p, t);
            const int chunk = it.cp * 2 + (int)cta_rank;
            const int m0 = chunk * p.OUT;
            const int end_m = min(m0 + p.OUT, p.M_out);
            for (int j = it.j0; j < it.j1; ++j, ++n) {
                const int slot = n % SL;
                mbar_wait_sleep(&acc_full[slot], (uint32_t)(n / SL) & 1u, 100);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(slot * N);
                uint32_t r[K][16];
#pragma unroll
                for (int gi = 0; gi < K; ++gi) tmem_ld16_issue(lane_base + (uint32_t)(gi * 16), r[gi]);
#pragma unroll
                for (int gi = 0; gi < K; ++gi) tmem_ld16_wait(r[gi]);
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                // column group gi = 16 out-channels of tap t4 = gi: out[row - gi][o] += D[row][gi*16 + o]; group 0 writes
                // bias + D for every row, the others accumulate (one barrier per group orders the read-modify-writes)
                {
                    const uint32_t dst = osm + (uint32_t)row * (kW16OutStride * 4u);
#pragma unroll
                    for (int w4 = 0; w4 < 4; ++w4)
                        sts_v4(dst + 16u * w4,
                               make_float4(bv[4 * w4 + 0] + __uint_as_float(r[0][4 * w4 + 0]), bv[4 * w4 + 1] + __uint_as_float(r[0][4 * w4 + 1]),
                                           bv[4 * w4 + 2] + __uint_as_float(r[0][4 * w4 + 2]), bv[4 * w4 + 3] + __uint_as_float(r[0][4 * w4 + 3])));
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");
                if (et == 0) arrive_cluster(r_acce0 + (uint32_t)slot * 8u);     // every warp's TMEM reads of the slot are done
#pragma unroll
                for (int gi = 1; gi < K; ++gi) {
                    const int dest = row - gi;
                    if (dest >= 0) {
                        const uint32_t dst = osm + (uint32_t)dest * (kW16OutStride * 4u);
                        float4 a[4];
#pragma unroll
                        for (int w4 = 0; w4 < 4; ++w4) a[w4] = lds_v4(dst + 16u * w4);
#pragma unroll
                        for (int w4 = 0; w4 < 4; ++w4) {
                            a[w4].x += __uint_as_float(r[gi][4 * w4 + 0]); a[w4].y += __uint_as_float(r[gi][4 * w4 + 1]);
                            a[w4].z += __uint_as_float(r[gi][4 * w4 + 2]); a[w4].w += __uint_as_float(r[gi][4 * w4 + 3]);
                            sts_v4(dst + 16u * w4, a[w4]);
                        }
                    }
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                }
                if (m0 < p.M_out) {
                    // CB16 store: the haloed positions this chunk owns (halo rows / columns written as zeros), 16 B per
                    // (position, channel half), lanes along the contiguous position axis
                    const int lo = chunk == 0 ? 0 : m0 + off_h;
                    const int hi_pos = end_m == p.M_out ? plane_cb : end_m + off_h;
                    const int npos = hi_pos - lo;
                    const size_t base_off = ((((size_t)it.b * p.G + g) * 2) * ij_count + (size_t)it.i * p.d2 + j) * (size_t)plane_cb * 8;
                    __half *ob = p.out + base_off;
                    const __half *mb = p.mask + base_off;
                    for (int e = et; e < 2 * npos; e += 128) {
                        const int half = e >= npos ? 1 : 0;
                        const int hp = lo + (e - half * npos);
                        const int m = hp - off_h;
                        const size_t goff = ((size_t)half * ij_count * plane_cb + (size_t)hp) * 8;
                        __align__(16) __half hv[8];
                        if (m >= m0 && m < end_m && (m % p.P) < p.d4) {
                            const uint32_t src = osm + (uint32_t)((m - m0) * kW16OutStride + half * 8) * 4u;
                            const float4 a = lds_v4(src), c = lds_v4(src + 16u);
                            float v[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
                            if (p.mask) {
                                const uint4 mraw = *reinterpret_cast<const uint4 *>(mb + goff);
                                const __half *mh = reinterpret_cast<const __half *>(&mraw);
#pragma unroll
                                for (int c8 = 0; c8 < 8; ++c8) if (!(__half2float(mh[c8]) > 0.f)) v[c8] = 0.f;
                            }
#pragma unroll
                            for (int c8 = 0; c8 < 8; ++c8) {
                                float x = v[c8];
                                if (p.relu) x = fmaxf(x, 0.f);
                                hv[c8] = __float2half_rn((half * 8 + c8) < p.Cout ? x : 0.f);
                            }
                        } else {
#pragma unroll
                            for (int c8 = 0; c8 < 8; ++c8) hv[c8] = __float2half_rn(0.f);
                        }
                        *reinterpret_cast<uint4 *>(ob + goff) = *reinterpret_cast<const uint4 *>(hv);
                        if (p.out_pm) {      // position-major twin: 32 B per position, halves swapped by bit 7 of the in-plane offset
                            const size_t pm_off = ((((size_t)it.b * p.G + g) * ij_count + (size_t)it.i * p.d2 + j) * plane_cb + hp) * 16 +
                                                  (size_t)((half ^ ((hp >> 2) & 1)) * 8);
                            *reinterpret_cast<uint4 *>(p.out_pm + pm_off) = *reinterpret_cast<const uint4 *>(hv);
                        }
                    }
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");        // the sum buffer is free for the next output
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    // neither CTA may leave (or free TMEM) while the other can still be the target of shared MMAs / remote arrives
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
struct Walk16Plan {
    bool ok;
    int OUT, M_out, npairs, npos_pad, stages;
    uint32_t stage_bytes, w_bytes, w_off, out_off;
    size_t smem;
};

static Walk16Plan walk16_plan(int k, int d3, int d4) {
    Walk16Plan w{};
    w.ok = false;
    if (k != 3 && k != 5) return w;
    const int P = d4 + k - 1, N = k * 16;
    w.OUT = 128 - (k - 1);
    w.M_out = (d3 - 1) * P + d4;
    w.npairs = ((w.M_out + w.OUT - 1) / w.OUT + 1) / 2;
    int npos = 128 + (k - 1) * P;
    npos = (npos + 7) / 8 * 8;
    if (npos > 16383) return w;
    w.npos_pad = npos;
    w.stage_bytes = (uint32_t)(2 * npos * 16);
    w.w_bytes = (uint32_t)(k * k * k * 2 * (N / 2) * 16);          // one CTA's half of every (t1,t2) slab
    const uint32_t out_bytes = 128u * kW16OutStride * 4u;
    const uint32_t budget = 225u * 1024u;
    const uint32_t fixed = (w.w_bytes + 127u) / 128u * 128u + out_bytes;
    if (fixed + 3u * w.stage_bytes > budget) return w;
    const int stages = (int)((budget - fixed) / w.stage_bytes);
    w.stages = stages > kW16StagesMax ? kW16StagesMax : stages;
    w.w_off = (uint32_t)w.stages * w.stage_bytes;                  // a multiple of 256 B
    w.out_off = w.w_off + (w.w_bytes + 127u) / 128u * 128u;
    w.smem = (size_t)w.out_off + out_bytes;
    if (w.smem < 120u * 1024u) w.smem = 120u * 1024u;              // one CTA per SM: each pair takes all 512 TMEM columns
    w.ok = true;
    return w;
}

bool conv4d_tc_walk16_ok(int k, int d3, int d4) { return walk16_plan(k, d3, d4).ok; }

template <int K>
static int walk16_launch(const Walk16Args &a, const Walk16Plan &w, cudaStream_t st) {
    cudaError_t e = cudaFuncSetAttribute(conv4d_tc_walk16_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)w.smem);
    NC_CHECK(e == cudaSuccess, "conv4d(tc,walk16): smem attribute: %s", cudaGetErrorString(e));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)(2 * a.G * a.cpg)); cfg.blockDim = dim3(kW16Threads); cfg.dynamicSmemBytes = w.smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    e = cudaLaunchKernelEx(&cfg, conv4d_tc_walk16_kernel<K>, a);
    NC_CHECK(e == cudaSuccess, "conv4d(tc,walk16): cluster launch: %s", cudaGetErrorString(e));
    NC_LAUNCH_OK("conv4d_tc_walk16_kernel");
    return 0;
}

// act: CB16 input; out (+ mask): CB16; wp: weights packed [G][rank][k*k][k][2][N/2][8] (pack_w_tc_kernel, pairsplit = 2)
int conv4d_tc_walk16_run(const __half *act, const __half *wp, const float *bias, const __half *mask, __half *out, __half *out_pm,
                         int B, int G, int Gx, int Cout, int d1, int d2, int d3, int d4, int k, int relu, cudaStream_t st) {
    const Walk16Plan w = walk16_plan(k, d3, d4);
    if (!w.ok) return 2;
    Walk16Args a{};
    a.act = act; a.wp = wp; a.bias = bias; a.mask = mask; a.out = out; a.out_pm = out_pm;
    a.B = B; a.G = G; a.Gx = Gx; a.Cout = Cout;
    a.d1 = d1; a.d2 = d2; a.d3 = d3; a.d4 = d4; a.P = d4 + k - 1;
    a.OUT = w.OUT; a.M_out = w.M_out; a.npairs = w.npairs;
    a.npos_pad = w.npos_pad; a.stages = w.stages;
    a.relu = relu;
    a.stage_bytes = w.stage_bytes; a.w_bytes = w.w_bytes; a.w_off = w.w_off; a.out_off = w.out_off;
    // instruction descriptor: D=f32, A=B=f16, both K-major, N>>3 at bit 17, M>>4 at bit 24 (M = 256: the pair)
    a.idesc = (1u << 4) | ((uint32_t)((k * 16) >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
    // one persistent pair per two SMs, split evenly between the branch groups
    int pairs = kNumSMs / 2;
    a.cpg = pairs / G > 0 ? pairs / G : 1;
    // work items: (b, iA, chunk pair) x jA segments; segments (each re-stages k-1 columns) only when there are too few
    // items to balance the pairs
    const long long base_items = (long long)B * d1 * w.npairs;
    int nseg = 1;
    if (base_items < 8ll * a.cpg) nseg = (int)((8ll * a.cpg + base_items - 1) / base_items);
    if (nseg > d2) nseg = d2;
    a.seg_len = (d2 + nseg - 1) / nseg;
    a.nseg = (d2 + a.seg_len - 1) / a.seg_len;
    const long long items = base_items * a.nseg;
    NC_CHECK(items < (1ll << 31), "conv4d(tc,walk16): too many work items");
    a.items = (int)items;
    if (a.cpg > a.items) a.cpg = a.items;
    return k == 5 ? walk16_launch<5>(a, w, st) : walk16_launch<3>(a, w, st);
}

}  // namespace ncnet
