// Conv4d on the 5th-gen tensor cores (mode NCNET_CONV_TC16): fp16 operands, fp32 accumulation in
// TMEM, one implicit-GEMM kernel per layer (reference semantics: lib/conv4d.py:11-51).
//
// Activation layout ("CB16", dtype code NCNET_CB16): channel-blocked, zero-haloed fp16
//     [block of 16 channels][channel half][iA*jA][iB + 2p][P = jB + 2p][8]        p = k/2
// One (iB,jB) plane with its halo is ONE contiguous run of (iB+2p)*P*16 bytes, so a pipeline stage is
// filled by one cp.async.bulk per channel half, and every position is one 16-byte UMMA core-matrix
// row.  Kernels that produce CB16 write the halo zeros themselves.
//
// GEMM view.  For one (batch, group, iA, jA) the CTA owns a chunk of the (iB,jB) plane, flattened
// with the pitch P so that a tap shift is an affine row offset:
//     A (M x K)  = activation rows: M = 128 consecutive flattened positions, K = 16 channels,
//                  K-major, no swizzle -> a tap is just a different descriptor start address;
//     B (N x K)  = weights of one (d1,d2[,d3]) tap for N = (tap-in-N, out-channel) columns;
//     D (M x N)  = fp32 accumulators in TMEM, NT tiles of N columns, live for the whole CTA.
// The taps that are folded into N are summed in the epilogue with the matching row shift
// (out[m] = sum_g D[m + shift_g][g]), which keeps N large enough (80 for a 16->16, k=5 layer)
// that one 4 KB A tile is amortised over 5 taps -- with N=16 the MMA would be bound by the
// 128 B/clk shared-memory read rate at 22 % of the tensor peak.
//   mode M1 (Cin<=16, Cout>1): channels = input channels, N = (d4, o), K-steps per plane = d3 taps
//   mode M2 (Cout == 1)      : channels = input channels, N = (d3, d4), one K-step per plane
//   mode M4 (Cin == 1)       : channels = the (d2,d4) taps, 16 per channel block (k=5: 25 taps in 2 blocks; k=3: 9 in 1),
//                              so a staged plane carries every d2 tap: 5x2 stages per output plane instead of 25;
//                              K-steps = d3 taps (row-shifted views), N = 16 out-channels, no shifted-sum epilogue
// Pipeline: one elected lane of warp 0 streams the (d1,d2) planes + their weight slab with
// cp.async.bulk (TMA engine, mbarrier complete_tx) through a 3-stage ring; one elected lane of warp 1
// issues tcgen05.mma and commits to the stage's "empty" barrier; warps 2-5 drain TMEM (tcgen05.ld),
// apply the shifted sums, bias, ReLU / ReLU-mask, and store planar or CB16.
#include "conv4d_tc.cuh"
#include "conv4d_common.cuh"
#include "tc_common.cuh"
#include <cstdlib>

namespace ncnet {

constexpr int kTcThreads = 192;
constexpr int kStages = 4;         // ring depth limit (the planner picks 2..4)
constexpr int kMaxGroups = 32;      // column groups (taps folded into N)
constexpr int kMaxKSteps = 5;
constexpr int kOutStride = 20;      // floats per outbuf row (16 channels + pad: conflict-free float4 access)

struct TcArgs {
    const __half *act;      // CB16 input
    const __half *wp;       // [G][k*k][KSTEPS][2][N][8] fp16
    const float *bias;      // [Cout] or null
    const void *mask;       // null | planar tensor shaped like the output | CB16 tensor shaped like the output
    void *out;              // planar [B, G*Cout, vol] (fp32/fp16) or CB16
    int B, G, Gx, Cout, CoutPad;
    int d1, d2, d3, d4, KS, P;
    int R, nchunks;         // output k-rows per CTA, chunks per plane
    int NT, N, NG, KSTEPS;
    int pair;               // 1: CTA pairs (cluster of 2, tcgen05 cta_group::2): the two plane chunks share one M256 MMA stream
    int OG;                 // branch groups side by side in N (1, or 2 for mode M4 on a shared input)
    int inner, m4;          // inner stage loop: the k d2 taps, or (mode M4) the channel blocks of the (d2,d4) taps
    int npos_pad;           // positions per channel-half in a smem stage
    int stages;             // ring depth (2 or 3)
    uint32_t tmem_cols;     // TMEM columns to allocate (power of two >= NT*N)
    int relu;
    int out_mode;           // 0 planar fp32, 1 planar fp16, 2 CB16
    int mask_mode;          // 0 none, 1 planar fp32, 2 planar fp16, 3 CB16
    uint32_t idesc;
    int aoff[kMaxKSteps];   // A row offset (positions) per K-step
    int gshift[kMaxGroups]; // epilogue row shift per column group
    uint32_t stage_bytes, data_bytes, w_bytes;
    __half *out_pm;         // optional position-major twin of a CB16 output (NCNET_CB16PM) or null
    uint32_t mask_off;      // CB16 mask: byte offset of its staging area in dynamic smem (0 = read it from global)
};

using namespace tc;

// MMA issue loop of one CTA.  NT_C / KS_C > 0: compile-time tile and K-step counts (straight-line issue);
// 0: run-time counts from the arguments.
// PAIR: issued by the leader CTA of a pair with cta_group::2 (M = 256: this CTA's 128 rows + the peer's, each CTA holds
// half of the N weight columns); waits for the peer's stage too and commits to both CTAs' barriers.
template <bool PAIR>
__device__ __forceinline__ void umma_issue(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    if (PAIR)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                     ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
    else
        umma_f16(d, a, b, idesc, acc);
}
template <bool PAIR>
__device__ __forceinline__ void umma_commit_to(uint32_t bar) {
    if (PAIR)
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                     ::"r"(bar), "h"((uint16_t)3) : "memory");
    else
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

template <int NT_C, int KS_C, bool PAIR = false>
__device__ __forceinline__ void mma_main(const TcArgs &p, int i, int j, uint8_t *stage0, uint64_t *full_bar, uint64_t *empty_bar,
                                         uint64_t *peer_full, uint32_t tmem_base, int &stage, uint32_t &phase) {
    constexpr int NTM = NT_C > 0 ? NT_C : 1, KSM = KS_C > 0 ? KS_C : 1;
    const int NT = NT_C > 0 ? NT_C : p.NT, KSTEPS = KS_C > 0 ? KS_C : p.KSTEPS;
    const int PAD = p.KS / 2;
    bool first = true;
    const uint32_t nloc = PAIR ? (uint32_t)p.N / 2u : (uint32_t)p.N;                              // weight columns held by this CTA
    const uint32_t lbo_a_field = (uint32_t)p.npos_pad << 16, lbo_b_field = nloc << 16;            // LBO in 16-byte units
    const uint32_t hi = (128u >> 4) | (1u << 14);                                                 // SBO = 128 B, descriptor version 1
    const uint32_t b_step = 2u * nloc;                                                            // 16-byte units per K-step of weights
    const uint32_t sm4 = smem_u32_pinned(stage0) >> 4, stage4 = p.stage_bytes >> 4, data4 = p.data_bytes >> 4;
    // low descriptor words relative to the stage base: A of (K-step, tile), B of K-step
    uint32_t a_rel[KSM][NTM], b_rel[KSM], d_col[NTM];
#pragma unroll
    for (int ks = 0; ks < KSM; ++ks) {
        b_rel[ks] = lbo_b_field | (data4 + (uint32_t)ks * b_step);
#pragma unroll
        for (int t = 0; t < NTM; ++t) a_rel[ks][t] = lbo_a_field | ((uint32_t)p.aoff[ks] + (uint32_t)t * 128u);
    }
#pragma unroll
    for (int t = 0; t < NTM; ++t) d_col[t] = tmem_base + (uint32_t)(t * p.N);
    // staged (d1,d2) planes of this CTA (the producer skips the out-of-range ones the same way)
    const int nplanes = (min(i + PAD, p.d1 - 1) - max(i - PAD, 0) + 1) *
                        (p.m4 ? p.inner : min(j + PAD, p.d2 - 1) - max(j - PAD, 0) + 1);
    const uint32_t full0 = smem_u32_pinned(full_bar), empty0 = smem_u32_pinned(empty_bar);
    const uint32_t peer0 = smem_u32_pinned(peer_full);
    if (elect_one()) {                      // one lane runs the whole loop: no per-stage election / reconvergence
        for (int it = 0; it < nplanes; ++it) {
            {
                mbar_wait_u32(full0 + (uint32_t)stage * 8u, phase);
                if (PAIR) mbar_wait_u32(peer0 + (uint32_t)stage * 8u, phase);       // the peer CTA's stage has landed too
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t s4 = sm4 + (uint32_t)stage * stage4;      // the address field never carries into LBO (smem < 256 KB)
                if (NT_C == 0) {
                    // run-time counts (any NT): descriptor words computed on the fly
                    for (int ks = 0; ks < KSTEPS; ++ks) {
                        const uint64_t bdesc = ((uint64_t)hi << 32) | (uint64_t)((lbo_b_field | (data4 + (uint32_t)ks * b_step)) + s4);
                        const uint32_t a0 = (lbo_a_field | (uint32_t)p.aoff[ks]) + s4;
                        const uint32_t acc = (first && ks == 0) ? 0u : 1u;
                        for (int t = 0; t < NT; ++t)
                                umma_issue<PAIR>(tmem_base + (uint32_t)(t * p.N), ((uint64_t)hi << 32) | (uint64_t)(a0 + (uint32_t)t * 128u), bdesc, p.idesc, acc);
                    }
                } else if (first) {
#pragma unroll
                    for (int ks = 0; ks < KSM; ++ks) {
                        const uint64_t bdesc = ((uint64_t)hi << 32) | (uint64_t)(b_rel[ks] + s4);
#pragma unroll
                        for (int t = 0; t < NTM; ++t)
                                umma_issue<PAIR>(d_col[t], ((uint64_t)hi << 32) | (uint64_t)(a_rel[ks][t] + s4), bdesc, p.idesc, ks == 0 ? 0u : 1u);
                    }
                } else {
#pragma unroll
                    for (int ks = 0; ks < KSM; ++ks) {
                        const uint64_t bdesc = ((uint64_t)hi << 32) | (uint64_t)(b_rel[ks] + s4);
#pragma unroll
                        for (int t = 0; t < NTM; ++t)
                                umma_issue<PAIR>(d_col[t], ((uint64_t)hi << 32) | (uint64_t)(a_rel[ks][t] + s4), bdesc, p.idesc, 1u);
                    }
                }
                // frees the smem slot when these MMAs have read it
                umma_commit_to<PAIR>(empty0 + (uint32_t)stage * 8u);
            }
            first = false;
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
    }
    __syncwarp();
}

// PAIR = true is a separate instantiation: a kernel that contains cta_group::2 / cluster instructions can only be launched
// as a cluster, so the plain launches use the PAIR = false code.
template <bool PAIR>
__global__ void __launch_bounds__(kTcThreads, 2)
conv4d_tc_kernel(const TcArgs p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t full_bar[kStages], empty_bar[kStages], peer_full[kStages], accum_bar, mask_bar;
    __shared__ uint32_t tmem_base_s;
    __shared__ int s_gshift[kMaxGroups];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t cta_rank = 0;          // position in the CTA pair (= chunk parity: the pair is the two chunks of one plane)
    if (PAIR) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta_rank));
    int bx = blockIdx.x;
    const int chunk = bx % p.nchunks; bx /= p.nchunks;
    const int j = bx % p.d2; bx /= p.d2;
    const int i = bx % p.d1; bx /= p.d1;
    const int g = bx % p.G;
    const int b = bx / p.G;
    const int PAD = p.KS / 2;
    const int k0 = chunk * p.R;
    const int rows_out = min(p.R, p.d3 - k0);
    const int rows_pad = p.d3 + 2 * PAD;                 // rows of a haloed plane
    const size_t plane_cb = (size_t)rows_pad * p.P;      // positions of a haloed plane
    const size_t ij_count = (size_t)p.d1 * p.d2;

    uint8_t *stage0 = smem;
    // The epilogue's output plane ALIASES the stage ring: it is first touched after accum_bar, i.e. after every
    // staged byte has been consumed, which leaves the whole budget to stages in flight (the main loop is bound by
    // bulk-copy latency x bytes in flight, not by the tensor pipe).
    const uint32_t osm = smem_u32_pinned(smem);       // all epilogue accesses go through this 32-bit shared address
    const int out_rows = p.NT * 128;

    // ---- one-time setup -------------------------------------------------------------------
    if (threadIdx.x < kMaxGroups) s_gshift[threadIdx.x] = p.gshift[threadIdx.x];
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); mbar_init(&peer_full[s], 1); }
        mbar_init(&accum_bar, 1);
        mbar_init(&mask_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {    // TMEM: NT accumulator tiles (256 columns or fewer lets two CTAs share the SM)
        if (PAIR) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(p.tmem_cols) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(p.tmem_cols) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    {
        // zero, once, the tail of each channel-half buffer behind the rows the bulk copies overwrite: it
        // is read by the last tiles' (discarded) rows and must stay finite
        const uint4 z = make_uint4(0, 0, 0, 0);
        const int copied = min(p.R + p.KS - 1, rows_pad - k0) * p.P;      // positions written by every stage's copy
        const int tail = p.npos_pad - copied;
        for (int e = threadIdx.x; e < p.stages * 2 * tail; e += kTcThreads) {
            const int s = e / (2 * tail), r = e - s * 2 * tail;
            const int half = r / tail, pos = copied + (r - half * tail);
            reinterpret_cast<uint4 *>(stage0 + (size_t)s * p.stage_bytes)[(size_t)half * p.npos_pad + pos] = z;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy zeros -> visible to TMA/UMMA
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (PAIR) {       // the peer's barriers must be initialised before remote arrives / multicast commits reach them
        asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
        asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;

    if (warp == 0) {
        // ===== producer =====  (a lone lane runs ~10 cycles per instruction: everything loop-invariant is hoisted)
        int stage = 0;
        uint32_t phase = 0;
        // the chunk's input rows (haloed coordinates k0 .. k0+R+KS-2), clipped to the plane
        const int rows_in = min(p.R + p.KS - 1, rows_pad - k0);
        const uint32_t half_bytes = (uint32_t)rows_in * p.P * 16u;
        const uint32_t wb = PAIR ? p.w_bytes / 2u : p.w_bytes;       // a pair's CTAs each hold half of the weight columns
        const uint32_t tx_bytes = 2u * half_bytes + wb;
        const size_t bg = (size_t)b * p.Gx + (p.Gx == 1 ? 0 : g);
        const size_t plane_halves = plane_cb * 8;
        // mode M4: the inner index selects the channel block (tensor block bg*inner + t2) of the SAME plane column j
        const size_t blk_halves = 2 * ij_count * plane_halves;
        const __half *src_base = p.act + (size_t)(p.m4 ? bg * p.inner : bg) * blk_halves + (size_t)k0 * p.P * 8;
        const size_t half_off = ij_count * plane_halves;
        const uint32_t w_halves = p.w_bytes / 2;
        const __half *w_base = p.wp + (size_t)g * p.KS * p.inner * w_halves;
        const uint32_t sm0 = smem_u32_pinned(stage0), full0 = smem_u32_pinned(&full_bar[0]);
        const uint32_t half1_off = (uint32_t)p.npos_pad * 16u;
        const uint32_t empty0 = smem_u32_pinned(&empty_bar[0]);
        if (elect_one()) {
        if (p.mask_mode == 3 && p.mask_off) {
            // the ReLU mask of this chunk's output rows (dgrad): staged now, read by the epilogue from shared memory
            const uint32_t mbytes = (uint32_t)rows_out * p.P * 16u, mb = smem_u32(&mask_bar);
            const __half *msrc = (const __half *)p.mask +
                ((((size_t)b * p.G + g) * 2) * ij_count + (size_t)i * p.d2 + j) * plane_cb * 8 + (size_t)(k0 + PAD) * p.P * 8;
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"(2u * mbytes) : "memory");
            bulk_g2s_u32(sm0 + p.mask_off, msrc, mbytes, mb);
            bulk_g2s_u32(sm0 + p.mask_off + mbytes, msrc + ij_count * plane_cb * 8, mbytes, mb);
        }
        for (int t1 = 0; t1 < p.KS; ++t1) {
            const int ii = i + t1 - PAD;
            if (ii < 0 || ii >= p.d1) continue;
            for (int t2 = 0; t2 < p.inner; ++t2) {
                const int jj = p.m4 ? j : j + t2 - PAD;
                if (jj < 0 || jj >= p.d2) continue;
                mbar_wait_u32(empty0 + (uint32_t)stage * 8u, phase ^ 1);
                {
                    const uint32_t sb = sm0 + (uint32_t)stage * p.stage_bytes, fb = full0 + (uint32_t)stage * 8u;
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fb), "r"(tx_bytes) : "memory");
                    const __half *src0 = src_base + (size_t)(ii * p.d2 + jj) * plane_halves + (p.m4 ? (size_t)t2 * blk_halves : 0);
                    bulk_g2s_u32(sb, src0, half_bytes, fb);
                    bulk_g2s_u32(sb + half1_off, src0 + half_off, half_bytes, fb);
                    bulk_g2s_u32(sb + p.data_bytes, w_base + (size_t)(t1 * p.inner + t2) * w_halves + (size_t)cta_rank * (wb / 2u), wb, fb);
                }
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
        }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===== MMA issuer =====
        // One lane issues every tcgen05.mma; it manages ~1 instruction per 10 cycles, while the tensor pipe can
        // retire an M128xN80xK16 MMA every 52 cycles.  The hot shapes therefore get a fully unrolled issue body
        // with the per-MMA descriptor offsets precomputed (one add per MMA); other shapes use the generic loop.
        bool done = false;
        int mstage = 0;
        uint32_t mphase = 0;
        if (PAIR) {
            done = true;
            if (cta_rank == 0) {
                if (p.NT == 3 && p.KSTEPS == 5) mma_main<3, 5, true>(p, i, j, stage0, full_bar, empty_bar, peer_full, tmem_base, mstage, mphase);
                else mma_main<0, 0, true>(p, i, j, stage0, full_bar, empty_bar, peer_full, tmem_base, mstage, mphase);
                if (elect_one()) umma_commit_to<true>(smem_u32(&accum_bar));
            } else if (elect_one()) {
                // peer CTA: no MMAs of its own; tell the leader when each of this CTA's stages has landed
                const int PADp = p.KS / 2;
                const int nplanes = (min(i + PADp, p.d1 - 1) - max(i - PADp, 0) + 1) * (min(j + PADp, p.d2 - 1) - max(j - PADp, 0) + 1);
                const uint32_t full0 = smem_u32_pinned(full_bar);
                uint32_t remote0;
                asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote0) : "r"(smem_u32(peer_full)), "r"(0));
                int stage = 0;
                uint32_t phase = 0;
                for (int it = 0; it < nplanes; ++it) {
                    mbar_wait_u32(full0 + (uint32_t)stage * 8u, phase);
                    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote0 + (uint32_t)stage * 8u) : "memory");
                    if (++stage == p.stages) { stage = 0; phase ^= 1; }
                }
            }
            __syncwarp();
        }
#define NC_TC_CASE(NT_, KS_) \
        if (!done && p.NT == NT_ && p.KSTEPS == KS_) { mma_main<NT_, KS_>(p, i, j, stage0, full_bar, empty_bar, peer_full, tmem_base, mstage, mphase); done = true; }
        NC_TC_CASE(3, 5) NC_TC_CASE(6, 5) NC_TC_CASE(2, 5) NC_TC_CASE(4, 5)
        NC_TC_CASE(1, 1) NC_TC_CASE(2, 1) NC_TC_CASE(3, 1) NC_TC_CASE(4, 1) NC_TC_CASE(5, 1) NC_TC_CASE(6, 1) NC_TC_CASE(7, 1) NC_TC_CASE(8, 1)
        NC_TC_CASE(3, 3) NC_TC_CASE(4, 3) NC_TC_CASE(2, 3)
#undef NC_TC_CASE
        if (!PAIR) {
        if (!done) mma_main<0, 0>(p, i, j, stage0, full_bar, empty_bar, peer_full, tmem_base, mstage, mphase);
        if (elect_one()) umma_commit(&accum_bar);
        }
        __syncwarp();
    } else {
        // ===== epilogue (warps 2..5 <-> TMEM lane quadrants warp%4) =====
        const int q = warp & 3;
        const int et = threadIdx.x - 64;                 // 0..127
        mbar_wait_sleep(&accum_bar, 0, 400);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        // OG > 1 (mode M4 on a shared input): the N columns hold OG branch groups of 16 out-channels side by side;
        // each is drained and stored as its own output group
        for (int og = 0; og < p.OG; ++og) {
        const int go = g * p.OG + og, Gt = p.G * p.OG;
        if (og) asm volatile("bar.sync 1, 128;" ::: "memory");       // the previous group's store has read the plane
        if (p.CoutPad >= 16) {
            // column group gi = 16 out-channels of one folded tap: out[m - shift_gi][o] += D[m][gi*16 + o].  The output
            // plane aliases the (now idle) ring.  Group 0 has shift 0 and touches every row: it WRITES bias + D, so
            // no initialisation pass is needed; the other groups accumulate.  Within a group every thread owns
            // distinct rows (all tiles), so one barrier per group orders the read-modify-writes; the TMEM loads of
            // two tiles are in flight together.
            float bv[16];
#pragma unroll
            for (int o = 0; o < 16; ++o) bv[o] = (p.bias && o < p.Cout) ? __ldg(p.bias + o) : 0.f;
            const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(og * 16);
            // iteration = (column group, pair of tiles); the TMEM loads of the next iteration are in flight while this
            // one is accumulated (tcgen05.ld has a few hundred cycles of latency)
            const int npair = (p.NT + 1) >> 1, nit = p.NG * npair;
            uint32_t r0[16], r1[16], n0[16], n1[16];
            tmem_ld16_issue(lane_base, r0);
            if (p.NT > 1) tmem_ld16_issue(lane_base + (uint32_t)p.N, r1);
            int gi = 0, t = 0;
            for (int it = 0; it < nit; ++it) {
                const bool two = t + 1 < p.NT;
                const int sh = s_gshift[gi];
                tmem_ld16_wait(r0);
                if (two) tmem_ld16_wait(r1);
                int gn = gi, tn = t + 2;
                if (tn >= p.NT) { tn = 0; ++gn; }
                if (it + 1 < nit) {
                    tmem_ld16_issue(lane_base + (uint32_t)(tn * p.N + gn * 16), n0);
                    if (tn + 1 < p.NT) tmem_ld16_issue(lane_base + (uint32_t)((tn + 1) * p.N + gn * 16), n1);
                }
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    if (h == 1 && !two) break;
                    const uint32_t(&r)[16] = h ? r1 : r0;
                    const int dest = (t + h) * 128 + q * 32 + lane - sh;
                    if (dest >= 0) {
                        const uint32_t dst = osm + (uint32_t)dest * (kOutStride * 4u);
                        if (gi == 0) {
#pragma unroll
                            for (int w4 = 0; w4 < 4; ++w4)
                                sts_v4(dst + 16u * w4,
                                       make_float4(bv[4 * w4 + 0] + __uint_as_float(r[4 * w4 + 0]), bv[4 * w4 + 1] + __uint_as_float(r[4 * w4 + 1]),
                                                   bv[4 * w4 + 2] + __uint_as_float(r[4 * w4 + 2]), bv[4 * w4 + 3] + __uint_as_float(r[4 * w4 + 3])));
                        } else {
                            float4 a[4];
#pragma unroll
                            for (int w4 = 0; w4 < 4; ++w4) a[w4] = lds_v4(dst + 16u * w4);
#pragma unroll
                            for (int w4 = 0; w4 < 4; ++w4) {
                                a[w4].x += __uint_as_float(r[4 * w4 + 0]); a[w4].y += __uint_as_float(r[4 * w4 + 1]);
                                a[w4].z += __uint_as_float(r[4 * w4 + 2]); a[w4].w += __uint_as_float(r[4 * w4 + 3]);
                                sts_v4(dst + 16u * w4, a[w4]);
                            }
                        }
                    }
                }
                if (gn != gi) asm volatile("bar.sync 1, 128;" ::: "memory");       // last tile pair of this group
                if (it + 1 < nit) {
                    tmem_ld16_wait(n0);
                    if (tn + 1 < p.NT) tmem_ld16_wait(n1);
#pragma unroll
                    for (int u = 0; u < 16; ++u) { r0[u] = n0[u]; r1[u] = n1[u]; }
                }
                gi = gn; t = tn;
            }
        } else {
            // single out-channel: column c of D holds tap (d3,d4) = (c / k, c % k) of output row m - shift_c.  The
            // columns are parked in shared memory as planes S[c][row] (lanes = consecutive rows: conflict-free, no
            // ordering between writers); the store below sums the k*k shifted reads of each output position.
            const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
            const bool wide = p.NG > 16;
            // tcgen05.ld has a few hundred cycles of latency: the next tile's loads are in flight while this one is stored
            uint32_t ra[16], rb[16], na[16], nb[16];
            tmem_ld16_issue(lane_base, ra);
            if (wide) tmem_ld16_issue(lane_base + 16u, rb);
            for (int t = 0; t < p.NT; ++t) {
                const int row = t * 128 + q * 32 + lane;
                tmem_ld16_wait(ra);
                if (wide) tmem_ld16_wait(rb);
                if (t + 1 < p.NT) {
                    tmem_ld16_issue(lane_base + (uint32_t)((t + 1) * p.N), na);
                    if (wide) tmem_ld16_issue(lane_base + (uint32_t)((t + 1) * p.N + 16), nb);
                }
#pragma unroll
                for (int u = 0; u < 16; ++u)
                    if (u < p.NG) sts_f32(osm + (uint32_t)(u * out_rows + row) * 4u, __uint_as_float(ra[u]));
                if (wide) {
#pragma unroll
                    for (int u = 0; u < 16; ++u)
                        if (16 + u < p.NG) sts_f32(osm + (uint32_t)((16 + u) * out_rows + row) * 4u, __uint_as_float(rb[u]));
                }
                if (t + 1 < p.NT) {
                    tmem_ld16_wait(na);
                    if (wide) tmem_ld16_wait(nb);
#pragma unroll
                    for (int u = 0; u < 16; ++u) { ra[u] = na[u]; rb[u] = nb[u]; }
                }
            }
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");

        if (p.out_mode == 2) {
            // ---- CB16 store: this chunk's rows of the haloed plane (halo columns/rows written as zeros), 16 B per
            //      (position, channel half), lanes along the contiguous position axis
            const int ra = chunk == 0 ? 0 : k0 + PAD;
            const int rb = chunk == p.nchunks - 1 ? rows_pad : k0 + PAD + rows_out;
            const int npos = (rb - ra) * p.P;
            const size_t base_off = ((((size_t)b * Gt + go) * 2) * ij_count + (size_t)i * p.d2 + j) * plane_cb * 8;
            __half *ob = (__half *)p.out + base_off;
            const __half *mb = (const __half *)p.mask + base_off;
            const bool smem_mask = p.mask_mode == 3 && p.mask_off;
            const uint32_t mbytes = (uint32_t)rows_out * p.P * 16u;
            if (smem_mask) mbar_wait(&mask_bar, 0);
            for (int e = et; e < 2 * npos; e += 128) {
                const int half = e / npos, pp = e - half * npos;
                const int kk = ra + pp / p.P, ll = pp % p.P;
                const int kr = kk - PAD - k0, lc = ll - PAD;
                const size_t goff = ((size_t)half * ij_count * plane_cb + (size_t)kk * p.P + ll) * 8;
                __align__(16) __half hv[8];
                if (kr >= 0 && kr < rows_out && lc >= 0 && lc < p.d4) {
                    const uint32_t src = osm + (uint32_t)((kr * p.P + lc) * kOutStride + half * 8) * 4u;
                    const float4 a = lds_v4(src), c = lds_v4(src + 16u);
                    float v[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
                    if (p.mask_mode == 3) {
                        const uint4 mraw = smem_mask ? lds_u4(osm + p.mask_off + (uint32_t)half * mbytes + (uint32_t)(kr * p.P + ll) * 16u)
                                                     : *reinterpret_cast<const uint4 *>(mb + goff);
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
                    const int pos = kk * p.P + ll;
                    const size_t pm_off = ((((size_t)b * Gt + go) * ij_count + (size_t)i * p.d2 + j) * plane_cb + pos) * 16 +
                                          (size_t)((half ^ ((pos >> 2) & 1)) * 8);
                    *reinterpret_cast<uint4 *>(p.out_pm + pm_off) = *reinterpret_cast<const uint4 *>(hv);
                }
            }
        } else {
            // ---- planar store: one warp per (channel, output row), lanes along l ----------------------
            const size_t plane_pos = (size_t)p.d3 * p.d4;
            const size_t vol = ij_count * plane_pos;
            const int ew = et >> 5;
            const int npairs = p.Cout * rows_out;
            const float bias0 = (p.CoutPad < 16 && p.bias) ? __ldg(p.bias) : 0.f;
            for (int pr = ew; pr < npairs; pr += 4) {
                const int o = pr / rows_out, rr = pr - o * rows_out;
                const size_t off0 = ((size_t)b * Gt * p.Cout + (size_t)go * p.Cout + o) * vol +
                                    ((size_t)i * p.d2 + j) * plane_pos + (size_t)(k0 + rr) * p.d4;
                for (int l = lane; l < p.d4; l += 32) {
                    const int row = rr * p.P + l;
                    float v;
                    if (p.CoutPad >= 16) v = lds_f32(osm + (uint32_t)(row * kOutStride + o) * 4u);
                    else {
                        // sum of the k*k column planes at their tap shifts (d3 * P + d4): independent loads
                        const uint32_t src = osm + (uint32_t)row * 4u, col = (uint32_t)out_rows * 4u;
                        float a0 = bias0, a1 = 0.f, a2 = 0.f, a3 = 0.f, a4 = 0.f;
                        if (p.KS == 5) {
#pragma unroll
                            for (int t3 = 0; t3 < 5; ++t3) {
                                const uint32_t s3 = src + (uint32_t)(t3 * 5) * col + (uint32_t)(t3 * p.P) * 4u;
                                a0 += lds_f32(s3); a1 += lds_f32(s3 + col + 4u); a2 += lds_f32(s3 + 2u * col + 8u);
                                a3 += lds_f32(s3 + 3u * col + 12u); a4 += lds_f32(s3 + 4u * col + 16u);
                            }
                        } else {
                            for (int c = 0; c < p.NG; ++c) a0 += lds_f32(src + (uint32_t)c * col + (uint32_t)s_gshift[c] * 4u);
                        }
                        v = (a0 + a1) + (a2 + a3) + a4;
                    }
                    if (p.relu) v = fmaxf(v, 0.f);
                    const size_t off = off0 + l;
                    if (p.mask_mode == 1) { if (!(((const float *)p.mask)[off] > 0.f)) v = 0.f; }
                    else if (p.mask_mode == 2) { if (!(__half2float(((const __half *)p.mask)[off]) > 0.f)) v = 0.f; }
                    if (p.out_mode == 1) ((__half *)p.out)[off] = __float2half_rn(v);
                    else ((float *)p.out)[off] = v;
                }
            }
        }
        }   // og
    }

    // ---- teardown ------------------------------------------------------------------------------
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (PAIR) {       // neither CTA of a pair may leave (or free TMEM) while the other can still be read by the shared MMAs
        asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
        asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    }
    if (warp == 0) {
        if (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
        else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
    }
}


// ---------------------------------------------------------------------------------------------
// Single-output-channel layers (mode M2) as a WALK along jA.
//
// In the kernel above a CTA owns one output plane (i, j) and stages its k*k input planes (i+t1-p, j+t2-p): every
// input plane travels from L2 to shared memory k*k times, and at Cout = 1 (N = 16 or 32 columns per MMA) that traffic,
// not the tensor pipe, bounds the layer (measured: 12.7 GB through the L2->SM crossbar for a 0.54 GB input at the
// PF-Pascal shape; the InLoc 16->1 layer moves 9x its 3.8 GB input).  Here a CTA owns (b, group, i, row chunk) and
// walks the input columns jj = 0 .. d2-1.  A staged plane (i+t1-p, jj) is used at once for the k outputs
// j = jj-p .. jj+p it contributes to (weights of tap t2 = jj-j+p), whose accumulators are k live slots of TMEM
// (slot = j mod k); each input plane is staged k times instead of k*k.  Output j is complete when column j+p is done:
// the issuer commits to the slot's "full" barrier and goes on with column j+p+1, whose MMAs touch the other k-1 live
// outputs first and the new output j+p+1+... (which reuses the freed slot) last, so the epilogue warps' drain of the
// slot overlaps the tensor pipe's work.  The epilogue is the M2 one: the k*k tap columns are parked in shared memory
// and each output position sums its k*k shifted reads, plus bias and ReLU, and stores planar.
// ---------------------------------------------------------------------------------------------
constexpr int kWalkStagesMax = 8;
constexpr int kWalkSlotsMax = 5;

struct WalkArgs {
    const __half *act;      // CB16 input
    const __half *wp;       // [G][k*k][2][N][8] fp16 (mode M2 packing)
    const float *bias;
    void *out;              // planar [B, G, vol]
    int B, G, Gx;
    int d1, d2, d3, d4, KS, P;
    int R, nchunks, NT, N, NG;
    int npos_pad, stages;
    int relu, out_f16;
    uint32_t idesc, stage_bytes, w_bytes, out_off, w_off;
};

// MMA issue loop of the walk (one lane).  K_C / NT_C > 0: compile-time kernel size and tile count -- the per-MMA work
// of the issuing lane is then two adds (it runs ~10 cycles per instruction, an M128xN32xK16 MMA retires in 40).
template <int K_C, int NT_C>
__device__ __forceinline__ void walk_issue(const WalkArgs &p, int t1_lo, int t1_hi, uint32_t sm0, uint32_t tmem_base,
                                           uint32_t full0, uint32_t empty0, uint32_t accf0, uint32_t acce0) {
    constexpr int KM = K_C > 0 ? K_C : 5, NTM = NT_C > 0 ? NT_C : 1;
    const int K = K_C > 0 ? K_C : p.KS, NT = NT_C > 0 ? NT_C : p.NT, PAD = K / 2;
    const uint32_t lbo_a = (uint32_t)p.npos_pad << 16, lbo_b = (uint32_t)p.N << 16;
    const uint64_t hi = (uint64_t)((128u >> 4) | (1u << 14)) << 32;                 // SBO = 128 B, descriptor version 1
    const uint32_t sm4 = sm0 >> 4, stage4 = p.stage_bytes >> 4, w4 = (sm0 + p.w_off) >> 4;
    const uint32_t slab4 = (uint32_t)p.N * 2u;                                      // 16-byte units per (t1,t2) weight slab
    const uint32_t slot_cols = (uint32_t)(NT * p.N);
    uint32_t a_rel[NTM];
#pragma unroll
    for (int t = 0; t < NTM; ++t) a_rel[t] = lbo_a | ((uint32_t)t * 128u);
    int stage = 0;
    uint32_t phase = 0;
    int slot_lo = (K - PAD) % K;                     // slot of output j = jj - PAD, kept incrementally (j mod K)
    for (int jj = 0; jj < p.d2; ++jj) {
        for (int t1 = t1_lo; t1 <= t1_hi; ++t1) {
            mbar_wait_u32(full0 + (uint32_t)stage * 8u, phase);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t s4 = sm4 + (uint32_t)stage * stage4;
            const uint32_t wrow = lbo_b | (w4 + (uint32_t)(t1 * K + (K - 1)) * slab4);     // tap t2 = K-1-dd for output dd
            if (NT_C > 0 && jj >= PAD && jj + PAD < p.d2) {
                // interior column: all k outputs j = jj-PAD+dd are live and only the last (j = jj+PAD) can be new.  A tile's A
                // is shared by the k outputs: tiles outer, outputs inner, A held in the tensor core's collector (fill, use ...,
                // lastuse: 16 / 22 cycles per N = 16 / 32 MMA instead of 40); straight-line issue.  A new output whose slot
                // is still being drained (its previous owner completed one column ago) goes last, without the hint.
                uint32_t dcol[KM];
                {
                    int slot = slot_lo;
#pragma unroll
                    for (int dd = 0; dd < KM; ++dd) { dcol[dd] = tmem_base + (uint32_t)slot * slot_cols; if (++slot == K) slot = 0; }
                }
                const bool nw = t1 == t1_lo;
                if (nw) {
#pragma unroll
                    for (int t = 0; t < NTM; ++t) {
                        const uint64_t adesc = hi | (uint64_t)(a_rel[t] + s4);
                        umma_f16_coll<1, true>(dcol[0] + (uint32_t)(t * p.N), adesc, hi | (uint64_t)wrow, p.idesc, 1u);
#pragma unroll
                        for (int dd = 1; dd < KM - 2; ++dd)
                            umma_f16_coll<2, true>(dcol[dd] + (uint32_t)(t * p.N), adesc, hi | (uint64_t)(wrow - (uint32_t)dd * slab4), p.idesc, 1u);
                        umma_f16_coll<3, true>(dcol[KM - 2] + (uint32_t)(t * p.N), adesc, hi | (uint64_t)(wrow - (uint32_t)(KM - 2) * slab4), p.idesc, 1u);
                    }
                    const int j = jj + PAD;
                    if (j >= K) {
                        mbar_wait_u32(acce0 + (uint32_t)((dcol[KM - 1] - tmem_base) / slot_cols) * 8u, (uint32_t)((j / K) - 1) & 1u);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    }
                    const uint64_t bdesc = hi | (uint64_t)(wrow - (uint32_t)(KM - 1) * slab4);
#pragma unroll
                    for (int t = 0; t < NTM; ++t)
                        umma_f16_coll<0, false>(dcol[KM - 1] + (uint32_t)(t * p.N), hi | (uint64_t)(a_rel[t] + s4), bdesc, p.idesc, 0u);
                } else {
#pragma unroll
                    for (int t = 0; t < NTM; ++t) {
                        const uint64_t adesc = hi | (uint64_t)(a_rel[t] + s4);
                        umma_f16_coll<1, true>(dcol[0] + (uint32_t)(t * p.N), adesc, hi | (uint64_t)wrow, p.idesc, 1u);
#pragma unroll
                        for (int dd = 1; dd < KM - 1; ++dd)
                            umma_f16_coll<2, true>(dcol[dd] + (uint32_t)(t * p.N), adesc, hi | (uint64_t)(wrow - (uint32_t)dd * slab4), p.idesc, 1u);
                        umma_f16_coll<3, true>(dcol[KM - 1] + (uint32_t)(t * p.N), adesc, hi | (uint64_t)(wrow - (uint32_t)(KM - 1) * slab4), p.idesc, 1u);
                    }
                }
            } else {
                int slot = slot_lo;
    #pragma unroll
                for (int dd = 0; dd < KM; ++dd) {        // outputs j = jj - PAD + dd, ascending: the newest comes last
                    if (dd < K) {
                        const int j = jj - PAD + dd;
                        if (j >= 0 && j < p.d2) {
                            const bool is_new = (jj == 0 || dd == K - 1) && t1 == t1_lo;
                            if (is_new && j >= K) {      // the slot's previous output (j - K) must have been drained
                                mbar_wait_u32(acce0 + (uint32_t)slot * 8u, (uint32_t)((j / K) - 1) & 1u);
                                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                            }
                            const uint64_t bdesc = hi | (uint64_t)(wrow - (uint32_t)dd * slab4);
                            const uint32_t dcol = tmem_base + (uint32_t)slot * slot_cols;
                            const uint32_t acc = is_new ? 0u : 1u;
                            if (NT_C > 0) {
    #pragma unroll
                                for (int t = 0; t < NTM; ++t)
                                    umma_f16(dcol + (uint32_t)(t * p.N), hi | (uint64_t)(a_rel[t] + s4), bdesc, p.idesc, acc);
                            } else {
                                for (int t = 0; t < NT; ++t)
                                    umma_f16(dcol + (uint32_t)(t * p.N), hi | (uint64_t)((lbo_a | ((uint32_t)t * 128u)) + s4), bdesc, p.idesc, acc);
                            }
                        }
                        if (++slot == K) slot = 0;
                    }
                }
            }
            umma_commit_to<false>(empty0 + (uint32_t)stage * 8u);
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
        // outputs whose last contributing column this was: j = jj - PAD, and at the last column every later one
        {
            int slot = slot_lo;
#pragma unroll
            for (int dd = 0; dd < KM; ++dd) {
                if (dd < K) {
                    const int j = jj - PAD + dd;
                    if (j >= 0 && j < p.d2 && min(p.d2 - 1, j + PAD) == jj) umma_commit_to<false>(accf0 + (uint32_t)slot * 8u);
                    if (++slot == K) slot = 0;
                }
            }
        }
        if (++slot_lo == K) slot_lo = 0;
    }
}

__global__ void __launch_bounds__(kTcThreads, 1)
conv4d_tc_walk_kernel(const WalkArgs p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t full_bar[kWalkStagesMax], empty_bar[kWalkStagesMax], w_bar;
    __shared__ __align__(8) uint64_t acc_full[kWalkSlotsMax], acc_empty[kWalkSlotsMax];
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int bx = blockIdx.x;
    const int chunk = bx % p.nchunks; bx /= p.nchunks;
    const int i = bx % p.d1; bx /= p.d1;
    const int g = bx % p.G;
    const int b = bx / p.G;
    const int K = p.KS, PAD = K / 2;
    const int k0 = chunk * p.R;
    const int rows_out = min(p.R, p.d3 - k0);
    const int rows_pad = p.d3 + 2 * PAD;
    const size_t plane_cb = (size_t)rows_pad * p.P;
    const size_t ij_count = (size_t)p.d1 * p.d2;
    const int out_rows = p.NT * 128;

    if (warp == 1 && lane == 0) {
        for (int s = 0; s < p.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int s = 0; s < K; ++s) { mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], 1); }
        mbar_init(&w_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    {   // the tail of each channel-half buffer behind the copied rows is read by the last tiles' (discarded) rows
        const uint4 z = make_uint4(0, 0, 0, 0);
        const int copied = min(p.R + K - 1, rows_pad - k0) * p.P;
        const int tail = p.npos_pad - copied;
        for (int e = threadIdx.x; e < p.stages * 2 * tail; e += kTcThreads) {
            const int s = e / (2 * tail), r = e - s * 2 * tail;
            const int half = r / tail, pos = copied + (r - half * tail);
            reinterpret_cast<uint4 *>(smem + (size_t)s * p.stage_bytes)[(size_t)half * p.npos_pad + pos] = z;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;
    const uint32_t sm0 = smem_u32_pinned(smem);
    const uint32_t full0 = smem_u32_pinned(full_bar), empty0 = smem_u32_pinned(empty_bar);
    const uint32_t accf0 = smem_u32_pinned(acc_full), acce0 = smem_u32_pinned(acc_empty);
    const int t1_lo = max(0, PAD - i), t1_hi = min(K - 1, p.d1 - 1 - i + PAD);      // valid d1 taps of this output row

    if (warp == 0) {
        // ===== producer: the weights once, then the planes (i+t1-p, jj) column by column =====
        if (elect_one()) {
            const uint32_t wb = smem_u32(&w_bar);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(wb), "r"(p.w_bytes) : "memory");
            bulk_g2s_u32(sm0 + p.w_off, p.wp + (size_t)g * (p.w_bytes / 2), p.w_bytes, wb);
            const int rows_in = min(p.R + K - 1, rows_pad - k0);
            const uint32_t half_bytes = (uint32_t)rows_in * p.P * 16u;
            const size_t bg = (size_t)b * p.Gx + (p.Gx == 1 ? 0 : g);
            const size_t plane_halves = plane_cb * 8;
            const __half *src_base = p.act + bg * 2 * ij_count * plane_halves + (size_t)k0 * p.P * 8;
            const size_t half_off = ij_count * plane_halves;
            const uint32_t half1_off = (uint32_t)p.npos_pad * 16u;
            int stage = 0;
            uint32_t phase = 0;
            for (int jj = 0; jj < p.d2; ++jj)
                for (int t1 = t1_lo; t1 <= t1_hi; ++t1) {
                    const int ii = i + t1 - PAD;
                    mbar_wait_u32(empty0 + (uint32_t)stage * 8u, phase ^ 1);
                    const uint32_t sb = sm0 + (uint32_t)stage * p.stage_bytes, fb = full0 + (uint32_t)stage * 8u;
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fb), "r"(2u * half_bytes) : "memory");
                    const __half *src0 = src_base + (size_t)(ii * p.d2 + jj) * plane_halves;
                    bulk_g2s_u32(sb, src0, half_bytes, fb);
                    bulk_g2s_u32(sb + half1_off, src0 + half_off, half_bytes, fb);
                    if (++stage == p.stages) { stage = 0; phase ^= 1; }
                }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (elect_one()) {
            mbar_wait(&w_bar, 0);
            if (K == 5 && p.NT == 3) walk_issue<5, 3>(p, t1_lo, t1_hi, sm0, tmem_base, full0, empty0, accf0, acce0);
            else if (K == 5 && p.NT == 2) walk_issue<5, 2>(p, t1_lo, t1_hi, sm0, tmem_base, full0, empty0, accf0, acce0);
            else if (K == 3 && p.NT == 10) walk_issue<3, 10>(p, t1_lo, t1_hi, sm0, tmem_base, full0, empty0, accf0, acce0);
            else if (K == 3 && p.NT == 9) walk_issue<3, 9>(p, t1_lo, t1_hi, sm0, tmem_base, full0, empty0, accf0, acce0);
            else walk_issue<0, 0>(p, t1_lo, t1_hi, sm0, tmem_base, full0, empty0, accf0, acce0);
        }
        __syncwarp();
    } else {
        // ===== epilogue: one output plane (chunk) per step of the walk =====
        const int q = warp & 3;
        const int et = threadIdx.x - 64;
        const uint32_t osm = sm0 + p.out_off;
        const bool wide = p.NG > 16;
        const float bias0 = p.bias ? __ldg(p.bias) : 0.f;
        const size_t plane_pos = (size_t)p.d3 * p.d4;
        const size_t vol = ij_count * plane_pos;
        for (int j = 0; j < p.d2; ++j) {
            const int slot = j % K;
            mbar_wait_sleep(&acc_full[slot], (uint32_t)(j / K) & 1u, 200);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(slot * p.NT * p.N);
            // the next tile's TMEM loads are in flight while this tile's columns are parked
            uint32_t ra[16], rb[16], na[16], nb[16];
            tmem_ld16_issue(lane_base, ra);
            if (wide) tmem_ld16_issue(lane_base + 16u, rb);
            for (int t = 0; t < p.NT; ++t) {
                const int row = t * 128 + q * 32 + lane;
                tmem_ld16_wait(ra);
                if (wide) tmem_ld16_wait(rb);
                if (t + 1 < p.NT) {
                    tmem_ld16_issue(lane_base + (uint32_t)((t + 1) * p.N), na);
                    if (wide) tmem_ld16_issue(lane_base + (uint32_t)((t + 1) * p.N + 16), nb);
                }
#pragma unroll
                for (int u = 0; u < 16; ++u)
                    if (u < p.NG) sts_f32(osm + (uint32_t)(u * out_rows + row) * 4u, __uint_as_float(ra[u]));
                if (wide) {
#pragma unroll
                    for (int u = 0; u < 16; ++u)
                        if (16 + u < p.NG) sts_f32(osm + (uint32_t)((16 + u) * out_rows + row) * 4u, __uint_as_float(rb[u]));
                }
                if (t + 1 < p.NT) {
                    tmem_ld16_wait(na);
                    if (wide) tmem_ld16_wait(nb);
#pragma unroll
                    for (int u = 0; u < 16; ++u) { ra[u] = na[u]; rb[u] = nb[u]; }
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (et == 0) {       // the slot may be overwritten: every warp's TMEM reads have completed
                asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(acce0 + (uint32_t)slot * 8u) : "memory");
            }
            // planar store: one warp per output row, lanes along l; the k*k tap columns summed at their shifts
            const int ew = et >> 5;
            for (int rr = ew; rr < rows_out; rr += 4) {
                const size_t off0 = ((size_t)b * p.G + g) * vol + ((size_t)i * p.d2 + j) * plane_pos + (size_t)(k0 + rr) * p.d4;
                for (int l = lane; l < p.d4; l += 32) {
                    const uint32_t src = osm + (uint32_t)(rr * p.P + l) * 4u, col = (uint32_t)out_rows * 4u;
                    float a0 = bias0, a1 = 0.f, a2 = 0.f;
                    if (K == 5) {
                        float a3 = 0.f, a4 = 0.f;
#pragma unroll
                        for (int t3 = 0; t3 < 5; ++t3) {
                            const uint32_t s3 = src + (uint32_t)(t3 * 5) * col + (uint32_t)(t3 * p.P) * 4u;
                            a0 += lds_f32(s3); a1 += lds_f32(s3 + col + 4u); a2 += lds_f32(s3 + 2u * col + 8u);
                            a3 += lds_f32(s3 + 3u * col + 12u); a4 += lds_f32(s3 + 4u * col + 16u);
                        }
                        a0 = (a0 + a1) + (a2 + a3) + a4;
                    } else {
#pragma unroll
                        for (int t3 = 0; t3 < 3; ++t3) {
                            const uint32_t s3 = src + (uint32_t)(t3 * 3) * col + (uint32_t)(t3 * p.P) * 4u;
                            a0 += lds_f32(s3); a1 += lds_f32(s3 + col + 4u); a2 += lds_f32(s3 + 2u * col + 8u);
                        }
                        a0 = (a0 + a1) + a2;
                    }
                    if (p.relu) a0 = fmaxf(a0, 0.f);
                    if (p.out_f16) ((__half *)p.out)[off0 + l] = __float2half_rn(a0);
                    else ((float *)p.out)[off0 + l] = a0;
                }
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");        // the column planes are free for the next output
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
}

// ---------------------------------------------------------------------------------------------
// CB16 packing
// ---------------------------------------------------------------------------------------------
// planar [nblk, C, vol] (T) -> CB16 [nblk][2][ij][d3+2p][P][8], halo = 0.
//   tap_mode 0: channels = the tensor's C (<= 16) channels
//   tap_mode 1: C == 1, channel t = x[.., l + t - p]                 (taps along the last axis)
//   tap_mode 2: C == 1, channel t = x[.., j - t + p, ..]             (taps along the second axis, mirrored)
template <typename T>
__global__ void __launch_bounds__(256)
cb16_pack_kernel(const T *__restrict__ x, __half *__restrict__ cb, int C, int d1, int d2, int d3, int d4, int pad,
                 int tap_mode, int taps) {
    const size_t blk = blockIdx.y;
    const int half = blockIdx.z;
    const int P = d4 + 2 * pad, rows = d3 + 2 * pad;
    const size_t ij = (size_t)d1 * d2, plane = (size_t)d3 * d4, vol = ij * plane;
    const size_t npad = ij * rows * P;
    for (size_t pp = (size_t)blockIdx.x * blockDim.x + threadIdx.x; pp < npad; pp += (size_t)gridDim.x * blockDim.x) {
        const int ll = (int)(pp % P);
        const int kk = (int)((pp / P) % rows);
        const size_t ijx = pp / ((size_t)P * rows);
        const int k = kk - pad, l = ll - pad;
        __align__(16) __half v[8];
        if (k >= 0 && k < d3 && l >= 0 && l < d4) {
            const size_t pos = ijx * plane + (size_t)k * d4 + l;
            if (tap_mode == 0) {
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const int ch = half * 8 + c;
                    v[c] = __float2half_rn(ch < C ? ldf(x + (blk * C + ch) * vol + pos) : 0.f);
                }
            } else if (tap_mode == 1) {
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const int t = half * 8 + c, lt = l + t - pad;
                    v[c] = __float2half_rn((t < taps && lt >= 0 && lt < d4) ? ldf(x + blk * vol + pos + (t - pad)) : 0.f);
                }
            } else {
                const int jx = (int)(ijx % d2);
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const int t = half * 8 + c, jt = jx - t + pad;
                    v[c] = __float2half_rn((t < taps && jt >= 0 && jt < d2)
                                               ? ldf(x + blk * vol + pos + (ptrdiff_t)(pad - t) * (ptrdiff_t)plane) : 0.f);
                }
            }
        } else {
#pragma unroll
            for (int c = 0; c < 8; ++c) v[c] = __float2half_rn(0.f);
        }
        *reinterpret_cast<uint4 *>(cb + ((blk * 2 + half) * npad + pp) * 8) = *reinterpret_cast<const uint4 *>(v);
    }
}

// Same mapping for planes whose haloed row fits one warp (P <= 32): one warp per haloed row, lanes along the row, 32-bit
// index math; the tap modes along the last axis load each element once and exchange it with shuffles.
template <typename T>
__global__ void __launch_bounds__(256)
cb16_pack_rows_kernel(const T *__restrict__ x, __half *__restrict__ cb, int C, int d1, int d2, int d3, int d4, int pad,
                      int tap_mode, int taps) {
    const unsigned blk = blockIdx.y;
    const int half = blockIdx.z;
    const int P = d4 + 2 * pad, rows = d3 + 2 * pad;
    const unsigned ij = (unsigned)(d1 * d2), nrows = ij * (unsigned)rows;
    const size_t plane = (size_t)d3 * d4, vol = (size_t)ij * plane;
    const int lane = threadIdx.x & 31;
    const unsigned r = blockIdx.x * 8u + (threadIdx.x >> 5);
    if (r >= nrows) return;
    const unsigned ijx = r / (unsigned)rows;
    const int kk = (int)(r - ijx * (unsigned)rows);
    const int k = kk - pad, l = lane - pad;
    const bool inside = k >= 0 && k < d3 && l >= 0 && l < d4;
    __align__(16) __half v[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) v[c] = __float2half_rn(0.f);
    if (tap_mode == 1) {
        // lane holds x[row][lane - 2p]; channel c of haloed column ll is the value of lane ll + c
        const int ks = k;
        const int col = lane - 2 * pad;
        float e = 0.f;
        if (ks >= 0 && ks < d3 && col >= 0 && col < d4) e = ldf(x + (size_t)blk * vol + (size_t)ijx * plane + (size_t)ks * d4 + col);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const float o = __shfl_down_sync(0xffffffffu, e, c);
            if (c < taps && lane + c < 32 && inside) v[c] = __float2half_rn(o);
        }
    } else if (inside) {
        const size_t pos = (size_t)ijx * plane + (size_t)k * d4 + l;
        if (tap_mode == 0) {
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const int ch = half * 8 + c;
                if (ch < C) v[c] = __float2half_rn(ldf(x + ((size_t)blk * C + ch) * vol + pos));
            }
        } else {
            const int jx = (int)(ijx % (unsigned)d2);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const int t = half * 8 + c, jt = jx - t + pad;
                if (t < taps && jt >= 0 && jt < d2)
                    v[c] = __float2half_rn(ldf(x + (size_t)blk * vol + pos + (ptrdiff_t)(pad - t) * (ptrdiff_t)plane));
            }
        }
    }
    if (lane < P) {
        const size_t npad = (size_t)nrows * P;
        *reinterpret_cast<uint4 *>(cb + (((size_t)blk * 2 + half) * npad + (size_t)r * P + lane) * 8) = *reinterpret_cast<const uint4 *>(v);
    }
}

// Rows wider than a warp: the same warp-per-row mapping walked in 32-column segments, direct (L1-served) tap loads.
template <typename T>
__global__ void __launch_bounds__(256)
cb16_pack_wide_rows_kernel(const T *__restrict__ x, __half *__restrict__ cb, int C, int d1, int d2, int d3, int d4, int pad,
                      int tap_mode, int taps) {
    const unsigned blk = blockIdx.y;
    const int half = blockIdx.z;
    const int P = d4 + 2 * pad, rows = d3 + 2 * pad;
    const unsigned ij = (unsigned)(d1 * d2), nrows = ij * (unsigned)rows;
    const size_t plane = (size_t)d3 * d4, vol = (size_t)ij * plane;
    const int lane = threadIdx.x & 31;
    const unsigned r = blockIdx.x * 8u + (threadIdx.x >> 5);
    if (r >= nrows) return;
    const unsigned ijx = r / (unsigned)rows;
    const int kk = (int)(r - ijx * (unsigned)rows);
    const int k = kk - pad;
    const size_t npad = (size_t)nrows * P;
    __half *dst = cb + (((size_t)blk * 2 + half) * npad + (size_t)r * P) * 8;
    for (int ll0 = 0; ll0 < P; ll0 += 32) {
        const int ll = ll0 + lane, l = ll - pad;
        const bool inside = k >= 0 && k < d3 && l >= 0 && l < d4;
        __align__(16) __half v[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) v[c] = __float2half_rn(0.f);
        if (tap_mode == 1) {
            const int ks = k;
            const bool row_ok = ks >= 0 && ks < d3;
            const T *xr = x + (size_t)blk * vol + (size_t)ijx * plane + (size_t)(row_ok ? ks : 0) * d4;
            if (row_ok && inside) {
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const int col = ll - 2 * pad + c;
                    if (c < taps && col >= 0 && col < d4) v[c] = __float2half_rn(ldf(xr + col));
                }
            }
        } else if (inside) {
            const size_t pos = (size_t)ijx * plane + (size_t)k * d4 + l;
            if (tap_mode == 0) {
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const int ch = half * 8 + c;
                    if (ch < C) v[c] = __float2half_rn(ldf(x + ((size_t)blk * C + ch) * vol + pos));
                }
            } else {
                const int jx = (int)(ijx % (unsigned)d2);
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const int t = half * 8 + c, jt = jx - t + pad;
                    if (t < taps && jt >= 0 && jt < d2)
                        v[c] = __float2half_rn(ldf(x + (size_t)blk * vol + pos + (ptrdiff_t)(pad - t) * (ptrdiff_t)plane));
                }
            }
        }
        if (ll < P) *reinterpret_cast<uint4 *>(dst + (size_t)ll * 8) = *reinterpret_cast<const uint4 *>(v);
    }
}

// tap_mode 4 (forward mode M4): single-channel planar [nblk, vol] -> nblk*ngrp CB16 blocks.  Channel c of block
// (blk, gr) at haloed position (i, j, kk, ll) is the (d2,d4) tap (t2,t4) = divmod(gr*16 + c, k):
//     x[blk, i, j + t2 - p, kk - p, ll + t4 - p]          (zero outside the volume or for gr*16 + c >= k*k),
// i.e. the d4 tap is taken relative to the column index ll itself, so a row-shifted view (the d3 taps) of this
// plane lines up with the output position (kr, lc) at row kr + t3, column lc.  One warp per haloed row.
template <typename T, int k>
__global__ void __launch_bounds__(256)
cb16_pack_taps2d_kernel(const T *__restrict__ x, __half *__restrict__ cb, int d1, int d2, int d3, int d4) {
    constexpr int pad = k / 2, ngrp = (k * k + 15) / 16;          // compile-time: the tap divisions below fold away
    const unsigned blk_out = blockIdx.y;
    const unsigned blk = blk_out / (unsigned)ngrp;
    const int gr = (int)(blk_out - blk * (unsigned)ngrp), half = blockIdx.z;
    const int P = d4 + 2 * pad, rows = d3 + 2 * pad;
    const unsigned ij = (unsigned)(d1 * d2), nrows = ij * (unsigned)rows;
    const size_t plane = (size_t)d3 * d4, vol = (size_t)ij * plane;
    const int lane = threadIdx.x & 31;
    const unsigned r = blockIdx.x * 8u + (threadIdx.x >> 5);
    if (r >= nrows) return;
    const unsigned ijx = r / (unsigned)rows;
    const int kk = (int)(r - ijx * (unsigned)rows), kq = kk - pad;
    const int jx = (int)(ijx % (unsigned)d2);
    __half *dst = cb + (((size_t)blk_out * 2 + half) * ((size_t)nrows * P) + (size_t)r * P) * 8;
    const bool row_ok = kq >= 0 && kq < d3;
    const T *xrow = x + (size_t)blk * vol + (size_t)ijx * plane + (size_t)(row_ok ? kq : 0) * d4;
    // the 8 channels of this thread: tap offsets and plane pointers are the same for every column of the row
    int lt0[8];
    const T *src[8];
#pragma unroll
    for (int c8 = 0; c8 < 8; ++c8) {
        const int c = gr * 16 + half * 8 + c8;
        const int t2 = c / k, t4 = c - t2 * k;
        const int jt = jx + t2 - pad;
        const bool ok = row_ok && c < k * k && jt >= 0 && jt < d2;
        lt0[c8] = ok ? t4 - pad : -(1 << 20);                      // pushes the column out of range when the tap is invalid
        src[c8] = xrow + (ok ? (ptrdiff_t)(t2 - pad) * (ptrdiff_t)plane : 0);
    }
    for (int ll = lane; ll < P; ll += 32) {
        __align__(16) __half v[8];
#pragma unroll
        for (int c8 = 0; c8 < 8; ++c8) {
            const int lt = ll + lt0[c8];
            v[c8] = __float2half_rn((lt >= 0 && lt < d4) ? ldf(src[c8] + lt) : 0.f);
        }
        *reinterpret_cast<uint4 *>(dst + (size_t)ll * 8) = *reinterpret_cast<const uint4 *>(v);
    }
}

// Narrow rows (P <= 32) of tap_mode 4: block = one (i,j) plane of one (output block, channel half); the eight
// channels' (d2,d4) taps are compile-time (BASE = first tap of this channel octet), each source row segment is loaded
// once per d2 tap and handed to the channels with shuffles.
template <typename T, int k, int BASE>
__device__ __forceinline__ void taps2d_rows(const T *__restrict__ xp, __half *__restrict__ dst, int jx, int d2, int d3, int d4,
                                            size_t plane, int warp, int lane) {
    constexpr int pad = k / 2, T2A = BASE / k;
    const int P = d4 + 2 * pad, rows = d3 + 2 * pad;
    const int col = lane - pad;
    const bool col_ok = col >= 0 && col < d4;
    for (int kk = warp; kk < rows; kk += 8) {
        const int kq = kk - pad;
        const bool row_ok = kq >= 0 && kq < d3;
        float e[3];
#pragma unroll
        for (int u = 0; u < 3; ++u) {
            const int jt = jx + T2A + u - pad;
            e[u] = (row_ok && col_ok && T2A + u < k && jt >= 0 && jt < d2)
                       ? ldf(xp + (ptrdiff_t)(T2A + u - pad) * (ptrdiff_t)plane + (size_t)kq * d4 + col) : 0.f;
        }
        __align__(16) __half v[8];
#pragma unroll
        for (int c8 = 0; c8 < 8; ++c8) {
            constexpr int dummy = 0; (void)dummy;
            const int c = BASE + c8;                     // compile-time after unrolling
            const int t2 = c / k, t4 = c - t2 * k;
            float o = 0.f;
            if (c < k * k) {
                const int srcl = lane + t4;              // column (lane + t4 - p) is held by lane (lane + t4)
                const float got = __shfl_sync(0xffffffffu, e[t2 - T2A], srcl & 31);
                o = srcl < 32 ? got : 0.f;
            }
            v[c8] = __float2half_rn(o);
        }
        if (lane < P) *reinterpret_cast<uint4 *>(dst + ((size_t)kk * P + lane) * 8) = *reinterpret_cast<const uint4 *>(v);
    }
}

template <typename T, int k>
__global__ void __launch_bounds__(256)
cb16_pack_taps2d_narrow_kernel(const T *__restrict__ x, __half *__restrict__ cb, int d1, int d2, int d3, int d4) {
    constexpr int pad = k / 2, ngrp = (k * k + 15) / 16;
    const int jx = blockIdx.x, ix = blockIdx.y;
    const unsigned z = blockIdx.z, blk_out = z >> 1, half = z & 1u;
    const unsigned blk = blk_out / ngrp, gr = blk_out - blk * ngrp;
    const int P = d4 + 2 * pad, rows = d3 + 2 * pad;
    const size_t ij = (size_t)d1 * d2, plane = (size_t)d3 * d4;
    const size_t ijx = (size_t)ix * d2 + jx;
    const T *xp = x + ((size_t)blk * ij + ijx) * plane;
    __half *dst = cb + ((((size_t)blk_out * 2 + half) * ij + ijx) * rows * P) * 8;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int base = (int)(gr * 16 + half * 8);
    if (base == 0) taps2d_rows<T, k, 0>(xp, dst, jx, d2, d3, d4, plane, warp, lane);
    else if (base == 8) taps2d_rows<T, k, 8>(xp, dst, jx, d2, d3, d4, plane, warp, lane);
    else if (base == 16) taps2d_rows<T, k, 16>(xp, dst, jx, d2, d3, d4, plane, warp, lane);
    else taps2d_rows<T, k, 24>(xp, dst, jx, d2, d3, d4, plane, warp, lane);
}

// CB16 -> planar [nblk, C, vol] (T); interior only
template <typename T>
__global__ void __launch_bounds__(256)
cb16_unpack_kernel(const __half *__restrict__ cb, T *__restrict__ x, int C, int d1, int d2, int d3, int d4, int pad) {
    const size_t blk = blockIdx.y;
    const int P = d4 + 2 * pad, rows = d3 + 2 * pad;
    const size_t ij = (size_t)d1 * d2, plane = (size_t)d3 * d4, vol = ij * plane;
    const size_t npad = ij * rows * P;
    for (size_t pos = (size_t)blockIdx.x * blockDim.x + threadIdx.x; pos < vol; pos += (size_t)gridDim.x * blockDim.x) {
        const int l = (int)(pos % d4);
        const int k = (int)((pos / d4) % d3);
        const size_t ijx = pos / plane;
        const size_t pp = (ijx * rows + (k + pad)) * P + (l + pad);
        for (int ch = 0; ch < C; ++ch)
            stf(x + (blk * C + ch) * vol + pos, __half2float(cb[((blk * 2 + (ch >> 3)) * npad + pp) * 8 + (ch & 7)]));
    }
}

size_t cb16_bytes(size_t nblk, int d1, int d2, int d3, int d4, int k) {
    const int p = k / 2;
    return nblk * 2 * (size_t)d1 * d2 * (d3 + 2 * p) * (d4 + 2 * p) * 16;
}

int cb16_pack(const void *x, int dtype, void *cb, int nblk, int C, int d1, int d2, int d3, int d4, int k, int tap_mode,
              cudaStream_t st) {
    NC_CHECK(nblk > 0 && nblk <= 65535 && C >= 1 && C <= 16, "cb16_pack: bad shape (blocks %d, channels %d)", nblk, C);
    NC_CHECK(tap_mode == 0 || C == 1, "cb16_pack: tap modes need a single-channel input");
    const int pad = k / 2;
    if (tap_mode == 4) {
        const int ngrp = (k * k + 15) / 16;
        const size_t nrows4 = (size_t)d1 * d2 * (d3 + 2 * pad);
        NC_CHECK((nrows4 + 7) / 8 < (1ull << 31) && (long long)nblk * ngrp <= 65535, "cb16_pack: grid too large");
        NC_CHECK(dtype == NCNET_F32 || dtype == NCNET_F16, "cb16_pack: source dtype %d must be planar fp32/fp16", dtype);
        if (d4 + 2 * pad <= 32 && d1 <= 65535 && (long long)nblk * ngrp * 2 <= 65535) {
            dim3 g2(d2, d1, nblk * ngrp * 2);
            if (dtype == NCNET_F32) {
                if (k == 5) cb16_pack_taps2d_narrow_kernel<float, 5><<<g2, 256, 0, st>>>((const float *)x, (__half *)cb, d1, d2, d3, d4);
                else cb16_pack_taps2d_narrow_kernel<float, 3><<<g2, 256, 0, st>>>((const float *)x, (__half *)cb, d1, d2, d3, d4);
            } else {
                if (k == 5) cb16_pack_taps2d_narrow_kernel<__half, 5><<<g2, 256, 0, st>>>((const __half *)x, (__half *)cb, d1, d2, d3, d4);
                else cb16_pack_taps2d_narrow_kernel<__half, 3><<<g2, 256, 0, st>>>((const __half *)x, (__half *)cb, d1, d2, d3, d4);
            }
            NC_LAUNCH_OK("cb16_pack_taps2d_narrow_kernel");
            return 0;
        }
        dim3 grid((unsigned)((nrows4 + 7) / 8), nblk * ngrp, 2);
        if (dtype == NCNET_F32) {
            if (k == 5) cb16_pack_taps2d_kernel<float, 5><<<grid, 256, 0, st>>>((const float *)x, (__half *)cb, d1, d2, d3, d4);
            else cb16_pack_taps2d_kernel<float, 3><<<grid, 256, 0, st>>>((const float *)x, (__half *)cb, d1, d2, d3, d4);
        } else {
            if (k == 5) cb16_pack_taps2d_kernel<__half, 5><<<grid, 256, 0, st>>>((const __half *)x, (__half *)cb, d1, d2, d3, d4);
            else cb16_pack_taps2d_kernel<__half, 3><<<grid, 256, 0, st>>>((const __half *)x, (__half *)cb, d1, d2, d3, d4);
        }
        NC_LAUNCH_OK("cb16_pack_taps2d_kernel");
        return 0;
    }
    const size_t npad = (size_t)d1 * d2 * (d3 + 2 * pad) * (d4 + 2 * pad);
    // tap modes 1 and 2 put at most k <= 8 taps into the channels: the second channel half is never read
    const int halves = (tap_mode == 1 || tap_mode == 2) ? 1 : 2;
    NC_CHECK(dtype == NCNET_F32 || dtype == NCNET_F16, "cb16_pack: source dtype %d must be planar fp32/fp16", dtype);
    const size_t nrows = (size_t)d1 * d2 * (d3 + 2 * pad);
    if ((nrows + 7) / 8 < (1ull << 31)) {
        dim3 grid((unsigned)((nrows + 7) / 8), nblk, halves);
        const bool narrow = d4 + 2 * pad <= 32;
        if (dtype == NCNET_F32) {
            if (narrow) cb16_pack_rows_kernel<float><<<grid, 256, 0, st>>>((const float *)x, (__half *)cb, C, d1, d2, d3, d4, pad, tap_mode, k);
            else cb16_pack_wide_rows_kernel<float><<<grid, 256, 0, st>>>((const float *)x, (__half *)cb, C, d1, d2, d3, d4, pad, tap_mode, k);
        } else {
            if (narrow) cb16_pack_rows_kernel<__half><<<grid, 256, 0, st>>>((const __half *)x, (__half *)cb, C, d1, d2, d3, d4, pad, tap_mode, k);
            else cb16_pack_wide_rows_kernel<__half><<<grid, 256, 0, st>>>((const __half *)x, (__half *)cb, C, d1, d2, d3, d4, pad, tap_mode, k);
        }
        NC_LAUNCH_OK("cb16_pack_rows_kernel");
        return 0;
    }
    const unsigned gx = (unsigned)((npad + 255) / 256 < 8192 ? (npad + 255) / 256 : 8192);
    dim3 grid(gx, nblk, halves);
    if (dtype == NCNET_F32)
        cb16_pack_kernel<float><<<grid, 256, 0, st>>>((const float *)x, (__half *)cb, C, d1, d2, d3, d4, pad, tap_mode, k);
    else
        cb16_pack_kernel<__half><<<grid, 256, 0, st>>>((const __half *)x, (__half *)cb, C, d1, d2, d3, d4, pad, tap_mode, k);
    NC_LAUNCH_OK("cb16_pack_kernel");
    return 0;
}

int cb16_unpack(const void *cb, void *x, int dtype, int nblk, int C, int d1, int d2, int d3, int d4, int k, cudaStream_t st) {
    NC_CHECK(nblk > 0 && nblk <= 65535 && C >= 1 && C <= 16, "cb16_unpack: bad shape");
    const size_t vol = (size_t)d1 * d2 * d3 * d4;
    const unsigned gx = (unsigned)((vol + 255) / 256 < 8192 ? (vol + 255) / 256 : 8192);
    dim3 grid(gx, nblk);
    if (dtype == NCNET_F32)
        cb16_unpack_kernel<float><<<grid, 256, 0, st>>>((const __half *)cb, (float *)x, C, d1, d2, d3, d4, k / 2);
    else if (dtype == NCNET_F16)
        cb16_unpack_kernel<__half><<<grid, 256, 0, st>>>((const __half *)cb, (__half *)x, C, d1, d2, d3, d4, k / 2);
    else
        return fail("cb16_unpack: destination dtype %d must be planar fp32/fp16", dtype);
    NC_LAUNCH_OK("cb16_unpack_kernel");
    return 0;
}

struct TcPackW {
    PackDesc d;
    int mode;       // 1,2,3,4
    int KSTEPS, N, CoutPad, inner, OG;
    int pairsplit;  // 1: per (d1,d2) slab the N columns are split in two contiguous halves [rank][K-step][2][N/2][8] (CTA pairs)
                    // 2: the split is outermost within a group: [G][rank][k*k slabs][K-step][2][N/2][8] (resident weights, walk16)
};

// -> [G][k*inner][KSTEPS][2][N][8] fp16   (inner = k d2 taps; mode 4: channel blocks of the (d2,d4) taps)
__global__ void pack_w_tc_kernel(const float *__restrict__ w, __half *__restrict__ wp, TcPackW a) {
    const int k = a.d.k;
    const long long total = (long long)(a.d.G_e / a.OG) * k * a.inner * a.KSTEPS * 2 * a.N * 8;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        long long r = idx;
        const int c8 = (int)(r % 8); r /= 8;
        int n, half, ks, rank_hi = -1;
        if (a.pairsplit) {
            const int nh = a.N / 2;
            n = (int)(r % nh); r /= nh;
            half = (int)(r % 2); r /= 2;
            ks = (int)(r % a.KSTEPS); r /= a.KSTEPS;
            if (a.pairsplit == 1) { n += (int)(r % 2) * nh; r /= 2; }
            else rank_hi = nh;
        } else {
            n = (int)(r % a.N); r /= a.N;
            half = (int)(r % 2); r /= 2;
            ks = (int)(r % a.KSTEPS); r /= a.KSTEPS;
        }
        int t2 = (int)(r % a.inner); r /= a.inner;
        const int t1 = (int)(r % k); r /= k;
        if (rank_hi > 0) { n += (int)(r % 2) * rank_hi; r /= 2; }
        int g = (int)r;
        const int kch = half * 8 + c8;
        int oe, ci, t3, t4;
        bool ok = true;
        if (a.mode == 4) {          // N = (branch group when merged, o), K-step = t3, channel = (t2,t4) tap of block t2
            const int c = t2 * 16 + kch;
            ok = c < k * k;
            t2 = c / k; t4 = c - t2 * k; t3 = ks; oe = n % 16; ci = 0;
            g = g * a.OG + n / 16;
        } else if (a.mode == 1) {          // N = (t4, o), K-step = t3, channel = ci
            t3 = ks; t4 = n / a.CoutPad; oe = n % a.CoutPad; ci = kch;
        } else {                    // mode 2: N = (t3, t4), single out channel, channel = ci
            t3 = n / k; t4 = n % k; oe = 0; ci = kch;
            ok = n < k * k;
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
struct TcPlan {
    int mode, KSTEPS, N, NG, CoutPad, P, R, nchunks, NT, npos_pad, stages, inner, OG;
    uint32_t tmem_cols;
    uint32_t data_bytes, w_bytes, stage_bytes, mask_off;
    size_t smem_bytes;
    bool ok;
};

static TcPlan tc_plan(int Cin_e, int Cout_e, int k, int d3, int d4, bool cb_mask = false, int og = 1) {
    TcPlan pl{};
    pl.ok = false;
    if (Cin_e > 16 || Cout_e > 16 || (k != 3 && k != 5)) return pl;
    pl.inner = k;
    if (Cin_e == 1 && Cout_e > 1) { pl.mode = 4; pl.CoutPad = 16; pl.NG = 1; pl.KSTEPS = k; pl.inner = (k * k + 15) / 16; }
    else if (Cout_e == 1)          { pl.mode = 2; pl.CoutPad = 1; pl.NG = k * k; pl.KSTEPS = 1; }
    else                           { pl.mode = 1; pl.CoutPad = 16; pl.NG = k; pl.KSTEPS = k; }
    pl.N = pl.mode == 2 ? ((k * k + 15) / 16) * 16 : pl.NG * 16;
    pl.OG = pl.mode == 4 ? og : 1;
    pl.N *= pl.OG;
    pl.P = d4 + k - 1;
    const int pad = k / 2;
    const int max_aoff = (pl.mode == 1 || pl.mode == 4) ? (k - 1) * pl.P : 0;
    // largest epilogue row shift
    const int max_shift = pl.mode == 4 ? 0 : pl.mode == 1 ? (k - 1) : (k - 1) * pl.P + (k - 1);
    pl.w_bytes = (uint32_t)pl.KSTEPS * 2u * pl.N * 16u;
    // budget: one CTA per SM may use all 512 TMEM columns and ~220 KB; two co-resident CTAs (whose
    // epilogue/prologue then overlap the other's MMA main loop) get 256 columns and ~108 KB each
    auto fit = [&](int R, int stages, int max_cols, size_t max_smem, TcPlan &o) {
        const int mrows = (R - 1) * o.P + d4 + max_shift;
        const int NT = (mrows + 127) / 128;
        if (NT * o.N > max_cols) return false;
        int npos = NT * 128 + max_aoff;
        const int need = (R + k - 1) * o.P;
        if (need > npos) npos = need;
        npos = ((npos + 7) / 8) * 8;
        if (npos > 16383) return false;                      // 14-bit LBO field (16 B units)
        const uint32_t data = 2u * npos * 16u;
        const uint32_t stage = ((data + o.w_bytes + 127) / 128) * 128;
        const size_t outb = o.CoutPad >= 16 ? (size_t)NT * 128 * kOutStride * sizeof(float) : (size_t)NT * 128 * o.NG * sizeof(float);
        size_t smem = (size_t)stages * stage;
        if (smem < outb) smem = outb;                        // the output plane aliases the ring
        smem = (smem + 127) / 128 * 128;
        o.mask_off = 0;
        if (cb_mask) {                                       // staging area for the chunk's CB16 ReLU mask
            o.mask_off = (uint32_t)smem;
            smem += ((size_t)2 * R * o.P * 16 + 127) / 128 * 128;
        }
        if (smem > max_smem) return false;
        o.R = R; o.NT = NT; o.npos_pad = npos; o.data_bytes = data; o.stage_bytes = stage; o.smem_bytes = smem;
        o.stages = stages;
        const int cols = NT * o.N;
        o.tmem_cols = cols <= 32 ? 32 : (cols <= 64 ? 64 : (cols <= 128 ? 128 : (cols <= 256 ? 256 : 512)));
        return true;
    };
    auto search = [&](int stages, int max_cols, size_t max_smem) {
        for (int R = d3; R >= 1; --R) {
            if (!fit(R, stages, max_cols, max_smem, pl)) continue;
            pl.nchunks = (d3 + R - 1) / R;
            const int balanced = (d3 + pl.nchunks - 1) / pl.nchunks;    // equal chunks; a smaller R always fits too
            return fit(balanced, stages, max_cols, max_smem, pl);
        }
        return false;
    };
    // For a budget, take the fewest plane chunks and then the deepest ring that still fits.
    auto best = [&](int max_cols, size_t max_smem, TcPlan &out) {
        bool found = false;
        const TcPlan keep = pl;
        for (int stages = 2; stages <= kStages; ++stages) {
            pl = keep;
            if (!search(stages, max_cols, max_smem)) continue;
            if (!found || pl.nchunks <= out.nchunks) { out = pl; found = true; }
        }
        pl = keep;
        return found;
    };
    // prefer two CTAs per SM (one's epilogue/prologue overlaps the other's main loop) unless that costs more than
    // twice the plane chunks of the one-CTA plan
    TcPlan one = pl, two = pl;
    const bool ok1 = best(512, 220 * 1024, one);
    const bool ok2 = best(256, 108 * 1024, two);
    if (ok2 && (!ok1 || two.nchunks <= 2 * one.nchunks)) { two.ok = true; return two; }
    if (ok1) { one.ok = true; return one; }
    return pl;
}

static size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

struct WalkPlan {
    bool ok;
    int N, NT, R, nchunks, npos_pad, stages;
    uint32_t stage_bytes, w_bytes, w_off, out_off;
    size_t smem;
};

// plan of conv4d_tc_walk_kernel: k live accumulator slots of NT tiles x N columns in the 512 TMEM columns
static WalkPlan walk_plan(int k, int d3, int d4) {
    WalkPlan w{};
    w.ok = false;
    if (k != 3 && k != 5) return w;
    w.N = (k * k + 15) / 16 * 16;
    const int nt_max = 512 / (k * w.N);
    const int P = d4 + k - 1, max_shift = (k - 1) * P + (k - 1);
    auto tiles = [&](int R) { return ((R - 1) * P + d4 + max_shift + 127) / 128; };
    int R = d3;
    while (R >= 1 && tiles(R) > nt_max) --R;
    if (R < 1) return w;
    w.nchunks = (d3 + R - 1) / R;
    w.R = (d3 + w.nchunks - 1) / w.nchunks;
    w.NT = tiles(w.R);
    int npos = w.NT * 128;
    if ((w.R + k - 1) * P > npos) npos = (w.R + k - 1) * P;
    npos = (npos + 7) / 8 * 8;
    if (npos > 16383) return w;
    w.npos_pad = npos;
    w.stage_bytes = ((uint32_t)(2 * npos * 16) + 127u) / 128u * 128u;
    w.w_bytes = (uint32_t)(k * k * w.N * 32);
    const uint32_t out_bytes = ((uint32_t)(w.NT * 128 * k * k * 4) + 127u) / 128u * 128u;
    const uint32_t fixed = ((w.w_bytes + 127u) / 128u * 128u) + out_bytes;
    const uint32_t budget = 220u * 1024u;
    if (fixed + 2u * w.stage_bytes > budget) return w;
    int stages = (int)((budget - fixed) / w.stage_bytes);
    w.stages = stages > kWalkStagesMax ? kWalkStagesMax : stages;
    w.w_off = (uint32_t)w.stages * w.stage_bytes;
    w.out_off = w.w_off + ((w.w_bytes + 127u) / 128u * 128u);
    w.smem = (size_t)w.out_off + out_bytes;
    if (w.smem < 120u * 1024u) w.smem = 120u * 1024u;      // one CTA per SM: each takes all 512 TMEM columns
    w.ok = true;
    return w;
}

size_t conv4d_tc_scratch_bytes(int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int symmetric) {
    const int G = symmetric ? 2 : 1;
    (void)Cin; (void)Cout;
    // a CB16 copy of a planar input (up to G blocks per batch item) + the re-laid-out weights
    const size_t act = cb16_bytes((size_t)B * G * 2, d1, d2, d3, d4, k);     // mode M4 packs up to 2 channel blocks per input
    const size_t wts = (size_t)G * k * k * k * 2 * 96 * 16;
    return align256(act) + align256(wts) + 512;
}

bool conv4d_tc_supports(int Cin_e, int Cout_e, int k, int d3, int d4) { return tc_plan(Cin_e, Cout_e, k, d3, d4).ok; }

// runs one effective convolution on the tensor cores; returns 2 when the shape/layout is outside the envelope
static int run_tc(const void *x, int x_dtype, const float *w, const float *bias, const void *mask, int mask_dtype, void *y,
                  int y_dtype, void *scratch, PackDesc d, int B, int Gx, int d1, int d2, int d3, int d4, int relu,
                  cudaStream_t st) {
    // NCNET_CB16PM = CB16 followed by its position-major twin: read as CB16, written as both
    __half *out_pm = nullptr;
    if (x_dtype == NCNET_CB16PM) x_dtype = NCNET_CB16;
    if (mask_dtype == NCNET_CB16PM) mask_dtype = NCNET_CB16;
    if (y_dtype == NCNET_CB16PM) {
        y_dtype = NCNET_CB16;
        out_pm = (__half *)((uint8_t *)y + align256(cb16_bytes((size_t)B * d.G_e, d1, d2, d3, d4, d.k)));
    }
    // both branch groups read the same single-channel input: their out-channels sit side by side in N
    const int og = (d.G_e == 2 && Gx == 1 && d.Cin_e == 1 && !d.dgrad && !mask) ? 2 : 1;
    const TcPlan pl = tc_plan(d.Cin_e, d.Cout_e, d.k, d3, d4, mask && mask_dtype == NCNET_CB16, og);
    if (!pl.ok) return 2;
    if (x_dtype == NCNET_CB16 && pl.mode == 4) return 2;   // the tap im2col needs the planar single-channel input
    if (y_dtype == NCNET_CB16 && pl.CoutPad < 16) return 2;
    if (mask && ((mask_dtype == NCNET_CB16) != (y_dtype == NCNET_CB16))) return 2;
    // many-channel layers: persistent CTA pairs walking along jA with the weights resident (conv4d_tc_walk16.cu) ...
    // (k = 5, CB16 output: a column step of k*k*k MMAs leaves the four epilogue warps time to drain a chunk; at k = 3 and
    // for planar outputs the per-plane kernel below is the faster one)
    const bool w16 = pl.mode == 1 && d.k == 5 && y_dtype == NCNET_CB16 && (!mask || mask_dtype == NCNET_CB16) &&
                     conv4d_tc_walk16_ok(d.k, d3, d4) && !getenv("NCNET_NO_W16");
    // ... else CTA pairs per output plane: the two chunks of a plane run as one cluster, one M256 MMA stream
    const int pair = w16 ? 2 : (pl.mode == 1 && pl.nchunks == 2 && pl.tmem_cols <= 256 && (pl.N / 2) % 8 == 0) ? 1 : 0;
    uint8_t *sc = (uint8_t *)scratch;
    const __half *act;
    size_t used = 0;
    if (x_dtype == NCNET_CB16) {
        act = (const __half *)x;
    } else {
        int rc = cb16_pack(x, x_dtype, sc, B * Gx, d.Cin_e, d1, d2, d3, d4, d.k, pl.mode == 4 ? 4 : 0, st);
        if (rc) return rc;
        act = (const __half *)sc;
        used = align256(cb16_bytes((size_t)B * Gx * (pl.mode == 4 ? pl.inner : 1), d1, d2, d3, d4, d.k));
    }
    __half *wp = (__half *)(sc + used);
    // single-input-channel layers writing CB16: persistent walk along iA with the d1 taps folded into N (conv4d_tc_walk1.cu)
    if (pl.mode == 4 && y_dtype == NCNET_CB16 && (!mask || mask_dtype == NCNET_CB16) && conv4d_tc_walk1_ok(d.k, og, d3, d4) &&
        !getenv("NCNET_NO_W1"))
        return conv4d_tc_walk1_run(act, w, wp, d, bias, (const __half *)mask, (__half *)y, out_pm, B, og, Gx, d1, d2, d3, d4, relu, st);
    {
        TcPackW a{};
        a.d = d; a.mode = pl.mode; a.KSTEPS = pl.KSTEPS; a.N = pl.N; a.CoutPad = pl.CoutPad; a.inner = pl.inner; a.OG = pl.OG;
        a.pairsplit = pair;
        const long long total = (long long)(d.G_e / pl.OG) * d.k * pl.inner * pl.KSTEPS * 2 * pl.N * 8;
        const int blocks = (int)((total + 255) / 256 < 1184 ? (total + 255) / 256 : 1184);
        pack_w_tc_kernel<<<blocks, 256, 0, st>>>(w, wp, a);
        NC_LAUNCH_OK("pack_w_tc_kernel");
    }
    if (w16)
        return conv4d_tc_walk16_run(act, wp, bias, (const __half *)mask, (__half *)y, out_pm, B, d.G_e, Gx, d.Cout_e, d1, d2, d3, d4, d.k,
                                    relu, st);
    if (pl.mode == 2 && !mask && (y_dtype == NCNET_F32 || y_dtype == NCNET_F16)) {
        // single output channel: walk along jA with k live output planes in TMEM (each input plane staged k times, not k*k)
        const WalkPlan wk = walk_plan(d.k, d3, d4);
        if (wk.ok) {
            WalkArgs q{};
            q.act = act; q.wp = wp; q.bias = bias; q.out = y;
            q.B = B; q.G = d.G_e; q.Gx = Gx;
            q.d1 = d1; q.d2 = d2; q.d3 = d3; q.d4 = d4; q.KS = d.k; q.P = pl.P;
            q.R = wk.R; q.nchunks = wk.nchunks; q.NT = wk.NT; q.N = wk.N; q.NG = d.k * d.k;
            q.npos_pad = wk.npos_pad; q.stages = wk.stages; q.relu = relu; q.out_f16 = y_dtype == NCNET_F16;
            q.idesc = (1u << 4) | ((uint32_t)(wk.N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
            q.stage_bytes = wk.stage_bytes; q.w_bytes = wk.w_bytes; q.out_off = wk.out_off; q.w_off = wk.w_off;
            cudaError_t e = cudaFuncSetAttribute(conv4d_tc_walk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wk.smem);
            NC_CHECK(e == cudaSuccess, "conv4d(tc,walk): smem attribute: %s", cudaGetErrorString(e));
            const long long ctas = (long long)B * d.G_e * d1 * wk.nchunks;
            NC_CHECK(ctas < (1ll << 31), "conv4d(tc,walk): grid too large");
            conv4d_tc_walk_kernel<<<(unsigned)ctas, kTcThreads, wk.smem, st>>>(q);
            NC_LAUNCH_OK("conv4d_tc_walk_kernel");
            return 0;
        }
    }
    TcArgs a{};
    a.act = act; a.wp = wp; a.bias = bias; a.mask = mask; a.out = y;
    a.B = B; a.G = d.G_e / pl.OG; a.OG = pl.OG; a.Gx = Gx; a.Cout = d.Cout_e; a.CoutPad = pl.CoutPad;
    a.d1 = d1; a.d2 = d2; a.d3 = d3; a.d4 = d4; a.KS = d.k; a.P = pl.P;
    a.R = pl.R; a.nchunks = pl.nchunks; a.NT = pl.NT; a.N = pl.N; a.NG = pl.NG; a.KSTEPS = pl.KSTEPS;
    a.inner = pl.inner; a.m4 = pl.mode == 4;
    a.npos_pad = pl.npos_pad; a.relu = relu; a.stages = pl.stages; a.tmem_cols = pl.tmem_cols;
    a.out_mode = y_dtype == NCNET_CB16 ? 2 : (y_dtype == NCNET_F16 ? 1 : 0);
    a.mask_mode = !mask ? 0 : (mask_dtype == NCNET_CB16 ? 3 : (mask_dtype == NCNET_F16 ? 2 : 1));
    a.stage_bytes = pl.stage_bytes; a.data_bytes = pl.data_bytes; a.w_bytes = pl.w_bytes; a.mask_off = pl.mask_off;
    a.out_pm = out_pm;
    // instruction descriptor: D=f32 (bits 4-5 = 1), A=B=f16 (0), both K-major, N>>3 at bit 17, M>>4 at bit 24
    a.idesc = (1u << 4) | ((uint32_t)(pl.N >> 3) << 17) | ((uint32_t)((pair ? 256 : 128) >> 4) << 24);
    a.pair = pair;
    const int pad = d.k / 2;
    for (int ks = 0; ks < kMaxKSteps; ++ks) a.aoff[ks] = ((pl.mode == 1 || pl.mode == 4) && ks < d.k) ? ks * pl.P : 0;
    for (int gi = 0; gi < kMaxGroups; ++gi) {
        int s = 0;
        if (gi < pl.NG) {
            if (pl.mode == 4) s = 0;                                    // one group, taps are in the channels / K-steps
            else if (pl.mode == 1) s = gi;                              // group = d4 tap
            else s = (gi / d.k) * pl.P + (gi % d.k);                    // mode 2: group = (d3, d4)
        }
        a.gshift[gi] = s;
    }
    cudaError_t e = pair ? cudaFuncSetAttribute(conv4d_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes)
                         : cudaFuncSetAttribute(conv4d_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes);
    NC_CHECK(e == cudaSuccess, "conv4d(tc): smem attribute: %s", cudaGetErrorString(e));
    const long long ctas = (long long)B * (d.G_e / pl.OG) * d1 * d2 * pl.nchunks;
    NC_CHECK(ctas < (1ll << 31), "conv4d(tc): grid too large");
    if (pair) {
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3((unsigned)ctas); cfg.blockDim = dim3(kTcThreads); cfg.dynamicSmemBytes = pl.smem_bytes; cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        e = cudaLaunchKernelEx(&cfg, conv4d_tc_kernel<true>, a);
        NC_CHECK(e == cudaSuccess, "conv4d(tc): cluster launch: %s", cudaGetErrorString(e));
    } else {
        conv4d_tc_kernel<false><<<(unsigned)ctas, kTcThreads, pl.smem_bytes, st>>>(a);
    }
    NC_LAUNCH_OK("conv4d_tc_kernel");
    return 0;
}

int conv4d_tc_fwd(const void *x, int x_dtype, const float *w, const float *bias, void *y, int y_dtype, void *scratch, int B,
                  int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int w_layout, int symmetric, int shared_input,
                  int relu, cudaStream_t st) {
    PackDesc d{};
    d.Cin = Cin; d.Cout = Cout; d.k = k; d.w_layout = w_layout; d.dgrad = 0; d.merge_groups = 0;
    d.G_e = symmetric ? 2 : 1; d.Cin_e = Cin; d.Cout_e = Cout;
    const int Gx = (symmetric && !shared_input) ? 2 : 1;
    return run_tc(x, x_dtype, w, bias, nullptr, 0, y, y_dtype, scratch, d, B, Gx, d1, d2, d3, d4, relu, st);
}

int conv4d_tc_dgrad(const void *gy, int gy_dtype, const float *w, const void *mask, int mask_dtype, void *gx, int gx_dtype,
                    void *scratch, int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int w_layout,
                    int symmetric, int shared_input, cudaStream_t st) {
    PackDesc d{};
    d.Cin = Cin; d.Cout = Cout; d.k = k; d.w_layout = w_layout; d.dgrad = 1;
    const int G = symmetric ? 2 : 1;
    int Gx;
    if (G == 2 && shared_input) { d.merge_groups = 1; d.G_e = 1; d.Cin_e = 2 * Cout; d.Cout_e = Cin; Gx = 1; }
    else { d.merge_groups = 0; d.G_e = G; d.Cin_e = Cout; d.Cout_e = Cin; Gx = G; }
    return run_tc(gy, gy_dtype, w, nullptr, mask, mask_dtype, gx, gx_dtype, scratch, d, B, Gx, d1, d2, d3, d4, 0, st);
}

}  // namespace ncnet
