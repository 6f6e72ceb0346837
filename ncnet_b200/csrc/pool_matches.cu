// maxpool4d (reference lib/model.py:177-191) and corr_to_matches (lib/point_tnf.py:12-80).
// Both are HBM-bound index kernels: one read of the volume is the algorithmic cost.
#include "common.cuh"

namespace ncnet {

// ---------------------------------------------------------------------------------------------
// maxpool4d: one thread per pooled cell; a warp reads k^3 row segments of 32*k contiguous elements
// ---------------------------------------------------------------------------------------------
template <typename T, typename I, int K>
__global__ void __launch_bounds__(256)
maxpool4d_kernel(const T *__restrict__ x, T *__restrict__ val, I *__restrict__ mi, I *__restrict__ mj,
                 I *__restrict__ mk, I *__restrict__ ml, int h1, int h2, int h3, int h4) {
    const long long n = (long long)h1 * h2 * h3 * h4;
    const long long o = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= n) return;
    long long r = o;
    const int l = (int)(r % h4); r /= h4;
    const int kk = (int)(r % h3); r /= h3;
    const int j = (int)(r % h2); r /= h2;
    const int i = (int)r;
    const long long H4 = (long long)h4 * K, H3 = (long long)h3 * K, H2 = (long long)h2 * K;
    float best = 0.f;
    int code = -1;
#pragma unroll
    for (int a = 0; a < K; ++a)
#pragma unroll
        for (int b = 0; b < K; ++b)
#pragma unroll
            for (int c = 0; c < K; ++c) {
                const T *row = x + ((((long long)(i * K + a)) * H2 + (j * K + b)) * H3 + (kk * K + c)) * H4 + (long long)l * K;
#pragma unroll
                for (int d = 0; d < K; ++d) {
                    const float v = ldf(row + d);
                    if (code < 0 || v > best) { best = v; code = ((a * K + b) * K + c) * K + d; }
                }
            }
    stf(val + o, best);
    ml[o] = (I)(code % K);
    mk[o] = (I)((code / K) % K);
    mj[o] = (I)((code / (K * K)) % K);
    mi[o] = (I)(code / (K * K * K));
}

template <typename T, typename I>
static int launch_pool(const void *x, void *val, void *mi, void *mj, void *mk, void *ml, int h1, int h2, int h3, int h4,
                       int k, cudaStream_t st) {
    const long long n = (long long)h1 * h2 * h3 * h4;
    const unsigned grid = (unsigned)((n + 255) / 256);
#define NC_POOL(KK)                                                                                          \
    maxpool4d_kernel<T, I, KK><<<grid, 256, 0, st>>>((const T *)x, (T *)val, (I *)mi, (I *)mj, (I *)mk, (I *)ml, \
                                                       h1, h2, h3, h4)
    switch (k) {
        case 1: NC_POOL(1); break;
        case 2: NC_POOL(2); break;
        case 3: NC_POOL(3); break;
        case 4: NC_POOL(4); break;
        default: return fail("maxpool4d: k_size %d not supported (1..4)", k);
    }
#undef NC_POOL
    NC_LAUNCH_OK("maxpool4d_kernel");
    return 0;
}

// ---------------------------------------------------------------------------------------------
// corr_to_matches
// ---------------------------------------------------------------------------------------------
struct MatchArgs {
    const void *corr;
    int B, fs1, fs2, fs3, fs4, do_softmax, invert, k_size, idx_bytes;
    const void *d_iA, *d_jA, *d_iB, *d_jB;
    const float *lut_xA, *lut_yA, *lut_xB, *lut_yB;
    float *xA, *yA, *xB, *yB, *score;
    long long *iA, *jA, *iB, *jB;
};

__device__ __forceinline__ int load_delta(const void *p, size_t off, int idx_bytes) {
    return idx_bytes == 1 ? (int)((const unsigned char *)p)[off] : (int)((const long long *)p)[off];
}

// writes one match: `red` = index along the reduced axis, `own` = the cell the match belongs to
__device__ __forceinline__ void emit_match(const MatchArgs &p, int b, int own, int red, float sc) {
    int ia, ja, ib, jb;
    if (p.invert) { ia = own / p.fs2; ja = own % p.fs2; ib = red / p.fs4; jb = red % p.fs4; }
    else          { ia = red / p.fs2; ja = red % p.fs2; ib = own / p.fs4; jb = own % p.fs4; }
    if (p.d_iA) {   // relocalisation (batch 1): restore the pre-pooling resolution
        const size_t off = (((size_t)ia * p.fs2 + ja) * p.fs3 + ib) * p.fs4 + jb;
        const int dia = load_delta(p.d_iA, off, p.idx_bytes), dja = load_delta(p.d_jA, off, p.idx_bytes);
        const int dib = load_delta(p.d_iB, off, p.idx_bytes), djb = load_delta(p.d_jB, off, p.idx_bytes);
        ia = ia * p.k_size + dia; ja = ja * p.k_size + dja; ib = ib * p.k_size + dib; jb = jb * p.k_size + djb;
    }
    const size_t N = p.invert ? (size_t)p.fs1 * p.fs2 : (size_t)p.fs3 * p.fs4;
    const size_t o = (size_t)b * N + own;
    p.xA[o] = p.lut_xA[ja]; p.yA[o] = p.lut_yA[ia]; p.xB[o] = p.lut_xB[jb]; p.yB[o] = p.lut_yB[ib];
    p.score[o] = sc;
    p.iA[o] = ia; p.jA[o] = ja; p.iB[o] = ib; p.jB[o] = jb;
}

// arg-max over the strided (source) axis: 32 columns x 8 row slices per CTA, coalesced along columns
template <typename T>
__global__ void __launch_bounds__(256)
matches_over_rows_kernel(MatchArgs p) {
    const int nA = p.fs1 * p.fs2, nB = p.fs3 * p.fs4;
    const int b = blockIdx.y;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int col = blockIdx.x * 32 + tx;
    const T *x = (const T *)p.corr + (size_t)b * nA * nB;
    __shared__ float s_m[8][33], s_s[8][33];
    __shared__ unsigned long long s_k[8][33];
    const bool ok = col < nB;
    // pass 1: running max (+ running sum of exp for the soft-max); 8 independent loads per thread in flight
    constexpr int U = 8;
    float m = -INFINITY, s = 0.f;
    unsigned long long best = 0ull;
    if (ok)
        for (int a0 = ty; a0 < nA; a0 += 8 * U) {
            float v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int a = a0 + 8 * u;
                v[u] = a < nA ? ldf(x + (size_t)a * nB + col) : -INFINITY;
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int a = a0 + 8 * u;
                if (a >= nA) break;
                if (p.do_softmax) {
                    if (v[u] > m) { s = s * expf(m - v[u]); m = v[u]; }
                    s += expf(v[u] - m);
                } else {
                    const unsigned long long k = pack_key(v[u], (uint32_t)a);
                    best = k > best ? k : best;
                }
            }
        }
    float sc = 0.f;
    if (p.do_softmax) {
        s_m[ty][tx] = m; s_s[ty][tx] = s;
        __syncthreads();
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < 8; ++w) M = fmaxf(M, s_m[w][tx]);
        // exact two-step sum would need another pass; rescaled partial sums agree to fp32 rounding
        float S = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) S += (s_m[w][tx] == -INFINITY) ? 0.f : s_s[w][tx] * expf(s_m[w][tx] - M);
        // pass 2: arg-max of the soft-max values themselves (ties -> first index)
        if (ok)
            for (int a0 = ty; a0 < nA; a0 += 8 * U) {
                float v[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int a = a0 + 8 * u;
                    v[u] = a < nA ? ldf(x + (size_t)a * nB + col) : -INFINITY;
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int a = a0 + 8 * u;
                    if (a >= nA) break;
                    if (v[u] < M - (sizeof(T) == 2 ? 1e-2f : 1e-4f)) continue;      // cannot reach the maximal rounded soft-max
                    const float q = rnd<T>(expf(v[u] - M) / S);
                    const unsigned long long k = pack_key(q, (uint32_t)a);
                    best = k > best ? k : best;
                }
            }
    }
    s_k[ty][tx] = best;
    __syncthreads();
    if (ty == 0 && ok) {
#pragma unroll
        for (int w = 1; w < 8; ++w) best = s_k[w][tx] > best ? s_k[w][tx] : best;
        sc = key_val(best);
        emit_match(p, b, col, (int)key_idx(best), sc);
    }
}

// ---------------------------------------------------------------------------------------------
// The same reduction for volumes with few batch items and many rows (InLoc): one CTA per (32*VEC-column tile, row
// chunk), VEC adjacent columns per lane with one 4/8/16-byte load, per-column results combined across the row chunks
// through scratch memory.  soft-max: kernel 1 leaves (max, sum of exp) per (chunk, column), kernel 2 merges them (in
// chunk order) and takes the arg-max of the rounded soft-max values with a 64-bit atomicMax of (value, first index);
// without soft-max kernel 1 does the atomicMax directly.  A last kernel turns the keys into matches.
// ---------------------------------------------------------------------------------------------
constexpr int kSplitU = 8;          // rows in flight per warp iteration

template <typename T, int VEC, int PHASE>        // PHASE 0: raw arg-max; 1: soft-max statistics; 2: soft-max arg-max
__global__ void __launch_bounds__(256)
matches_rows_split_kernel(MatchArgs p, unsigned long long *__restrict__ keys, float2 *__restrict__ part, int rows_per_chunk) {
    const int nA = p.fs1 * p.fs2, nB = p.fs3 * p.fs4;
    const int b = blockIdx.z, chunk = blockIdx.y, R = gridDim.y;
    const int lane = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c0 = (blockIdx.x * 32 + lane) * VEC;
    const bool ok = c0 < nB;                                    // nB is a multiple of VEC
    const T *x = (const T *)p.corr + (size_t)b * nA * nB;
    const int r0 = chunk * rows_per_chunk, r1 = min(nA, r0 + rows_per_chunk);
    float M[VEC], S[VEC];
    unsigned long long best[VEC];
#pragma unroll
    for (int u = 0; u < VEC; ++u) { M[u] = -INFINITY; S[u] = 0.f; best[u] = 0ull; }
    if (PHASE == 2 && ok) {                                     // merge the chunks' statistics, in chunk order
#pragma unroll
        for (int u = 0; u < VEC; ++u) {
            float mm = -INFINITY;
            for (int q = 0; q < R; ++q) mm = fmaxf(mm, part[((size_t)b * R + q) * nB + c0 + u].x);
            float ss = 0.f;
            for (int q = 0; q < R; ++q) {
                const float2 t = part[((size_t)b * R + q) * nB + c0 + u];
                ss += (t.x == -INFINITY) ? 0.f : t.y * expf(t.x - mm);
            }
            M[u] = mm; S[u] = ss;
        }
    }
    float m[VEC], sacc[VEC], thr[VEC];
    // PHASE 2: only values within `tau` of the column maximum can reach the maximal ROUNDED soft-max value (their
    // soft-max is within e^-tau of the largest: 1e-4 is ~1700 fp32 ulps, 1e-2 is ~10 fp16 ulps), so the exp / divide /
    // round / 64-bit key work is done for those few; everything else costs one compare
    const float tau = sizeof(T) == 2 ? 1e-2f : 1e-4f;
#pragma unroll
    for (int u = 0; u < VEC; ++u) { m[u] = -INFINITY; sacc[u] = 0.f; thr[u] = M[u] - tau; }
    float bv[VEC];
    int bi[VEC];
#pragma unroll
    for (int u = 0; u < VEC; ++u) { bv[u] = -INFINITY; bi[u] = 0; }
    if (ok)
        for (int a0 = r0 + ty; a0 < r1; a0 += 8 * kSplitU) {
            float v[kSplitU][VEC];
#pragma unroll
            for (int k = 0; k < kSplitU; ++k) {
                const int a = a0 + 8 * k;
                if (a < r1) load_vec<T, VEC>(x + (size_t)a * nB + c0, v[k]);
                else {
#pragma unroll
                    for (int u = 0; u < VEC; ++u) v[k][u] = -INFINITY;
                }
            }
            if (PHASE == 1) {                     // one rescale of the running sum per kSplitU rows
#pragma unroll
                for (int u = 0; u < VEC; ++u) {
                    float mn = m[u];
#pragma unroll
                    for (int k = 0; k < kSplitU; ++k) mn = fmaxf(mn, v[k][u]);
                    float acc = sacc[u] * expf(m[u] - mn);        // first iteration: 0 * exp(-inf) = 0
#pragma unroll
                    for (int k = 0; k < kSplitU; ++k) acc += expf(v[k][u] - mn);      // rows past the end: exp(-inf) = 0
                    sacc[u] = acc; m[u] = mn;
                }
            } else {
#pragma unroll
                for (int k = 0; k < kSplitU; ++k) {
                    const int a = a0 + 8 * k;
#pragma unroll
                    for (int u = 0; u < VEC; ++u) {
                        if (PHASE == 0) {         // raw values: (value, first row) as a float / int pair
                            if (v[k][u] > bv[u]) { bv[u] = v[k][u]; bi[u] = a; }
                        } else if (v[k][u] >= thr[u]) {
                            const unsigned long long key = pack_key(rnd<T>(expf(v[k][u] - M[u]) / S[u]), (uint32_t)a);
                            best[u] = key > best[u] ? key : best[u];
                        }
                    }
                }
            }
        }
    if (PHASE == 0) {
#pragma unroll
        for (int u = 0; u < VEC; ++u) best[u] = bv[u] > -INFINITY ? pack_key(bv[u], (uint32_t)bi[u]) : 0ull;
    }
    // combine the 8 warps of the CTA
    __shared__ float s_m[8][32 * VEC + 1], s_s[8][32 * VEC + 1];
    __shared__ unsigned long long s_k[8][32 * VEC + 1];
#pragma unroll
    for (int u = 0; u < VEC; ++u) {
        if (PHASE == 1) { s_m[ty][lane * VEC + u] = m[u]; s_s[ty][lane * VEC + u] = sacc[u]; }
        else s_k[ty][lane * VEC + u] = best[u];
    }
    __syncthreads();
    for (int cc = threadIdx.x; cc < 32 * VEC; cc += 256) {
        const int col = blockIdx.x * 32 * VEC + cc;
        if (col >= nB) continue;
        if (PHASE == 1) {
            float mm = -INFINITY;
#pragma unroll
            for (int w = 0; w < 8; ++w) mm = fmaxf(mm, s_m[w][cc]);
            float ss = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) ss += (s_m[w][cc] == -INFINITY) ? 0.f : s_s[w][cc] * expf(s_m[w][cc] - mm);
            part[((size_t)b * R + chunk) * nB + col] = make_float2(mm, ss);
        } else {
            unsigned long long k = 0ull;
#pragma unroll
            for (int w = 0; w < 8; ++w) k = s_k[w][cc] > k ? s_k[w][cc] : k;
            if (k) atomicMax(keys + (size_t)b * nB + col, k);
        }
    }
}

__global__ void matches_emit_kernel(MatchArgs p, const unsigned long long *__restrict__ keys) {
    const int nB = p.fs3 * p.fs4;
    const int col = blockIdx.x * blockDim.x + threadIdx.x, b = blockIdx.y;
    if (col >= nB) return;
    const unsigned long long k = keys[(size_t)b * nB + col];
    emit_match(p, b, col, (int)key_idx(k), key_val(k));
}

struct SplitPlan { int vec, col_tiles, R, rows_per_chunk; size_t keys_bytes, total; };

static int vec_elems(uintptr_t bits, int elem, int cap) {
    const int bytes = (bits & 15) == 0 ? 16 : ((bits & 7) == 0 ? 8 : ((bits & 3) == 0 ? 4 : elem));
    const int v = bytes / elem;
    return v > cap ? cap : (v < 1 ? 1 : v);
}

// R > 1 only when the one-CTA-per-column-tile kernel would leave most SMs idle
static SplitPlan split_plan(int B, int nA, int nB, int elem) {
    SplitPlan s{};
    s.vec = vec_elems((uintptr_t)nB * elem, elem, 4);            // the base pointer is checked at launch
    s.col_tiles = (nB + 32 * s.vec - 1) / (32 * s.vec);
    int R = (2 * kNumSMs + s.col_tiles * B - 1) / (s.col_tiles * B);
    if (R > 32) R = 32;
    if (R < 1 || nA < 1024) R = 1;          // short reductions: one launch of the one-CTA-per-tile kernel is quicker
    s.rows_per_chunk = ((nA + R - 1) / R + 7) / 8 * 8;
    s.R = (nA + s.rows_per_chunk - 1) / s.rows_per_chunk;
    s.keys_bytes = ((size_t)B * nB * 8 + 255) & ~(size_t)255;
    s.total = s.keys_bytes + (size_t)B * s.R * nB * sizeof(float2);
    return s;
}

template <typename T, int VEC>
static int launch_split(const MatchArgs &p, const SplitPlan &sp, void *scratch, cudaStream_t st) {
    const int nB = p.fs3 * p.fs4;
    unsigned long long *keys = (unsigned long long *)scratch;
    float2 *part = (float2 *)((uint8_t *)scratch + sp.keys_bytes);
    cudaError_t e = cudaMemsetAsync(keys, 0, (size_t)p.B * nB * 8, st);
    NC_CHECK(e == cudaSuccess, "corr_to_matches: memset: %s", cudaGetErrorString(e));
    dim3 grid(sp.col_tiles, sp.R, p.B);
    if (p.do_softmax) {
        matches_rows_split_kernel<T, VEC, 1><<<grid, 256, 0, st>>>(p, keys, part, sp.rows_per_chunk);
        NC_LAUNCH_OK("matches_rows_split_kernel(stats)");
        matches_rows_split_kernel<T, VEC, 2><<<grid, 256, 0, st>>>(p, keys, part, sp.rows_per_chunk);
    } else {
        matches_rows_split_kernel<T, VEC, 0><<<grid, 256, 0, st>>>(p, keys, part, sp.rows_per_chunk);
    }
    NC_LAUNCH_OK("matches_rows_split_kernel");
    matches_emit_kernel<<<dim3((nB + 255) / 256, p.B), 256, 0, st>>>(p, keys);
    NC_LAUNCH_OK("matches_emit_kernel");
    return 0;
}

template <typename T>
static int launch_split_t(const MatchArgs &p, const SplitPlan &sp, void *scratch, cudaStream_t st) {
    int vec = sp.vec;
    while (vec > 1 && ((uintptr_t)p.corr % (vec * sizeof(T))) != 0) vec >>= 1;      // misaligned base pointer
    SplitPlan q = sp;
    q.vec = vec;
    q.col_tiles = (p.fs3 * p.fs4 + 32 * vec - 1) / (32 * vec);
    switch (vec) {
        case 4: return launch_split<T, 4>(p, q, scratch, st);
        case 2: return launch_split<T, 2>(p, q, scratch, st);
        default: return launch_split<T, 1>(p, q, scratch, st);
    }
}

// arg-max over the contiguous (target) axis: one warp per row
// STAGE: the warp copies its row into shared memory while it takes the maximum, the later passes read it from there
// (one DRAM read of the volume; without it the rows of all resident warps exceed L2 at InLoc sizes and are re-read).
template <typename T, int VEC, bool STAGE>
__global__ void __launch_bounds__(256)
matches_over_cols_kernel(MatchArgs p) {
    extern __shared__ __align__(16) unsigned char cols_smem[];
    const int nA = p.fs1 * p.fs2, nB = p.fs3 * p.fs4;
    const int b = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= nA) return;
    const T *x = (const T *)p.corr + ((size_t)b * nA + row) * nB;
    if (STAGE) {
        T *mine = reinterpret_cast<T *>(cols_smem) + (size_t)(threadIdx.x >> 5) * (((size_t)nB + 7) & ~(size_t)7);
        float m = -INFINITY;
        for (int c4 = lane * VEC; c4 < nB; c4 += 4 * 32 * VEC) {      // four loads in flight per lane
            float v[4][VEC];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int c = c4 + k * 32 * VEC;
                if (c < nB) load_vec<T, VEC>(x + c, v[k]);
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int c = c4 + k * 32 * VEC;
                if (c < nB) {
                    store_vec<T, VEC>(mine + c, v[k]);
#pragma unroll
                    for (int u = 0; u < VEC; ++u) m = fmaxf(m, v[k][u]);
                }
            }
        }
        __syncwarp();
        x = mine;                                  // generic pointer into shared memory from here on
        unsigned long long best = 0ull;
        if (p.do_softmax) {
            m = warp_max(m);
            float s = 0.f;
            for (int c = lane * VEC; c < nB; c += 32 * VEC) {
                float v[VEC];
                load_vec_smem<T, VEC>(x + c, v);
#pragma unroll
                for (int u = 0; u < VEC; ++u) s += expf(v[u] - m);
            }
            s = warp_sum(s);
            const float thr = m - (sizeof(T) == 2 ? 1e-2f : 1e-4f);      // see matches_rows_split_kernel
            for (int c = lane * VEC; c < nB; c += 32 * VEC) {
                float v[VEC];
                load_vec_smem<T, VEC>(x + c, v);
#pragma unroll
                for (int u = 0; u < VEC; ++u)
                    if (v[u] >= thr) {
                        const unsigned long long k = pack_key(rnd<T>(expf(v[u] - m) / s), (uint32_t)(c + u));
                        best = k > best ? k : best;
                    }
            }
        } else {
            float bv = -INFINITY;
            int bi = 0;
            for (int c = lane * VEC; c < nB; c += 32 * VEC) {
                float v[VEC];
                load_vec_smem<T, VEC>(x + c, v);
#pragma unroll
                for (int u = 0; u < VEC; ++u)
                    if (v[u] > bv) { bv = v[u]; bi = c + u; }
            }
            best = bv > -INFINITY ? pack_key(bv, (uint32_t)bi) : 0ull;
        }
        best = warp_max_u64(best);
        if (lane == 0) emit_match(p, b, row, (int)key_idx(best), key_val(best));
        return;
    }
    unsigned long long best = 0ull;
    // the row is walked in groups of VEC adjacent columns per lane (one 4/8/16-byte load; nB is a multiple of VEC)
    if (p.do_softmax) {
        float m = -INFINITY;
        for (int c = lane * VEC; c < nB; c += 32 * VEC) {
            float v[VEC];
            load_vec<T, VEC>(x + c, v);
#pragma unroll
            for (int u = 0; u < VEC; ++u) m = fmaxf(m, v[u]);
        }
        m = warp_max(m);
        float s = 0.f;
        for (int c = lane * VEC; c < nB; c += 32 * VEC) {
            float v[VEC];
            load_vec<T, VEC>(x + c, v);
#pragma unroll
            for (int u = 0; u < VEC; ++u) s += expf(v[u] - m);
        }
        s = warp_sum(s);
        for (int c = lane * VEC; c < nB; c += 32 * VEC) {
            float v[VEC];
            load_vec<T, VEC>(x + c, v);
#pragma unroll
            for (int u = 0; u < VEC; ++u) {
                const unsigned long long k = pack_key(rnd<T>(expf(v[u] - m) / s), (uint32_t)(c + u));
                best = k > best ? k : best;
            }
        }
    } else {
        for (int c = lane * VEC; c < nB; c += 32 * VEC) {
            float v[VEC];
            load_vec<T, VEC>(x + c, v);
#pragma unroll
            for (int u = 0; u < VEC; ++u) {
                const unsigned long long k = pack_key(v[u], (uint32_t)(c + u));
                best = k > best ? k : best;
            }
        }
    }
    best = warp_max_u64(best);
    if (lane == 0) emit_match(p, b, row, (int)key_idx(best), key_val(best));
}

}  // namespace ncnet

using namespace ncnet;

extern "C" int ncnet_maxpool4d_fwd(const void *x, void *value, void *max_i, void *max_j, void *max_k, void *max_l,
                                   int H1, int H2, int H3, int H4, int k, int idx_bytes, int dtype, void *stream) {
    NC_CHECK(x && value && max_i && max_j && max_k && max_l, "maxpool4d: null pointer");
    NC_CHECK(k >= 1 && H1 > 0 && H2 > 0 && H3 > 0 && H4 > 0, "maxpool4d: bad shape");
    NC_CHECK(H1 % k == 0 && H2 % k == 0 && H3 % k == 0 && H4 % k == 0,
             "maxpool4d: dims (%d,%d,%d,%d) not divisible by k_size %d", H1, H2, H3, H4, k);
    NC_CHECK(idx_bytes == 1 || idx_bytes == 8, "maxpool4d: idx_bytes must be 1 or 8");
    cudaStream_t st = (cudaStream_t)stream;
    const int h1 = H1 / k, h2 = H2 / k, h3 = H3 / k, h4 = H4 / k;
    NC_CHECK((long long)h1 * h2 * h3 * h4 < (1ll << 39), "maxpool4d: volume too large");
    if (dtype == NCNET_F32) {
        if (idx_bytes == 8) return launch_pool<float, long long>(x, value, max_i, max_j, max_k, max_l, h1, h2, h3, h4, k, st);
        return launch_pool<float, unsigned char>(x, value, max_i, max_j, max_k, max_l, h1, h2, h3, h4, k, st);
    }
    if (dtype == NCNET_F16) {
        if (idx_bytes == 8) return launch_pool<__half, long long>(x, value, max_i, max_j, max_k, max_l, h1, h2, h3, h4, k, st);
        return launch_pool<__half, unsigned char>(x, value, max_i, max_j, max_k, max_l, h1, h2, h3, h4, k, st);
    }
    return fail("maxpool4d: unsupported dtype %d", dtype);
}

extern "C" size_t ncnet_corr_to_matches_scratch_bytes(int B, int fs1, int fs2, int fs3, int fs4, int invert, int dtype) {
    if (B < 1 || fs1 < 1 || fs2 < 1 || fs3 < 1 || fs4 < 1 || invert) return 0;
    const SplitPlan sp = split_plan(B, fs1 * fs2, fs3 * fs4, dtype == NCNET_F16 ? 2 : 4);
    return sp.R > 1 ? sp.total : 0;
}

extern "C" int ncnet_corr_to_matches(const void *corr, int B, int fs1, int fs2, int fs3, int fs4, int do_softmax,
                                     int invert, int k_size, const void *delta_iA, const void *delta_jA,
                                     const void *delta_iB, const void *delta_jB, int idx_bytes, const float *lut_xA,
                                     const float *lut_yA, const float *lut_xB, const float *lut_yB, float *xA,
                                     float *yA, float *xB, float *yB, float *score, int64_t *iA, int64_t *jA,
                                     int64_t *iB, int64_t *jB, void *scratch, int dtype, void *stream) {
    NC_CHECK(corr && lut_xA && lut_yA && lut_xB && lut_yB && xA && yA && xB && yB && score && iA && jA && iB && jB,
             "corr_to_matches: null pointer");
    NC_CHECK(B > 0 && B <= 65535 && fs1 > 0 && fs2 > 0 && fs3 > 0 && fs4 > 0 && k_size >= 1, "corr_to_matches: bad shape");
    const bool has_delta = delta_iA || delta_jA || delta_iB || delta_jB;
    if (has_delta) {
        NC_CHECK(delta_iA && delta_jA && delta_iB && delta_jB, "corr_to_matches: all four delta tensors are required");
        NC_CHECK(B == 1, "corr_to_matches: relocalisation (delta4d) is batch-1 only, like the reference");
        NC_CHECK(idx_bytes == 1 || idx_bytes == 8, "corr_to_matches: idx_bytes must be 1 or 8");
    }
    MatchArgs p{};
    p.corr = corr; p.B = B; p.fs1 = fs1; p.fs2 = fs2; p.fs3 = fs3; p.fs4 = fs4;
    p.do_softmax = do_softmax; p.invert = invert; p.k_size = k_size; p.idx_bytes = idx_bytes;
    p.d_iA = delta_iA; p.d_jA = delta_jA; p.d_iB = delta_iB; p.d_jB = delta_jB;
    p.lut_xA = lut_xA; p.lut_yA = lut_yA; p.lut_xB = lut_xB; p.lut_yB = lut_yB;
    p.xA = xA; p.yA = yA; p.xB = xB; p.yB = yB; p.score = score;
    p.iA = (long long *)iA; p.jA = (long long *)jA; p.iB = (long long *)iB; p.jB = (long long *)jB;
    cudaStream_t st = (cudaStream_t)stream;
    const int nA = fs1 * fs2, nB = fs3 * fs4;
    if (dtype != NCNET_F32 && dtype != NCNET_F16) return fail("corr_to_matches: unsupported dtype %d", dtype);
    const int elem = dtype == NCNET_F16 ? 2 : 4;
    if (!invert) {
        const SplitPlan sp = split_plan(B, nA, nB, elem);
        if (sp.R > 1) {          // few batch items, many rows: split the rows over CTAs (scratch from the query above)
            NC_CHECK(scratch, "corr_to_matches: this shape needs ncnet_corr_to_matches_scratch_bytes() bytes of scratch");
            return dtype == NCNET_F32 ? launch_split_t<float>(p, sp, scratch, st) : launch_split_t<__half>(p, sp, scratch, st);
        }
        dim3 grid((nB + 31) / 32, B);
        if (dtype == NCNET_F32) matches_over_rows_kernel<float><<<grid, 256, 0, st>>>(p);
        else matches_over_rows_kernel<__half><<<grid, 256, 0, st>>>(p);
    } else {
        dim3 grid((nA + 7) / 8, B);
        const int vec = vec_elems((uintptr_t)corr | ((uintptr_t)nB * elem), elem, 8);
        // stage the rows in shared memory when 8 of them fit (rows padded to 16 bytes)
        const size_t row_b = (((size_t)nB + 7) & ~(size_t)7) * elem;
        const bool stage = 8 * row_b <= 200 * 1024;
        const size_t smem = stage ? 8 * row_b : 0;
#define NC_COLS(T, V)                                                                                                   \
    do {                                                                                                                \
        if (stage) {                                                                                                    \
            cudaError_t e__ = cudaFuncSetAttribute(matches_over_cols_kernel<T, V, true>,                                \
                                                   cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);             \
            NC_CHECK(e__ == cudaSuccess, "corr_to_matches: smem attribute: %s", cudaGetErrorString(e__));                \
            matches_over_cols_kernel<T, V, true><<<grid, 256, smem, st>>>(p);                                           \
        } else {                                                                                                        \
            matches_over_cols_kernel<T, V, false><<<grid, 256, 0, st>>>(p);                                             \
        }                                                                                                               \
    } while (0)
        if (dtype == NCNET_F32) {
            switch (vec) { case 4: NC_COLS(float, 4); break; case 2: NC_COLS(float, 2); break; default: NC_COLS(float, 1); }
        } else {
            switch (vec) { case 8: NC_COLS(__half, 8); break; case 4: NC_COLS(__half, 4); break; case 2: NC_COLS(__half, 2); break;
                           default: NC_COLS(__half, 1); }
        }
#undef NC_COLS
    }
    NC_LAUNCH_OK("corr_to_matches kernel");
    return 0;
}
