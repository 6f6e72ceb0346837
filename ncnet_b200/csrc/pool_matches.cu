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

// arg-max over the contiguous (target) axis: one warp per row
template <typename T>
__global__ void __launch_bounds__(256)
matches_over_cols_kernel(MatchArgs p) {
    const int nA = p.fs1 * p.fs2, nB = p.fs3 * p.fs4;
    const int b = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= nA) return;
    const T *x = (const T *)p.corr + ((size_t)b * nA + row) * nB;
    unsigned long long best = 0ull;
    if (p.do_softmax) {
        float m = -INFINITY;
        for (int c = lane; c < nB; c += 32) m = fmaxf(m, ldf(x + c));
        m = warp_max(m);
        float s = 0.f;
        for (int c = lane; c < nB; c += 32) s += expf(ldf(x + c) - m);
        s = warp_sum(s);
        for (int c = lane; c < nB; c += 32) {
            const float v = rnd<T>(expf(ldf(x + c) - m) / s);
            const unsigned long long k = pack_key(v, (uint32_t)c);
            best = k > best ? k : best;
        }
    } else {
        for (int c = lane; c < nB; c += 32) {
            const unsigned long long k = pack_key(ldf(x + c), (uint32_t)c);
            best = k > best ? k : best;
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

extern "C" int ncnet_corr_to_matches(const void *corr, int B, int fs1, int fs2, int fs3, int fs4, int do_softmax,
                                     int invert, int k_size, const void *delta_iA, const void *delta_jA,
                                     const void *delta_iB, const void *delta_jB, int idx_bytes, const float *lut_xA,
                                     const float *lut_yA, const float *lut_xB, const float *lut_yB, float *xA,
                                     float *yA, float *xB, float *yB, float *score, int64_t *iA, int64_t *jA,
                                     int64_t *iB, int64_t *jB, int dtype, void *stream) {
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
    if (!invert) {
        dim3 grid((nB + 31) / 32, B);
        if (dtype == NCNET_F32) matches_over_rows_kernel<float><<<grid, 256, 0, st>>>(p);
        else matches_over_rows_kernel<__half><<<grid, 256, 0, st>>>(p);
    } else {
        dim3 grid((nA + 7) / 8, B);
        if (dtype == NCNET_F32) matches_over_cols_kernel<float><<<grid, 256, 0, st>>>(p);
        else matches_over_cols_kernel<__half><<<grid, 256, 0, st>>>(p);
    }
    NC_LAUNCH_OK("corr_to_matches kernel");
    return 0;
}
