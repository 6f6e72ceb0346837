// FeatureCorrelation (reference lib/model.py:89-120) and featureL2Norm (lib/model.py:14-17).
//
// out[b,a,t] = sum_c fa[b,c,a] * fb[b,c,t]: a batched GEMM, M=nA, N=nB, K=C, both operands
// MN-major (K strided).  This file holds the fp32 CUDA-core version (exact fp32 products,
// fp32 accumulate) used for fp32 features; one strided kernel also serves the two backward
// products.  Epilogue option: ReLU + the singleton-dim L2 norm of normalization=True.
#include "common.cuh"

namespace ncnet {

constexpr int BM = 64, BN = 64, BK = 16;

struct GemmArgs {
    const void *A; const void *B; void *C;
    int M, N, K;
    long long sam, sak, sbk, sbn;       // element strides: A(m,k), B(k,n); C is row-major [M,N]
    long long batchA, batchB, batchC;
    int relu_l2norm;
};

// 256 threads, 4x4 micro-tile.  AK / BKc: operand is contiguous along K (else along M / N).
template <typename T, bool AK, bool BKc>
__global__ void __launch_bounds__(256)
gemm_strided_kernel(GemmArgs p) {
    __shared__ float As[BK][BM + 4];
    __shared__ float Bs[BK][BN + 4];
    const int b = blockIdx.z;
    const T *A = (const T *)p.A + (size_t)b * p.batchA;
    const T *Bm = (const T *)p.B + (size_t)b * p.batchB;
    T *C = (T *)p.C + (size_t)b * p.batchC;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int tid = threadIdx.x;
    const int tm = (tid / 16) * 4, tn = (tid % 16) * 4;
    float acc[4][4] = {};
    for (int k0 = 0; k0 < p.K; k0 += BK) {
#pragma unroll
        for (int e = tid; e < BM * BK; e += 256) {
            int mm, kk;
            if (AK) { kk = e % BK; mm = e / BK; } else { mm = e % BM; kk = e / BM; }
            const int m = m0 + mm, k = k0 + kk;
            As[kk][mm] = (m < p.M && k < p.K) ? ldf(A + (long long)m * p.sam + (long long)k * p.sak) : 0.f;
        }
#pragma unroll
        for (int e = tid; e < BN * BK; e += 256) {
            int nn, kk;
            if (BKc) { kk = e % BK; nn = e / BK; } else { nn = e % BN; kk = e / BN; }
            const int n = n0 + nn, k = k0 + kk;
            Bs[kk][nn] = (n < p.N && k < p.K) ? ldf(Bm + (long long)k * p.sbk + (long long)n * p.sbn) : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            const float4 a4 = *reinterpret_cast<const float4 *>(&As[kk][tm]);
            const float4 b4 = *reinterpret_cast<const float4 *>(&Bs[kk][tn]);
            const float a[4] = {a4.x, a4.y, a4.z, a4.w}, bb[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int m = m0 + tm + i;
        if (m >= p.M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = n0 + tn + j;
            if (n >= p.N) continue;
            float v = acc[i][j];
            if (p.relu_l2norm) {
                v = rnd<T>(fmaxf(rnd<T>(v), 0.f));
                const float nrm = rnd<T>(sqrtf(rnd<T>(rnd<T>(v * v) + 1e-6f)));
                v = v / nrm;
            }
            stf(C + (size_t)m * p.N + n, v);
        }
    }
}

template <typename T>
static int launch_gemm(const GemmArgs &p, int batch, bool ak, bool bk, cudaStream_t st) {
    dim3 grid((p.N + BN - 1) / BN, (p.M + BM - 1) / BM, batch);
    NC_CHECK(grid.y <= 65535 && batch <= 65535, "corr4d: grid too large");
    if (!ak && !bk) gemm_strided_kernel<T, false, false><<<grid, 256, 0, st>>>(p);
    else if (ak && bk) gemm_strided_kernel<T, true, true><<<grid, 256, 0, st>>>(p);
    else if (ak && !bk) gemm_strided_kernel<T, true, false><<<grid, 256, 0, st>>>(p);
    else gemm_strided_kernel<T, false, true><<<grid, 256, 0, st>>>(p);
    NC_LAUNCH_OK("gemm_strided_kernel");
    return 0;
}

template <typename T>
static int corr_fwd(const void *fa, const void *fb, void *out, int B, int C, int nA, int nB, int norm, cudaStream_t st) {
    GemmArgs p{};
    p.A = fa; p.B = fb; p.C = out; p.M = nA; p.N = nB; p.K = C;
    p.sam = 1; p.sak = nA; p.sbk = nB; p.sbn = 1;
    p.batchA = (long long)C * nA; p.batchB = (long long)C * nB; p.batchC = (long long)nA * nB;
    p.relu_l2norm = norm;
    return launch_gemm<T>(p, B, false, false, st);
}

template <typename T>
static int corr_bwd(const void *go, const void *fa, const void *fb, void *gfa, void *gfb, int B, int C, int nA, int nB,
                    cudaStream_t st) {
    if (gfa) {   // gfa[c,a] = sum_t go[a,t] * fb[c,t]
        GemmArgs p{};
        p.A = fb; p.B = go; p.C = gfa; p.M = C; p.N = nA; p.K = nB;
        p.sam = nB; p.sak = 1; p.sbk = 1; p.sbn = nB;
        p.batchA = (long long)C * nB; p.batchB = (long long)nA * nB; p.batchC = (long long)C * nA;
        int rc = launch_gemm<T>(p, B, true, true, st);
        if (rc) return rc;
    }
    if (gfb) {   // gfb[c,t] = sum_a fa[c,a] * go[a,t]
        GemmArgs p{};
        p.A = fa; p.B = go; p.C = gfb; p.M = C; p.N = nB; p.K = nA;
        p.sam = nA; p.sak = 1; p.sbk = nB; p.sbn = 1;
        p.batchA = (long long)C * nA; p.batchB = (long long)nA * nB; p.batchC = (long long)C * nB;
        int rc = launch_gemm<T>(p, B, true, false, st);
        if (rc) return rc;
    }
    return 0;
}

// featureL2Norm: one thread per (b, pos), channels strided by HW (coalesced across threads)
template <typename T>
__global__ void __launch_bounds__(256)
l2norm_kernel(const T *__restrict__ x, T *__restrict__ y, int C, int HW) {
    const int b = blockIdx.y;
    const int pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= HW) return;
    const T *xb = x + (size_t)b * C * HW + pos;
    T *yb = y + (size_t)b * C * HW + pos;
    float s = 0.f;
    for (int c = 0; c < C; ++c) {
        const float v = ldf(xb + (size_t)c * HW);
        s += rnd<T>(v * v);          // elementwise square in storage precision, fp32 running sum
    }
    const float nrm = rnd<T>(sqrtf(rnd<T>(rnd<T>(s) + 1e-6f)));
    for (int c = 0; c < C; ++c) stf(yb + (size_t)c * HW, ldf(xb + (size_t)c * HW) / nrm);
}

}  // namespace ncnet

namespace ncnet {
size_t corr4d_tc_scratch_bytes(int B, int C, int nA, int nB);
int corr4d_tc_fwd(const void *fa, const void *fb, void *out, void *scratch, int B, int C, int nA, int nB, int dtype,
                  int relu_l2norm, int l2norm_inputs, cudaStream_t st);
int corr4d_pool_tc_fwd(const void *fa, const void *fb, void *value, void *const idx[4], void *scratch, int C, int hA, int wA,
                       int hB, int wB, int idx_bytes, int dtype, int relu_l2norm, int l2norm_inputs, cudaStream_t st);
}  // namespace ncnet

using namespace ncnet;

extern "C" size_t ncnet_corr4d_tc_scratch_bytes(int B, int C, int nA, int nB) {
    if (B < 1 || C < 1 || nA < 1 || nB < 1) return 0;
    return corr4d_tc_scratch_bytes(B, C, nA, nB);
}

extern "C" int ncnet_corr4d_fwd_tc(const void *fa, const void *fb, void *out, void *scratch, int B, int C, int nA, int nB,
                                   int dtype, int relu_l2norm, int l2norm_inputs, void *stream) {
    NC_CHECK(fa && fb && out && scratch, "corr4d_fwd_tc: null pointer");
    NC_CHECK(B > 0 && C > 0 && nA > 0 && nB > 0, "corr4d_fwd_tc: bad shape");
    NC_CHECK(dtype == NCNET_F32 || dtype == NCNET_F16, "corr4d_fwd_tc: unsupported dtype %d", dtype);
    return corr4d_tc_fwd(fa, fb, out, scratch, B, C, nA, nB, dtype, relu_l2norm, l2norm_inputs, (cudaStream_t)stream);
}

extern "C" int ncnet_corr4d_pool_fwd_tc(const void *fa, const void *fb, void *value, void *max_i, void *max_j, void *max_k,
                                        void *max_l, void *scratch, int C, int hA, int wA, int hB, int wB, int k_size,
                                        int idx_bytes, int dtype, int relu_l2norm, int l2norm_inputs, void *stream) {
    NC_CHECK(fa && fb && value && max_i && max_j && max_k && max_l && scratch, "corr4d_pool_fwd_tc: null pointer");
    NC_CHECK(C > 0 && hA > 0 && wA > 0 && hB > 0 && wB > 0, "corr4d_pool_fwd_tc: bad shape");
    NC_CHECK(k_size == 2, "corr4d_pool_fwd_tc: k_size %d not supported by the fused kernel (2 only; use ncnet_corr4d_fwd_tc + "
             "ncnet_maxpool4d_fwd)", k_size);
    NC_CHECK(hA % 2 == 0 && wA % 2 == 0 && hB % 2 == 0 && wB % 2 == 0, "corr4d_pool_fwd_tc: feature map sides must be even");
    NC_CHECK(idx_bytes == 1 || idx_bytes == 8, "corr4d_pool_fwd_tc: idx_bytes must be 1 or 8");
    NC_CHECK(dtype == NCNET_F32 || dtype == NCNET_F16, "corr4d_pool_fwd_tc: unsupported dtype %d", dtype);
    void *const idx[4] = {max_i, max_j, max_k, max_l};
    return corr4d_pool_tc_fwd(fa, fb, value, idx, scratch, C, hA, wA, hB, wB, idx_bytes, dtype, relu_l2norm, l2norm_inputs,
                              (cudaStream_t)stream);
}

extern "C" int ncnet_feature_l2norm(const void *x, void *y, int B, int C, int HW, int dtype, void *stream) {
    NC_CHECK(x && y, "feature_l2norm: null pointer");
    NC_CHECK(B > 0 && C > 0 && HW > 0 && B <= 65535, "feature_l2norm: bad shape");
    cudaStream_t st = (cudaStream_t)stream;
    dim3 grid((HW + 255) / 256, B);
    if (dtype == NCNET_F32) l2norm_kernel<float><<<grid, 256, 0, st>>>((const float *)x, (float *)y, C, HW);
    else if (dtype == NCNET_F16) l2norm_kernel<__half><<<grid, 256, 0, st>>>((const __half *)x, (__half *)y, C, HW);
    else return fail("feature_l2norm: unsupported dtype %d", dtype);
    NC_LAUNCH_OK("l2norm_kernel");
    return 0;
}

extern "C" int ncnet_corr4d_fwd(const void *fa, const void *fb, void *out, int B, int C, int nA, int nB, int dtype,
                                int relu_l2norm, void *stream) {
    NC_CHECK(fa && fb && out, "corr4d_fwd: null pointer");
    NC_CHECK(B > 0 && C > 0 && nA > 0 && nB > 0, "corr4d_fwd: bad shape");
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == NCNET_F32) return corr_fwd<float>(fa, fb, out, B, C, nA, nB, relu_l2norm, st);
    if (dtype == NCNET_F16) return corr_fwd<__half>(fa, fb, out, B, C, nA, nB, relu_l2norm, st);
    return fail("corr4d_fwd: unsupported dtype %d", dtype);
}

extern "C" int ncnet_corr4d_bwd(const void *grad_out, const void *fa, const void *fb, void *grad_fa, void *grad_fb,
                                int B, int C, int nA, int nB, int dtype, void *stream) {
    NC_CHECK(grad_out && fa && fb && (grad_fa || grad_fb), "corr4d_bwd: null pointer");
    NC_CHECK(B > 0 && C > 0 && nA > 0 && nB > 0, "corr4d_bwd: bad shape");
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == NCNET_F32) return corr_bwd<float>(grad_out, fa, fb, grad_fa, grad_fb, B, C, nA, nB, st);
    if (dtype == NCNET_F16) return corr_bwd<__half>(grad_out, fa, fb, grad_fa, grad_fb, B, C, nA, nB, st);
    return fail("corr4d_bwd: unsupported dtype %d", dtype);
}
