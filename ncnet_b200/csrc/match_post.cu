// What the reference does with the matches after corr_to_matches, on the device:
//   * InLoc evaluation (eval_inloc.py:151-203): both matching directions concatenated, sorted by descending score,
//     duplicate (xA,yA,xB,yB) columns removed with np.unique(axis=1) on the HOST (two D2H/H2D round trips), recentred,
//     packed into rows [xA,yA,xB,yB,score];
//   * PF-Pascal evaluation (lib/point_tnf.py:96-167, lib/eval_util.py:12-51): key-point transfer by bilinear
//     interpolation of the match grid, pixel <-> unit coordinate changes, PCK.
// All of it is index / gather work on a few thousand matches; nothing here is bandwidth- or FLOP-relevant, the point
// is that the pipeline stays on the GPU and returns the reference's values bit for bit.  Every arithmetic expression
// below is written with the rounding-explicit intrinsics (__fmul_rn, ...) in the reference's evaluation order so that
// the compiler cannot contract a*b+c into an FMA, which PyTorch's elementwise ops do not do.
#include "common.cuh"

namespace ncnet {

struct MatchCols {
    const float *xA, *yA, *xB, *yB, *score;
    int n;
};

__device__ __forceinline__ void load_match(const MatchCols &a, const MatchCols &b, int i, float v[5]) {
    const MatchCols &s = i < a.n ? a : b;
    const int j = i < a.n ? i : i - a.n;
    v[0] = s.xA[j]; v[1] = s.yA[j]; v[2] = s.xB[j]; v[3] = s.yB[j]; v[4] = s.score[j];
}

constexpr int kPostTile = 256;

// The duplicate search and the ranking compare all pairs (i, j).  Both run on a 2-D grid of (i tile, j tile) CTAs so
// that even a few thousand matches fill the GPU; a CTA stages its j tile in shared memory, every thread owns one i.
// Pass 1: dup[i] = 1 when another match has the same four coordinates and wins the reference's tie-break (the copy
// that comes first after sorting by descending score; equal scores: the lower concatenated index).
__global__ void __launch_bounds__(kPostTile)
match_dup_kernel(MatchCols a, MatchCols b, int *__restrict__ dup, float *__restrict__ out, int *__restrict__ count) {
    __shared__ float t[5][kPostTile];
    const int n = a.n + b.n;
    const int i = blockIdx.x * kPostTile + threadIdx.x;
    const int j0 = blockIdx.y * kPostTile, j = j0 + threadIdx.x;
    float me[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
    if (i < n) load_match(a, b, i, me);
    if (j < n) {
        float v[5];
        load_match(a, b, j, v);
#pragma unroll
        for (int q = 0; q < 5; ++q) t[q][threadIdx.x] = v[q];
    }
    __syncthreads();
    const int lim = min(kPostTile, n - j0);
    bool hit = false;
    if (i < n)
        for (int u = 0; u < lim; ++u) {
            if (t[0][u] != me[0]) continue;                       // most pairs differ in the first coordinate
            const bool same = t[1][u] == me[1] && t[2][u] == me[2] && t[3][u] == me[3];
            const float s = t[4][u];
            hit |= same && (s > me[4] || (s == me[4] && j0 + u < i));
        }
    if (hit) dup[i] = 1;                                          // every writer stores the same value
    if (blockIdx.y == 0 && i < n) {
#pragma unroll
        for (int q = 0; q < 5; ++q) out[(size_t)i * 5 + q] = 0.f;
    }
    if (blockIdx.y == 0 && i == 0) *count = 0;
}

__device__ __forceinline__ bool coords_less(float p0, float p1, float p2, float p3, float x0, float x1, float x2, float x3) {
    // lexicographic order of the columns of [xA; yA; xB; yB], the order np.unique(axis=1) returns
    if (p0 != x0) return p0 < x0;
    if (p1 != x1) return p1 < x1;
    if (p2 != x2) return p2 < x2;
    return p3 < x3;
}

// v * (n-1) / n + 0.5/n, the three fp32 roundings of  y*(fs*k-1)/(fs*k)+0.5/(fs*k)  (eval_inloc.py:185-194)
__device__ __forceinline__ float recentre(float v, int n) {
    if (n <= 0) return v;
    return __fadd_rn(__fdiv_rn(__fmul_rn(v, (float)(n - 1)), (float)n), (float)(0.5 / (double)n));
}

// Pass 2: pos[i] += number of kept matches of this j tile with smaller coordinates (integer atomics: order-independent).
__global__ void __launch_bounds__(kPostTile)
match_rank_kernel(MatchCols a, MatchCols b, const int *__restrict__ dup, int *__restrict__ pos) {
    __shared__ float t[4][kPostTile];
    __shared__ int tk[kPostTile];
    const int n = a.n + b.n;
    const int i = blockIdx.x * kPostTile + threadIdx.x;
    const int j0 = blockIdx.y * kPostTile, j = j0 + threadIdx.x;
    float me[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
    bool mine = false;
    if (i < n) { load_match(a, b, i, me); mine = dup[i] == 0; }
    if (j < n) {
        float v[5];
        load_match(a, b, j, v);
#pragma unroll
        for (int q = 0; q < 4; ++q) t[q][threadIdx.x] = v[q];
        tk[threadIdx.x] = dup[j] == 0;
    }
    __syncthreads();
    const int lim = min(kPostTile, n - j0);
    int cnt = 0;
    if (mine)
        for (int u = 0; u < lim; ++u)
            cnt += (tk[u] && coords_less(t[0][u], t[1][u], t[2][u], t[3][u], me[0], me[1], me[2], me[3])) ? 1 : 0;
    if (mine && cnt) atomicAdd(pos + i, cnt);
}

// Pass 3: a kept match goes to row pos[i].  dup == nullptr: single direction, nothing removed, order kept.
__global__ void __launch_bounds__(kPostTile)
match_emit_kernel(MatchCols a, MatchCols b, const int *__restrict__ dup, const int *__restrict__ pos, float *__restrict__ out,
                  int *__restrict__ count, int nyA, int nxA, int nyB, int nxB) {
    const int n = a.n + b.n;
    const int i = blockIdx.x * kPostTile + threadIdx.x;
    if (i >= n) return;
    if (!dup && i == 0) *count = n;
    if (dup && dup[i]) return;
    float me[5];
    load_match(a, b, i, me);
    float *o = out + (size_t)(dup ? pos[i] : i) * 5;
    o[0] = recentre(me[0], nxA); o[1] = recentre(me[1], nyA); o[2] = recentre(me[2], nxB); o[3] = recentre(me[3], nyB);
    o[4] = me[4];
    if (dup) atomicAdd(count, 1);          // integer: order-independent
}

// ---------------------------------------------------------------------------------------------
// key-point transfer
// ---------------------------------------------------------------------------------------------
// to_unit: (x-1-(L-1)/2)*2/(L-1)   else: x*(L-1)/2+1+(L-1)/2      (lib/point_tnf.py:6-10), L per batch item and row
__global__ void points_axis_kernel(const float *__restrict__ P, const float *__restrict__ im_size, int im_stride,
                                   float *__restrict__ out, int B, int N, int to_unit) {
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long long)B * 2 * N) return;
    const int b = (int)(e / (2 * N)), row = (int)((e / N) % 2);
    const float L = im_size[(size_t)b * im_stride + (row == 0 ? 1 : 0)];     // row 0 (x) scales with w, row 1 (y) with h
    const float x = P[e];
    const float lm1 = __fsub_rn(L, 1.f), half = __fdiv_rn(lm1, 2.f);
    float r;
    if (to_unit) r = __fdiv_rn(__fmul_rn(__fsub_rn(__fsub_rn(x, 1.f), half), 2.f), lm1);
    else r = __fadd_rn(__fadd_rn(__fdiv_rn(__fmul_rn(x, lm1), 2.f), 1.f), half);
    out[e] = r;
}

// bilinearInterpPointTnf (lib/point_tnf.py:96-148).  grid: the fs values of np.linspace(-1,1,fs) as fp32.
// Reference quirk kept: the four neighbours are looked up in the FLATTENED [B*fs*fs] match arrays without a batch
// offset, i.e. always among batch item 0's matches (the evaluation runs with batch size 1).
__global__ void bilinear_point_tnf_kernel(const float *__restrict__ xA, const float *__restrict__ yA,
                                          const float *__restrict__ xB, const float *__restrict__ yB,
                                          const float *__restrict__ grid, int fs, long long total,
                                          const float *__restrict__ tp, float *__restrict__ out, int B, int N) {
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long long)B * N) return;
    const int b = (int)(e / N), n = (int)(e % N);
    const float tx = tp[((size_t)b * 2 + 0) * N + n], ty = tp[((size_t)b * 2 + 1) * N + n];
    int cx = 0, cy = 0;
    for (int g = 0; g < fs; ++g) {
        cx += __fsub_rn(tx, grid[g]) > 0.f ? 1 : 0;
        cy += __fsub_rn(ty, grid[g]) > 0.f ? 1 : 0;
    }
    const int xm = max(cx - 1, 0), ym = max(cy - 1, 0), xp = xm + 1, yp = ym + 1;
    auto at = [&](int x, int y) {
        long long idx = (long long)y * fs + x;
        return idx < total ? idx : total - 1;        // the reference raises IndexError here; clamp instead of reading out of bounds
    };
    const long long mm = at(xm, ym), pp = at(xp, yp), pm = at(xp, ym), mp = at(xm, yp);
    auto area = [&](long long idx) { return __fmul_rn(fabsf(__fsub_rn(tx, xB[idx])), fabsf(__fsub_rn(ty, yB[idx]))); };
    const float f_pp = area(mm), f_mm = area(pp), f_mp = area(pm), f_pm = area(mp);
    const float den = __fadd_rn(__fadd_rn(__fadd_rn(f_pp, f_mm), f_mp), f_pm);
    auto mix = [&](const float *Q) {
        float s = __fadd_rn(__fmul_rn(Q[mm], f_mm), __fmul_rn(Q[pp], f_pp));
        s = __fadd_rn(s, __fmul_rn(Q[mp], f_mp));
        s = __fadd_rn(s, __fmul_rn(Q[pm], f_pm));
        return __fdiv_rn(s, den);
    };
    out[((size_t)b * 2 + 0) * N + n] = mix(xA);
    out[((size_t)b * 2 + 1) * N + n] = mix(yA);
}

// pck (lib/eval_util.py:12-25): one warp per batch item
__global__ void pck_kernel(const float *__restrict__ src, const float *__restrict__ wrp, const float *__restrict__ L_pck,
                           float alpha, float *__restrict__ out, int N) {
    const int b = blockIdx.x, lane = threadIdx.x;
    const float *sx = src + (size_t)b * 2 * N, *sy = sx + N;
    const float *wx = wrp + (size_t)b * 2 * N, *wy = wx + N;
    int npts = 0;
    for (int n = lane; n < N; n += 32) npts += (sx[n] != -1.f && sy[n] != -1.f) ? 1 : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) npts += __shfl_xor_sync(0xffffffffu, npts, o);
    const float thr = __fmul_rn(L_pck[b], alpha);
    int ok = 0;
    for (int n = lane; n < npts; n += 32) {      // the first N_pts columns, like the reference's slice
        const float dx = __fsub_rn(sx[n], wx[n]), dy = __fsub_rn(sy[n], wy[n]);
        const float d = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
        ok += d <= thr ? 1 : 0;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ok += __shfl_xor_sync(0xffffffffu, ok, o);
    if (lane == 0) out[b] = __fdiv_rn((float)ok, (float)npts);      // 0/0 = nan, as torch.mean of an empty slice
}

}  // namespace ncnet

using namespace ncnet;

extern "C" size_t ncnet_matches_unique_scratch_bytes(int n) { return n > 0 ? (((size_t)n * 8 + 255) & ~(size_t)255) : 0; }

extern "C" int ncnet_matches_sort_unique(const float *xA, const float *yA, const float *xB, const float *yB,
                                         const float *score, int n1, const float *xA2, const float *yA2,
                                         const float *xB2, const float *yB2, const float *score2, int n2,
                                         int recentre_yA, int recentre_xA, int recentre_yB, int recentre_xB,
                                         float *out, int *count, void *scratch, void *stream) {
    NC_CHECK(n1 > 0 && n2 >= 0, "matches_sort_unique: bad sizes n1=%d n2=%d", n1, n2);
    NC_CHECK(xA && yA && xB && yB && score && out && count && (scratch || n2 == 0), "matches_sort_unique: null pointer");
    NC_CHECK(n2 == 0 || (xA2 && yA2 && xB2 && yB2 && score2), "matches_sort_unique: null pointer (second direction)");
    NC_CHECK((long long)n1 + n2 <= (1 << 20), "matches_sort_unique: more than 2^20 matches (the duplicate search is quadratic)");
    cudaStream_t st = (cudaStream_t)stream;
    MatchCols a{xA, yA, xB, yB, score, n1}, b{xA2, yA2, xB2, yB2, score2, n2};
    const int n = n1 + n2, blocks = (n + kPostTile - 1) / kPostTile;
    int *dup = nullptr, *pos = nullptr;
    if (n2 > 0) {
        dup = (int *)scratch; pos = dup + n;
        cudaError_t e = cudaMemsetAsync(scratch, 0, (size_t)n * 8, st);
        NC_CHECK(e == cudaSuccess, "matches_sort_unique: memset: %s", cudaGetErrorString(e));
        NC_CHECK(blocks <= 65535, "matches_sort_unique: too many matches for the grid");
        match_dup_kernel<<<dim3(blocks, blocks), kPostTile, 0, st>>>(a, b, dup, out, count);
        NC_LAUNCH_OK("match_dup_kernel");
        match_rank_kernel<<<dim3(blocks, blocks), kPostTile, 0, st>>>(a, b, dup, pos);
        NC_LAUNCH_OK("match_rank_kernel");
    }
    match_emit_kernel<<<blocks, kPostTile, 0, st>>>(a, b, dup, pos, out, count, recentre_yA, recentre_xA, recentre_yB, recentre_xB);
    NC_LAUNCH_OK("match_emit_kernel");
    return 0;
}

extern "C" int ncnet_points_axis_transform(const float *points, const float *im_size, int im_stride, float *out, int B,
                                           int N, int to_unit, void *stream) {
    NC_CHECK(points && im_size && out, "points_axis_transform: null pointer");
    NC_CHECK(B > 0 && N > 0 && im_stride >= 2, "points_axis_transform: bad shape");
    const long long total = (long long)B * 2 * N;
    points_axis_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(points, im_size, im_stride, out, B, N,
                                                                                          to_unit);
    NC_LAUNCH_OK("points_axis_kernel");
    return 0;
}

extern "C" int ncnet_bilinear_point_tnf(const float *xA, const float *yA, const float *xB, const float *yB,
                                        const float *grid, int fs, const float *target_points_norm, float *out, int B,
                                        int N, void *stream) {
    NC_CHECK(xA && yA && xB && yB && grid && target_points_norm && out, "bilinear_point_tnf: null pointer");
    NC_CHECK(B > 0 && N > 0 && fs > 1, "bilinear_point_tnf: bad shape");
    const long long total = (long long)B * N;
    bilinear_point_tnf_kernel<<<(unsigned)((total + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        xA, yA, xB, yB, grid, fs, (long long)B * fs * fs, target_points_norm, out, B, N);
    NC_LAUNCH_OK("bilinear_point_tnf_kernel");
    return 0;
}

extern "C" int ncnet_pck(const float *source_points, const float *warped_points, const float *L_pck, float alpha,
                         float *out, int B, int N, void *stream) {
    NC_CHECK(source_points && warped_points && L_pck && out, "pck: null pointer");
    NC_CHECK(B > 0 && N > 0, "pck: bad shape");
    pck_kernel<<<B, 32, 0, (cudaStream_t)stream>>>(source_points, warped_points, L_pck, alpha, out, N);
    NC_LAUNCH_OK("pck_kernel");
    return 0;
}
