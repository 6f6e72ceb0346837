// MutualMatching forward/backward  (reference lib/model.py:155-175)
//
// The volume [B,1,fs1,fs2,fs3,fs4] is a row-major matrix [nA x nB] per batch item.
//   r[a] = max_b x[a,b]   c[b] = max_a x[a,b]
//   y = x * ((x/(r+eps)) * (x/(c+eps)))          eps = 1e-5, evaluation order as written
// HBM-bound: algorithmic traffic is one read + one write of the volume.  Pass 1 reads
// the matrix once and leaves packed (max, first-argmax) keys for rows and columns; pass 2
// re-reads it (L2-resident up to ~60 MB volumes on the 126 MB L2) and writes y.
#include "common.cuh"

namespace ncnet {

constexpr float kMMEps = 1e-5f;
constexpr int kMMThreads = 256;
constexpr int kMMSlots = 16;                         // column slots per lane in pass 1
constexpr int kMMColsPerTile = 32 * kMMSlots;        // 512 columns per CTA tile
constexpr int kMMBwdSlots = 8;                       // backward keeps the narrower 256-column tile
constexpr int kMMBwdCols = 32 * kMMBwdSlots;

template <typename T>
__device__ __forceinline__ float mm_load(const T *x, size_t off, size_t group_stride, int groups) {
    float v = ldf(x + off);
    if (groups == 2) v = rnd<T>(v + ldf(x + off + group_stride));
    return v;
}

// the maximum of a packed key without touching its index half
__device__ __forceinline__ float key_val_hi(const unsigned long long *keys, size_t i) {
    return ord2f(reinterpret_cast<const uint32_t *>(keys)[2 * i + 1]);
}

// widest vector (in elements) such that every row start and the base pointers stay aligned to it
template <typename T>
static int mm_vec_for(int nB, const void *a, const void *b) {
    const uintptr_t bits = (uintptr_t)a | (uintptr_t)b | ((uintptr_t)nB * sizeof(T));
    int bytes = (bits & 15) == 0 ? 16 : ((bits & 7) == 0 ? 8 : ((bits & 3) == 0 ? 4 : (int)sizeof(T)));
    return bytes / (int)sizeof(T);
}

// Pass 1.  grid: (col tiles of 512, row strips, B); each warp walks rows warp, warp+8, ... of the strip and keeps,
// per lane, the running (max, first row) of its 16 columns as plain float/int pairs; keys are packed only for the
// final atomicMax.  A lane owns 16/VEC groups of VEC adjacent columns and loads each group with one 4/8/16-byte load.
// All loads of a row are issued before any use (a predicated-off dependent op between them serialised them once).
template <typename T, int GROUPS, int VEC>
__global__ void __launch_bounds__(kMMThreads)
mm_max_kernel(const T *__restrict__ x, unsigned long long *__restrict__ keys, int nA, int nB, int rows_per_block) {
    constexpr int NL = kMMSlots / VEC;                 // loads per lane per row
    const int b = blockIdx.z;
    const int row0 = blockIdx.y * rows_per_block;
    const int col0 = blockIdx.x * kMMColsPerTile;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t mat = (size_t)nA * nB;
    const T *xb = x + (size_t)b * mat * GROUPS;
    unsigned long long *rowkeys = keys + (size_t)b * (nA + nB);
    unsigned long long *colkeys = rowkeys + nA;

    float cv[kMMSlots];
    int ci[kMMSlots];
#pragma unroll
    for (int s = 0; s < kMMSlots; ++s) { cv[s] = -INFINITY; ci[s] = -1; }
    const int ncols = min(kMMColsPerTile, nB - col0);
    // slot s = (load q, element u): column offset q*32*VEC + lane*VEC + u.  With VEC > 1 nB is a multiple of VEC, so a
    // group of VEC columns is either entirely inside the matrix or entirely outside.
    const int row_end = min(row0 + rows_per_block, nA);
    constexpr int RU = 2;                              // rows in flight per warp iteration
    for (int a0 = row0 + warp; a0 < row_end; a0 += RU * (kMMThreads / 32)) {
        float v[RU][kMMSlots];
#pragma unroll
        for (int r = 0; r < RU; ++r) {                 // every load of both rows is issued before any use
            const int a = a0 + r * (kMMThreads / 32);
            const size_t roff = (size_t)a * nB + col0;
#pragma unroll
            for (int q = 0; q < NL; ++q) {
                const int c = q * 32 * VEC + lane * VEC;
                float t[VEC];
                if (a < row_end && c < ncols) load_vec<T, VEC>(xb + roff + c, t);
                else {
#pragma unroll
                    for (int u = 0; u < VEC; ++u) t[u] = -INFINITY;
                }
#pragma unroll
                for (int u = 0; u < VEC; ++u) v[r][q * VEC + u] = t[u];
            }
            if (GROUPS == 2) {                         // second branch, again as one batch of loads
#pragma unroll
                for (int q = 0; q < NL; ++q) {
                    const int c = q * 32 * VEC + lane * VEC;
                    float t[VEC];
                    if (a < row_end && c < ncols) load_vec<T, VEC>(xb + mat + roff + c, t);
                    else {
#pragma unroll
                        for (int u = 0; u < VEC; ++u) t[u] = 0.f;
                    }
#pragma unroll
                    for (int u = 0; u < VEC; ++u) v[r][q * VEC + u] = rnd<T>(v[r][q * VEC + u] + t[u]);
                }
            }
        }
        float rv[RU];
        int rc[RU];
#pragma unroll
        for (int r = 0; r < RU; ++r) {
            const int a = a0 + r * (kMMThreads / 32);
            rv[r] = -INFINITY; rc[r] = 0x7fffffff;
#pragma unroll
            for (int q = 0; q < NL; ++q)
#pragma unroll
                for (int u = 0; u < VEC; ++u) {
                    const int s = q * VEC + u;
                    if (v[r][s] > cv[s]) { cv[s] = v[r][s]; ci[s] = a; }
                    if (v[r][s] > rv[r]) { rv[r] = v[r][s]; rc[r] = col0 + q * 32 * VEC + lane * VEC + u; }
                }
        }
        // warp arg-max with first-index tie-break, the RU rows' shuffle chains interleaved
        float m[RU];
        int cand[RU];
#pragma unroll
        for (int r = 0; r < RU; ++r) m[r] = rv[r];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
#pragma unroll
            for (int r = 0; r < RU; ++r) m[r] = fmaxf(m[r], __shfl_xor_sync(0xffffffffu, m[r], o));
#pragma unroll
        for (int r = 0; r < RU; ++r) cand[r] = rv[r] == m[r] ? rc[r] : 0x7fffffff;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
#pragma unroll
            for (int r = 0; r < RU; ++r) cand[r] = min(cand[r], __shfl_xor_sync(0xffffffffu, cand[r], o));
#pragma unroll
        for (int r = 0; r < RU; ++r) {
            const int a = a0 + r * (kMMThreads / 32);
            if (lane == 0 && a < row_end && cand[r] != 0x7fffffff) atomicMax(rowkeys + a, pack_key(m[r], (uint32_t)cand[r]));
        }
    }

    __shared__ unsigned long long sm[kMMThreads / 32][kMMColsPerTile];
#pragma unroll
    for (int q = 0; q < NL; ++q)
#pragma unroll
        for (int u = 0; u < VEC; ++u) {
            const int s = q * VEC + u;
            sm[warp][q * 32 * VEC + lane * VEC + u] = ci[s] >= 0 ? pack_key(cv[s], (uint32_t)ci[s]) : 0ull;
        }
    __syncthreads();
    for (int cc = threadIdx.x; cc < ncols; cc += kMMThreads) {
        unsigned long long best = 0ull;
#pragma unroll
        for (int w = 0; w < kMMThreads / 32; ++w) {
            const unsigned long long k = sm[w][cc];
            best = k > best ? k : best;
        }
        if (best) atomicMax(colkeys + col0 + cc, best);
    }
}

// Pass 1 without the arg-max (inference: nothing will ask for the backward pass).  Same tiling; the inner loop is one
// max per element for the column and one for the row, four rows in flight per warp, keys carry index 0.
template <typename T, int GROUPS, int VEC>
__global__ void __launch_bounds__(kMMThreads)
mm_max_vals_kernel(const T *__restrict__ x, unsigned long long *__restrict__ keys, int nA, int nB, int rows_per_block) {
    constexpr int NL = kMMSlots / VEC;
    constexpr int RU = 4;
    const int b = blockIdx.z;
    const int row0 = blockIdx.y * rows_per_block;
    const int col0 = blockIdx.x * kMMColsPerTile;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t mat = (size_t)nA * nB;
    const T *xb = x + (size_t)b * mat * GROUPS;
    unsigned long long *rowkeys = keys + (size_t)b * (nA + nB);
    unsigned long long *colkeys = rowkeys + nA;
    float cv[kMMSlots];
#pragma unroll
    for (int s = 0; s < kMMSlots; ++s) cv[s] = -INFINITY;
    const int ncols = min(kMMColsPerTile, nB - col0);
    const int row_end = min(row0 + rows_per_block, nA);
    for (int a0 = row0 + warp; a0 < row_end; a0 += RU * (kMMThreads / 32)) {
        float v[RU][kMMSlots];
#pragma unroll
        for (int r = 0; r < RU; ++r) {
            const int a = a0 + r * (kMMThreads / 32);
            const size_t roff = (size_t)a * nB + col0;
#pragma unroll
            for (int q = 0; q < NL; ++q) {
                const int c = q * 32 * VEC + lane * VEC;
                float t[VEC];
                if (a < row_end && c < ncols) {
                    load_vec<T, VEC>(xb + roff + c, t);
                    if (GROUPS == 2) {
                        float w[VEC];
                        load_vec<T, VEC>(xb + mat + roff + c, w);
#pragma unroll
                        for (int u = 0; u < VEC; ++u) t[u] = rnd<T>(t[u] + w[u]);
                    }
                } else {
#pragma unroll
                    for (int u = 0; u < VEC; ++u) t[u] = -INFINITY;
                }
#pragma unroll
                for (int u = 0; u < VEC; ++u) v[r][q * VEC + u] = t[u];
            }
        }
        float m[RU];
#pragma unroll
        for (int r = 0; r < RU; ++r) {
            m[r] = -INFINITY;
#pragma unroll
            for (int s = 0; s < kMMSlots; ++s) { cv[s] = fmaxf(cv[s], v[r][s]); m[r] = fmaxf(m[r], v[r][s]); }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
#pragma unroll
            for (int r = 0; r < RU; ++r) m[r] = fmaxf(m[r], __shfl_xor_sync(0xffffffffu, m[r], o));
#pragma unroll
        for (int r = 0; r < RU; ++r) {
            const int a = a0 + r * (kMMThreads / 32);
            if (lane == r && a < row_end && m[r] > -INFINITY) atomicMax(rowkeys + a, pack_key(m[r], 0u));
        }
    }
    __shared__ float sm[kMMThreads / 32][kMMColsPerTile];
#pragma unroll
    for (int q = 0; q < NL; ++q)
#pragma unroll
        for (int u = 0; u < VEC; ++u) sm[warp][q * 32 * VEC + lane * VEC + u] = cv[q * VEC + u];
    __syncthreads();
    for (int cc = threadIdx.x; cc < ncols; cc += kMMThreads) {
        float best = -INFINITY;
#pragma unroll
        for (int w = 0; w < kMMThreads / 32; ++w) best = fmaxf(best, sm[w][cc]);
        if (best > -INFINITY) atomicMax(colkeys + col0 + cc, pack_key(best, 0u));
    }
}

// Pass 2: y = x * ((x * 1/(rowmax+eps)) * (x * 1/(colmax+eps))).  One thread owns VEC adjacent columns (one 4/8/16-byte
// load and store per row) and walks kMMApplyRows rows: the column reciprocals are computed once per thread, the row
// reciprocal once per warp-row (lane i, broadcast), so the inner loop is a load, four multiplies and a store per element.
// x/(max+eps) is evaluated as x * rcp_rn(max+eps) -- one correctly rounded reciprocal per row/column instead of two
// IEEE divisions per element (which made this pass ALU-bound); the product is within 1 ulp of the true quotient.
// The blocks run through the volume BACKWARDS (last batch item, last rows first): pass 1 streamed it forwards, so
// what it read last is what the 126 MB L2 still holds, and this pass starts there.
constexpr int kMMApplyRows = 16;

template <typename T, int GROUPS, int VEC>
__global__ void __launch_bounds__(256)
mm_apply_kernel(const T *__restrict__ x, T *__restrict__ y, const unsigned long long *__restrict__ keys,
                int nA, int nB) {
    const int b = gridDim.z - 1 - blockIdx.z;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int c0 = (blockIdx.x * 32 + lane) * VEC;
    const int a0 = ((gridDim.y - 1 - blockIdx.y) * 8 + warp) * kMMApplyRows;
    const size_t mat = (size_t)nA * nB;
    const T *xb = x + (size_t)b * mat * GROUPS;
    T *yb = y + (size_t)b * mat;
    const size_t kb = (size_t)b * (nA + nB);
    const bool col_ok = c0 < nB;                 // VEC > 1: nB is a multiple of VEC, the whole group is in range
    float crcp[VEC];
#pragma unroll
    for (int u = 0; u < VEC; ++u)
        crcp[u] = col_ok ? __frcp_rn(rnd<T>(key_val_hi(keys, kb + nA + c0 + u) + kMMEps)) : 0.f;
    // row reciprocals of this warp's rows: lane i computes row a0+i, broadcast below
    float my_rrcp = 0.f;
    if (lane < kMMApplyRows && a0 + lane < nA) my_rrcp = __frcp_rn(rnd<T>(key_val_hi(keys, kb + a0 + lane) + kMMEps));
    float v[kMMApplyRows][VEC];
#pragma unroll
    for (int r = 0; r < kMMApplyRows; ++r) {       // all row loads in flight before the first use
        const int a = a0 + r;
        if (a < nA && col_ok) {
            load_vec<T, VEC>(xb + (size_t)a * nB + c0, v[r]);
            if (GROUPS == 2) {
                float w[VEC];
                load_vec<T, VEC>(xb + mat + (size_t)a * nB + c0, w);
#pragma unroll
                for (int u = 0; u < VEC; ++u) v[r][u] = rnd<T>(v[r][u] + w[u]);
            }
        }
    }
#pragma unroll
    for (int r = 0; r < kMMApplyRows; ++r) {
        const float rrcp = __shfl_sync(0xffffffffu, my_rrcp, r);
        const int a = a0 + r;
        if (a >= nA || !col_ok) continue;
        float res[VEC];
#pragma unroll
        for (int u = 0; u < VEC; ++u) {
            const float qa = rnd<T>(v[r][u] * rrcp);      // x / (max over B + eps)   (corr4d_A)
            const float qb = rnd<T>(v[r][u] * crcp[u]);   // x / (max over A + eps)   (corr4d_B)
            res[u] = v[r][u] * rnd<T>(qa * qb);
        }
        store_vec<T, VEC>(yb + (size_t)a * nB + c0, res);
    }
}

// Pass 2 for fp16 volumes on packed half2 arithmetic.  The reference's half path rounds after every operation; a
// native half multiply rounds the exact product once, which is the same value as the fp32 product of two halves
// (exact: 22 significant bits) rounded to half -- so  qa*qb  and  x*(qa*qb)  are single HMUL2s, and the two quotients
// are fp32 multiplies by the row / column reciprocal packed to half2 by one conversion: ~5 instructions per element
// instead of ~23 (this pass was issue-bound on fp16 volumes).
template <int GROUPS, int VEC>
__global__ void __launch_bounds__(256)
mm_apply_half_kernel(const __half *__restrict__ x, __half *__restrict__ y, const unsigned long long *__restrict__ keys,
                     int nA, int nB) {
    static_assert(VEC % 2 == 0, "half2 lanes");
    constexpr int NP = VEC / 2;
    const int b = gridDim.z - 1 - blockIdx.z;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int c0 = (blockIdx.x * 32 + lane) * VEC;
    const int a0 = ((gridDim.y - 1 - blockIdx.y) * 8 + warp) * kMMApplyRows;
    const size_t mat = (size_t)nA * nB;
    const __half *xb = x + (size_t)b * mat * GROUPS;
    __half *yb = y + (size_t)b * mat;
    const size_t kb = (size_t)b * (nA + nB);
    const bool col_ok = c0 < nB;
    float crcp[VEC];
#pragma unroll
    for (int u = 0; u < VEC; ++u)
        crcp[u] = col_ok ? __frcp_rn(rnd<__half>(key_val_hi(keys, kb + nA + c0 + u) + kMMEps)) : 0.f;
    float my_rrcp = 0.f;
    if (lane < kMMApplyRows && a0 + lane < nA) my_rrcp = __frcp_rn(rnd<__half>(key_val_hi(keys, kb + a0 + lane) + kMMEps));
    union Vec { uint4 q; uint2 d; unsigned w; __half2 h[NP]; };
    Vec v[kMMApplyRows];
    auto ld = [&](const __half *p, Vec &o) {
        if constexpr (VEC == 8) o.q = __ldg(reinterpret_cast<const uint4 *>(p));
        else if constexpr (VEC == 4) o.d = __ldg(reinterpret_cast<const uint2 *>(p));
        else o.w = __ldg(reinterpret_cast<const unsigned *>(p));
    };
#pragma unroll
    for (int r = 0; r < kMMApplyRows; ++r) {       // all row loads in flight before the first use
        const int a = a0 + r;
        if (a < nA && col_ok) {
            ld(xb + (size_t)a * nB + c0, v[r]);
            if (GROUPS == 2) {
                Vec w;
                ld(xb + mat + (size_t)a * nB + c0, w);
#pragma unroll
                for (int q = 0; q < NP; ++q) v[r].h[q] = __hadd2(v[r].h[q], w.h[q]);
            }
        }
    }
#pragma unroll
    for (int r = 0; r < kMMApplyRows; ++r) {
        const float rrcp = __shfl_sync(0xffffffffu, my_rrcp, r);
        const int a = a0 + r;
        if (a >= nA || !col_ok) continue;
        Vec o;
#pragma unroll
        for (int q = 0; q < NP; ++q) {
            const float2 xf = __half22float2(v[r].h[q]);
            const __half2 qa = __floats2half2_rn(xf.x * rrcp, xf.y * rrcp);                       // x / (max over B + eps)
            const __half2 qb = __floats2half2_rn(xf.x * crcp[2 * q], xf.y * crcp[2 * q + 1]);     // x / (max over A + eps)
            o.h[q] = __hmul2(v[r].h[q], __hmul2(qa, qb));
        }
        __half *dst = yb + (size_t)a * nB + c0;
        if constexpr (VEC == 8) *reinterpret_cast<uint4 *>(dst) = o.q;
        else if constexpr (VEC == 4) *reinterpret_cast<uint2 *>(dst) = o.d;
        else *reinterpret_cast<unsigned *>(dst) = o.w;
    }
}

// Row / column sums without floating-point atomics (their arrival order would make the result differ from run to
// run).  The sum kernels run on a (column tile, row strip, batch) grid; every CTA owns one (tile, strip) cell and
// writes its partial row sums to rpart[b][tile][a] and its partial column sums to cpart[b][strip][c].  Every entry
// is written exactly once, so the buffers need no clearing, and the consumers add the partials in index order.
// Layout of the `sums` scratch (floats): [B*(nA+nB)] reduced sums | rpart | cpart.
struct MMSumLayout {
    int rpb, col_tiles, strips;
    size_t rpart_off, cpart_off, total;
};

static int rows_per_block_for(int B, int nA, int nB, int tile_cols);

static MMSumLayout mm_sum_layout(int B, int nA, int nB) {
    MMSumLayout L;
    L.rpb = rows_per_block_for(B, nA, nB, kMMBwdCols);
    L.col_tiles = (nB + kMMBwdCols - 1) / kMMBwdCols;
    L.strips = (nA + L.rpb - 1) / L.rpb;
    L.rpart_off = (size_t)B * ((size_t)nA + nB);
    L.cpart_off = L.rpart_off + (size_t)B * L.col_tiles * nA;
    L.total = L.cpart_off + (size_t)B * L.strips * nB;
    return L;
}

// backward pass 1: partial sums of Sr[a] = sum_b g*y, Sc[b] = sum_a g*y  (y recomputed in fp32)
template <typename T>
__global__ void __launch_bounds__(kMMThreads)
mm_bwd_sums_kernel(const T *__restrict__ x, const T *__restrict__ gy, const unsigned long long *__restrict__ keys,
                   float *__restrict__ rpart, float *__restrict__ cpart, int nA, int nB, int rows_per_block, int groups) {
    const int b = blockIdx.z;
    const int row0 = blockIdx.y * rows_per_block;
    const int col0 = blockIdx.x * kMMBwdCols;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t mat = (size_t)nA * nB;
    const T *xb = x + (size_t)b * mat * groups;
    const T *gb = gy + (size_t)b * mat;
    const unsigned long long *rowkeys = keys + (size_t)b * (nA + nB);
    const unsigned long long *colkeys = rowkeys + nA;
    float *rsum = rpart + ((size_t)b * gridDim.x + blockIdx.x) * nA;
    float *csum = cpart + ((size_t)b * gridDim.y + blockIdx.y) * nB;

    float cden[kMMBwdSlots], cacc[kMMBwdSlots];
#pragma unroll
    for (int s = 0; s < kMMBwdSlots; ++s) {
        const int c = col0 + s * 32 + lane;
        cden[s] = c < nB ? key_val(colkeys[c]) + kMMEps : 1.f;
        cacc[s] = 0.f;
    }
    const int row_end = min(row0 + rows_per_block, nA);
    for (int a = row0 + warp; a < row_end; a += kMMThreads / 32) {
        const float rden = key_val(rowkeys[a]) + kMMEps;
        const size_t roff = (size_t)a * nB;
        float racc = 0.f;
#pragma unroll
        for (int s = 0; s < kMMBwdSlots; ++s) {
            const int c = col0 + s * 32 + lane;
            if (c < nB) {
                const float v = mm_load(xb, roff + c, mat, groups);
                const float gyv = ldf(gb + roff + c);
                const float t = gyv * (v * ((v / rden) * (v / cden[s])));
                racc += t;
                cacc[s] += t;
            }
        }
        racc = warp_sum(racc);
        if (lane == 0) rsum[a] = racc;
    }
    __shared__ float sm[kMMThreads / 32][kMMBwdCols];
#pragma unroll
    for (int s = 0; s < kMMBwdSlots; ++s) sm[warp][s * 32 + lane] = cacc[s];
    __syncthreads();
    const int c = col0 + threadIdx.x;
    if (c < nB) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < kMMThreads / 32; ++w) t += sm[w][threadIdx.x];
        csum[c] = t;
    }
}

// backward pass 2
template <typename T>
__global__ void __launch_bounds__(256)
mm_bwd_apply_kernel(const T *__restrict__ x, const T *__restrict__ gy, const unsigned long long *__restrict__ keys,
                    const float *__restrict__ rpart, const float *__restrict__ cpart, int col_tiles, int strips,
                    T *__restrict__ gx, int nA, int nB, int groups) {
    const int b = blockIdx.y;
    const size_t mat = (size_t)nA * nB;
    const size_t e0 = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (e0 >= mat) return;
    const T *xb = x + (size_t)b * mat * groups;
    const T *gb = gy + (size_t)b * mat;
    T *ob = gx + (size_t)b * mat;
    const unsigned long long *rowkeys = keys + (size_t)b * (nA + nB);
    const unsigned long long *colkeys = rowkeys + nA;
    const float *rsum = rpart + (size_t)b * col_tiles * nA;     // [col_tiles][nA] partials of this batch item
    const float *csum = cpart + (size_t)b * strips * nB;        // [strips][nB]
    int a = (int)(e0 / nB), c = (int)(e0 - (size_t)a * nB);
    unsigned long long rk = rowkeys[a];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const size_t e = e0 + i;
        if (e < mat) {
            const float rden = key_val(rk) + kMMEps;
            const unsigned long long ck = colkeys[c];
            const float cden = key_val(ck) + kMMEps;
            const float v = mm_load(xb, e, mat, groups);
            const float g = ldf(gb + e);
            const float qa = v / rden, qb = v / cden;
            // y = v*(qa*qb):  d/dv through the three occurrences of v
            float d = g * (qa * qb) + (g * v * qb) / rden + (g * v * qa) / cden;
            // flow through the maxima lands on their (first) arg-max element
            // (one element per row and per column: the partial sums are added here, in index order)
            if ((int)key_idx(rk) == c) {
                float t = 0.f;
                for (int q = 0; q < col_tiles; ++q) t += rsum[(size_t)q * nA + a];
                d -= t / rden;
            }
            if ((int)key_idx(ck) == a) {
                float t = 0.f;
                for (int q = 0; q < strips; ++q) t += csum[(size_t)q * nB + c];
                d -= t / cden;
            }
            stf(ob + e, d);
        }
        if (++c == nB) {
            c = 0;
            ++a;
            if (a < nA) rk = rowkeys[a];
        }
    }
}

static int rows_per_block_for(int B, int nA, int nB, int tile_cols) {
    const int col_tiles = (nB + tile_cols - 1) / tile_cols;
    // aim for ~12 blocks per SM: two or three CTAs are resident per SM, and with only 4 blocks per SM the last,
    // partly filled wave was a third of the run (ncu: 600 CTAs on 444 slots).  A strip is a multiple of 32 rows
    // (8 warps x up to 4 rows in flight each).
    long long want = 12ll * kNumSMs;
    long long strips = (want + (long long)B * col_tiles - 1) / ((long long)B * col_tiles);
    int rpb = (int)((nA + strips - 1) / strips);
    rpb = ((rpb + 31) / 32) * 32;
    if (rpb < 32) rpb = 32;
    if (rpb > 256) rpb = 256;
    return rpb;
}

// which: bit 0 = pass 1 (maxima -> keys), bit 1 = pass 2 (keys -> y), bit 2 = pass 1 without the arg-max indices
template <typename T>
static int mm_fwd_parts(const void *x, void *y, uint64_t *keys, int B, int nA, int nB, int groups, int which,
                        cudaStream_t st);

template <typename T>
static int mm_fwd(const void *x, void *y, uint64_t *keys, int B, int nA, int nB, int groups, cudaStream_t st) {
    return mm_fwd_parts<T>(x, y, keys, B, nA, nB, groups, 3, st);
}

template <typename T>
static int mm_fwd_parts(const void *x, void *y, uint64_t *keys, int B, int nA, int nB, int groups, int which,
                        cudaStream_t st) {
    if (which & 1) {
    cudaError_t e = cudaMemsetAsync(keys, 0, (size_t)B * (nA + nB) * sizeof(uint64_t), st);
    NC_CHECK(e == cudaSuccess, "mutual_matching: memset keys: %s", cudaGetErrorString(e));
    const int rpb = rows_per_block_for(B, nA, nB, kMMColsPerTile);
    dim3 g1((nB + kMMColsPerTile - 1) / kMMColsPerTile, (nA + rpb - 1) / rpb, B);
    const int vec1 = mm_vec_for<T>(nB, x, x);
#define NC_MM_MAX(G, V)                                                                                                    \
    do {                                                                                                                   \
        if (which & 4) mm_max_vals_kernel<T, G, V><<<g1, kMMThreads, 0, st>>>((const T *)x, (unsigned long long *)keys, nA, nB, rpb); \
        else mm_max_kernel<T, G, V><<<g1, kMMThreads, 0, st>>>((const T *)x, (unsigned long long *)keys, nA, nB, rpb);      \
    } while (0)
#define NC_MM_MAX_V(G)                                                                         \
    switch (vec1) {                                                                            \
        case 8: if constexpr (sizeof(T) == 2) { NC_MM_MAX(G, 8); break; }                      \
        case 4: NC_MM_MAX(G, 4); break;                                                        \
        case 2: NC_MM_MAX(G, 2); break;                                                        \
        default: NC_MM_MAX(G, 1); break;                                                       \
    }
    if (groups == 2) { NC_MM_MAX_V(2) } else { NC_MM_MAX_V(1) }
#undef NC_MM_MAX_V
#undef NC_MM_MAX
    NC_LAUNCH_OK("mm_max_kernel");
    }
    if (!(which & 2)) return 0;
    const int vec = mm_vec_for<T>(nB, x, y);
    dim3 g2((nB + 32 * vec - 1) / (32 * vec), (nA + 8 * kMMApplyRows - 1) / (8 * kMMApplyRows), B);
    NC_CHECK(g2.y <= 65535, "mutual_matching: too many rows");
#define NC_MM_APPLY(G, V)                                                                                                     \
    do {                                                                                                                      \
        if constexpr (sizeof(T) == 2 && (V) % 2 == 0)                                                                         \
            mm_apply_half_kernel<G, ((V) % 2 == 0 ? (V) : 2)><<<g2, 256, 0, st>>>((const __half *)x, (__half *)y, (const unsigned long long *)keys, nA, nB); \
        else                                                                                                                  \
            mm_apply_kernel<T, G, V><<<g2, 256, 0, st>>>((const T *)x, (T *)y, (const unsigned long long *)keys, nA, nB);      \
    } while (0)
#define NC_MM_APPLY_V(G)                                                                       \
    switch (vec) {                                                                             \
        case 8: if constexpr (sizeof(T) == 2) { NC_MM_APPLY(G, 8); break; }                    \
        case 4: NC_MM_APPLY(G, 4); break;                                                      \
        case 2: NC_MM_APPLY(G, 2); break;                                                      \
        default: NC_MM_APPLY(G, 1); break;                                                     \
    }
    if (groups == 2) { NC_MM_APPLY_V(2) } else { NC_MM_APPLY_V(1) }
#undef NC_MM_APPLY_V
#undef NC_MM_APPLY
    NC_LAUNCH_OK("mm_apply_kernel");
    return 0;
}

template <typename T>
static int mm_bwd(const void *x, const void *gy, const uint64_t *keys, float *sums, void *gx, int B, int nA, int nB,
                  int groups, cudaStream_t st) {
    const MMSumLayout L = mm_sum_layout(B, nA, nB);
    NC_CHECK(L.strips <= 65535, "mutual_matching_bwd: too many row strips");
    dim3 g1(L.col_tiles, L.strips, B);
    mm_bwd_sums_kernel<T><<<g1, kMMThreads, 0, st>>>((const T *)x, (const T *)gy, (const unsigned long long *)keys,
                                                     sums + L.rpart_off, sums + L.cpart_off, nA, nB, L.rpb, groups);
    NC_LAUNCH_OK("mm_bwd_sums_kernel");
    const size_t mat = (size_t)nA * nB;
    dim3 g2((unsigned)((mat + 1023) / 1024), B);
    mm_bwd_apply_kernel<T><<<g2, 256, 0, st>>>((const T *)x, (const T *)gy, (const unsigned long long *)keys,
                                               sums + L.rpart_off, sums + L.cpart_off, L.col_tiles, L.strips,
                                               (T *)gx, nA, nB, groups);
    NC_LAUNCH_OK("mm_bwd_apply_kernel");
    return 0;
}

}  // namespace ncnet

using namespace ncnet;

extern "C" size_t ncnet_mutual_matching_keys_bytes(int B, int nA, int nB) {
    return (size_t)B * ((size_t)nA + nB) * sizeof(uint64_t);
}

extern "C" size_t ncnet_mutual_matching_sums_bytes(int B, int nA, int nB) {
    if (B <= 0 || nA <= 0 || nB <= 0) return 0;
    return mm_sum_layout(B, nA, nB).total * sizeof(float);
}

extern "C" int ncnet_mutual_matching_fwd(const void *x, void *y, uint64_t *keys, int B, int nA, int nB,
                                         int in_groups, int dtype, void *stream) {
    NC_CHECK(x && y && keys, "mutual_matching_fwd: null pointer");
    NC_CHECK(B > 0 && nA > 0 && nB > 0 && B <= 65535, "mutual_matching_fwd: bad shape B=%d nA=%d nB=%d", B, nA, nB);
    NC_CHECK(in_groups == 1 || in_groups == 2, "mutual_matching_fwd: in_groups must be 1 or 2");
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == NCNET_F32) return mm_fwd<float>(x, y, keys, B, nA, nB, in_groups, st);
    if (dtype == NCNET_F16) return mm_fwd<__half>(x, y, keys, B, nA, nB, in_groups, st);
    return fail("mutual_matching_fwd: unsupported dtype %d", dtype);
}

extern "C" int ncnet_mutual_matching_fwd_infer(const void *x, void *y, uint64_t *keys, int B, int nA, int nB,
                                               int in_groups, int dtype, void *stream) {
    NC_CHECK(x && y && keys, "mutual_matching_fwd_infer: null pointer");
    NC_CHECK(B > 0 && nA > 0 && nB > 0 && B <= 65535, "mutual_matching_fwd_infer: bad shape B=%d nA=%d nB=%d", B, nA, nB);
    NC_CHECK(in_groups == 1 || in_groups == 2, "mutual_matching_fwd_infer: in_groups must be 1 or 2");
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == NCNET_F32) return mm_fwd_parts<float>(x, y, keys, B, nA, nB, in_groups, 3 | 4, st);
    if (dtype == NCNET_F16) return mm_fwd_parts<__half>(x, y, keys, B, nA, nB, in_groups, 3 | 4, st);
    return fail("mutual_matching_fwd_infer: unsupported dtype %d", dtype);
}

extern "C" int ncnet_mutual_matching_max(const void *x, uint64_t *keys, int B, int nA, int nB, int in_groups, int dtype,
                                         void *stream) {
    NC_CHECK(x && keys, "mutual_matching_max: null pointer");
    NC_CHECK(B > 0 && nA > 0 && nB > 0 && B <= 65535, "mutual_matching_max: bad shape");
    NC_CHECK(in_groups == 1 || in_groups == 2, "mutual_matching_max: in_groups must be 1 or 2");
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == NCNET_F32) return mm_fwd_parts<float>(x, nullptr, keys, B, nA, nB, in_groups, 1, st);
    if (dtype == NCNET_F16) return mm_fwd_parts<__half>(x, nullptr, keys, B, nA, nB, in_groups, 1, st);
    return fail("mutual_matching_max: unsupported dtype %d", dtype);
}

extern "C" int ncnet_mutual_matching_apply(const void *x, const uint64_t *keys, void *y, int B, int nA, int nB,
                                           int in_groups, int dtype, void *stream) {
    NC_CHECK(x && keys && y, "mutual_matching_apply: null pointer");
    NC_CHECK(B > 0 && nA > 0 && nB > 0 && B <= 65535, "mutual_matching_apply: bad shape");
    NC_CHECK(in_groups == 1 || in_groups == 2, "mutual_matching_apply: in_groups must be 1 or 2");
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == NCNET_F32) return mm_fwd_parts<float>(x, y, (uint64_t *)keys, B, nA, nB, in_groups, 2, st);
    if (dtype == NCNET_F16) return mm_fwd_parts<__half>(x, y, (uint64_t *)keys, B, nA, nB, in_groups, 2, st);
    return fail("mutual_matching_apply: unsupported dtype %d", dtype);
}

extern "C" int ncnet_mutual_matching_bwd(const void *x, const void *grad_y, const uint64_t *keys, float *sums,
                                         void *grad_x, int B, int nA, int nB, int in_groups, int dtype,
                                         void *stream) {
    NC_CHECK(x && grad_y && keys && sums && grad_x, "mutual_matching_bwd: null pointer");
    NC_CHECK(B > 0 && nA > 0 && nB > 0 && B <= 65535, "mutual_matching_bwd: bad shape");
    NC_CHECK(in_groups == 1 || in_groups == 2, "mutual_matching_bwd: in_groups must be 1 or 2");
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == NCNET_F32) return mm_bwd<float>(x, grad_y, keys, sums, grad_x, B, nA, nB, in_groups, st);
    if (dtype == NCNET_F16) return mm_bwd<__half>(x, grad_y, keys, sums, grad_x, B, nA, nB, in_groups, st);
    return fail("mutual_matching_bwd: unsupported dtype %d", dtype);
}

// ---------------------------------------------------------------------------------------------
// Allocation-free variant of the forward pass.
//
// Who needs it: a host that is not allowed to allocate device memory at the point where it runs
// the matching head -- for instance a plugin whose "enqueue" contract forbids cudaMalloc, a
// real-time loop that must not take the allocator's locks, or a process whose whole device
// budget was handed to another framework's memory pool.  Such a caller can own x and y but has
// nowhere to get the (max, arg-max) key table from, so the table lives inside the library.
//
// The table is a file-scope device array: every GPU that loads this library carries these 2 MiB
// until the process exits.  Host code cannot use the array's name as a pointer (on the host the
// symbol is only a handle); the real device address is looked up once per device with
// cudaGetSymbolAddress while that device is current, and that address is what the kernels get.
// ---------------------------------------------------------------------------------------------
namespace ncnet {

constexpr size_t kStaticKeyCapacity = (size_t)1 << 18;      // 262 144 keys = 2 MiB
static __device__ __align__(256) unsigned long long g_static_keys[kStaticKeyCapacity];

constexpr int kMaxDevices = 64;
static unsigned long long *g_static_keys_addr[kMaxDevices] = {};    // per device, filled on first use

// Resolves (once per device) the device address of g_static_keys for the current device.
static int static_keys_for_current_device(unsigned long long **out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    NC_CHECK(e == cudaSuccess, "mutual_matching_fwd_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
    NC_CHECK(dev >= 0 && dev < kMaxDevices, "mutual_matching_fwd_noalloc: device ordinal %d out of range", dev);
    unsigned long long *p = __atomic_load_n(&g_static_keys_addr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        void *sym = nullptr;
        e = cudaGetSymbolAddress(&sym, g_static_keys);      // this device's copy of the array
        NC_CHECK(e == cudaSuccess && sym, "mutual_matching_fwd_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        p = static_cast<unsigned long long *>(sym);
        __atomic_store_n(&g_static_keys_addr[dev], p, __ATOMIC_RELEASE);   // racing threads store the same value
    }
    *out = p;
    return 0;
}

}  // namespace ncnet

extern "C" int ncnet_mutual_matching_fwd_noalloc(const void *x, void *y, int B, int nA, int nB, int in_groups,
                                                 int dtype, void *stream) {
    NC_CHECK(x && y, "mutual_matching_fwd_noalloc: null pointer");
    NC_CHECK(B > 0 && nA > 0 && nB > 0 && B <= 65535, "mutual_matching_fwd_noalloc: bad shape B=%d nA=%d nB=%d", B, nA, nB);
    NC_CHECK(in_groups == 1 || in_groups == 2, "mutual_matching_fwd_noalloc: in_groups must be 1 or 2");
    const size_t need = (size_t)B * ((size_t)nA + (size_t)nB);
    NC_CHECK(need <= kStaticKeyCapacity,
             "mutual_matching_fwd_noalloc: B*(nA+nB) = %zu keys exceeds the built-in table of %zu; use "
             "ncnet_mutual_matching_fwd with caller-provided keys",
             need, kStaticKeyCapacity);
    unsigned long long *keys = nullptr;
    if (static_keys_for_current_device(&keys)) return 1;
    cudaStream_t st = (cudaStream_t)stream;
    // pass 1 clears the `need` keys it is about to use (they still hold the previous call's maxima); the arg-max
    // indices are not returned by this entry point, so the values-only pass 1 serves it
    if (dtype == NCNET_F32) return mm_fwd_parts<float>(x, y, (uint64_t *)keys, B, nA, nB, in_groups, 3 | 4, st);
    if (dtype == NCNET_F16) return mm_fwd_parts<__half>(x, y, (uint64_t *)keys, B, nA, nB, in_groups, 3 | 4, st);
    return fail("mutual_matching_fwd_noalloc: unsupported dtype %d", dtype);
}

// ---------------------------------------------------------------------------------------------
// Weak-supervision matching score (reference train.py:123-134, normalization='softmax'), fused.
//   score = ( mean_{b,t} max_a softmax_a(x[b,a,t]) + mean_{b,a} max_t softmax_t(x[b,a,t]) ) / 2
// and max softmax = 1 / sum exp(x - max), so the score needs only the row/column maxima (pass 1 above) and the
// row/column sums of exp(x - max): two reads of the volume instead of four soft-max temporaries, and the
// gradient is one elementwise pass:  d/dx[a,t] = w_t * ([a = a*_t] - p_col) / S_t + w_a * ([t = t*_a] - p_row) / S_a.
// ---------------------------------------------------------------------------------------------
namespace ncnet {

template <typename T>
__global__ void __launch_bounds__(kMMThreads)
score_sumexp_kernel(const T *__restrict__ x, const unsigned long long *__restrict__ keys, float *__restrict__ rpart,
                    float *__restrict__ cpart, int nA, int nB, int rows_per_block) {
    const int b = blockIdx.z;
    const int row0 = blockIdx.y * rows_per_block;
    const int col0 = blockIdx.x * kMMBwdCols;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t mat = (size_t)nA * nB;
    const T *xb = x + (size_t)b * mat;
    const size_t kb = (size_t)b * (nA + nB);
    float *rsum = rpart + ((size_t)b * gridDim.x + blockIdx.x) * nA;
    float *csum = cpart + ((size_t)b * gridDim.y + blockIdx.y) * nB;
    float cmax[kMMBwdSlots], cacc[kMMBwdSlots];
#pragma unroll
    for (int s = 0; s < kMMBwdSlots; ++s) {
        const int c = col0 + s * 32 + lane;
        cmax[s] = c < nB ? key_val_hi(keys, kb + nA + c) : 0.f;
        cacc[s] = 0.f;
    }
    const int row_end = min(row0 + rows_per_block, nA);
    for (int a = row0 + warp; a < row_end; a += kMMThreads / 32) {
        const float rmax = key_val_hi(keys, kb + a);
        const size_t roff = (size_t)a * nB;
        float v[kMMBwdSlots];
#pragma unroll
        for (int s = 0; s < kMMBwdSlots; ++s) {
            const int c = col0 + s * 32 + lane;
            v[s] = c < nB ? ldf(xb + roff + c) : -INFINITY;
        }
        float racc = 0.f;
#pragma unroll
        for (int s = 0; s < kMMBwdSlots; ++s) {
            if (col0 + s * 32 + lane < nB) {
                racc += expf(v[s] - rmax);
                cacc[s] += expf(v[s] - cmax[s]);
            }
        }
        racc = warp_sum(racc);
        if (lane == 0) rsum[a] = racc;
    }
    __shared__ float sm[kMMThreads / 32][kMMBwdCols];
#pragma unroll
    for (int s = 0; s < kMMBwdSlots; ++s) sm[warp][s * 32 + lane] = cacc[s];
    __syncthreads();
    const int c = col0 + threadIdx.x;
    if (c < nB) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < kMMThreads / 32; ++w) t += sm[w][threadIdx.x];
        csum[c] = t;
    }
}

// partial sums -> sums[b][0..nA) (rows) and sums[b][nA..nA+nB) (columns), added in index order
__global__ void __launch_bounds__(256)
score_sums_reduce_kernel(float *__restrict__ sums, const float *__restrict__ rpart, const float *__restrict__ cpart,
                         int B, int nA, int nB, int col_tiles, int strips) {
    const int per = nA + nB;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)B * per) return;
    const int b = (int)(i / per), r = (int)(i - (long long)b * per);
    float t = 0.f;
    if (r < nA) {
        const float *p = rpart + (size_t)b * col_tiles * nA + r;
        for (int q = 0; q < col_tiles; ++q) t += p[(size_t)q * nA];
    } else {
        const float *p = cpart + (size_t)b * strips * nB + (r - nA);
        for (int q = 0; q < strips; ++q) t += p[(size_t)q * nB];
    }
    sums[i] = t;
}

// score = (sum_rows 1/S / (B*nA) + sum_cols 1/S / (B*nB)) / 2, one block, fixed summation order
__global__ void __launch_bounds__(1024)
score_reduce_kernel(const float *__restrict__ sums, const float *__restrict__ item_weight, float *__restrict__ score,
                    int B, int nA, int nB) {
    __shared__ double sm[32];
    const int per = nA + nB;
    const double wa = 0.5 / ((double)B * nA), wb = 0.5 / ((double)B * nB);
    double acc = 0.0;
    for (long long i = threadIdx.x; i < (long long)B * per; i += blockDim.x) {
        const int r = (int)(i % per);
        if (item_weight) {        // weighted sum over batch items instead of the mean
            const double w = (double)item_weight[i / per];
            acc += w * (r < nA ? 0.5 / nA : 0.5 / nB) / (double)sums[i];
        } else {
            acc += (r < nA ? wa : wb) / (double)sums[i];
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += sm[w];
        *score = (float)t;
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
score_bwd_kernel(const T *__restrict__ x, const unsigned long long *__restrict__ keys, const float *__restrict__ sums,
                 const float *__restrict__ gscore, const float *__restrict__ item_weight, T *__restrict__ gx, int B, int nA,
                 int nB) {
    const int b = blockIdx.z;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int c = blockIdx.x * 32 + lane;
    const int a0 = (blockIdx.y * 8 + warp) * kMMApplyRows;
    const size_t mat = (size_t)nA * nB;
    const size_t kb = (size_t)b * (nA + nB);
    const float g = __ldg(gscore);
    float wa = 0.5f * g / ((float)B * nA), wb = 0.5f * g / ((float)B * nB);
    if (item_weight) {
        const float w = __ldg(item_weight + b);
        wa = 0.5f * g * w / (float)nA; wb = 0.5f * g * w / (float)nB;
    }
    float cmax = 0.f, cinv = 0.f;
    int carg = -1;
    if (c < nB) {
        const unsigned long long k = keys[kb + nA + c];
        cmax = key_val(k); carg = (int)key_idx(k); cinv = 1.f / sums[kb + nA + c];
    }
    for (int r = 0; r < kMMApplyRows; ++r) {
        const int a = a0 + r;
        if (a >= nA) break;
        const unsigned long long rk = keys[kb + a];
        const float rmax = key_val(rk), rinv = 1.f / sums[kb + a];
        const int rarg = (int)key_idx(rk);
        if (c < nB) {
            const size_t off = (size_t)b * mat + (size_t)a * nB + c;
            const float v = ldf(x + off);
            const float pc = expf(v - cmax) * cinv, pr = expf(v - rmax) * rinv;
            const float d = wb * ((a == carg ? 1.f : 0.f) - pc) * cinv + wa * ((c == rarg ? 1.f : 0.f) - pr) * rinv;
            stf(gx + off, d);
        }
    }
}

template <typename T>
static int score_fwd(const void *x, uint64_t *keys, float *sums, const float *item_weight, float *score, int B, int nA, int nB,
                     cudaStream_t st) {
    int rc = mm_fwd_parts<T>(x, nullptr, keys, B, nA, nB, 1, 1, st);      // pass 1: maxima and first arg-maxima
    if (rc) return rc;
    const MMSumLayout L = mm_sum_layout(B, nA, nB);
    NC_CHECK(L.strips <= 65535, "match_score: too many row strips");
    dim3 g1(L.col_tiles, L.strips, B);
    score_sumexp_kernel<T><<<g1, kMMThreads, 0, st>>>((const T *)x, (const unsigned long long *)keys, sums + L.rpart_off,
                                                      sums + L.cpart_off, nA, nB, L.rpb);
    NC_LAUNCH_OK("score_sumexp_kernel");
    const long long nsum = (long long)B * (nA + nB);
    score_sums_reduce_kernel<<<(unsigned)((nsum + 255) / 256), 256, 0, st>>>(sums, sums + L.rpart_off, sums + L.cpart_off, B,
                                                                             nA, nB, L.col_tiles, L.strips);
    NC_LAUNCH_OK("score_sums_reduce_kernel");
    score_reduce_kernel<<<1, 1024, 0, st>>>(sums, item_weight, score, B, nA, nB);
    NC_LAUNCH_OK("score_reduce_kernel");
    return 0;
}

}  // namespace ncnet

extern "C" int ncnet_match_score_fwd(const void *corr, uint64_t *keys, float *sums, const float *item_weight, float *score,
                                     int B, int nA, int nB, int dtype, void *stream) {
    NC_CHECK(corr && keys && sums && score, "match_score_fwd: null pointer");
    NC_CHECK(B > 0 && nA > 0 && nB > 0 && B <= 65535, "match_score_fwd: bad shape");
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == NCNET_F32) return score_fwd<float>(corr, keys, sums, item_weight, score, B, nA, nB, st);
    if (dtype == NCNET_F16) return score_fwd<__half>(corr, keys, sums, item_weight, score, B, nA, nB, st);
    return fail("match_score_fwd: unsupported dtype %d", dtype);
}

extern "C" int ncnet_match_score_bwd(const void *corr, const uint64_t *keys, const float *sums, const float *item_weight,
                                     const float *grad_score, void *grad_corr, int B, int nA, int nB, int dtype,
                                     void *stream) {
    NC_CHECK(corr && keys && sums && grad_score && grad_corr, "match_score_bwd: null pointer");
    NC_CHECK(B > 0 && nA > 0 && nB > 0 && B <= 65535, "match_score_bwd: bad shape");
    cudaStream_t st = (cudaStream_t)stream;
    dim3 grid((nB + 31) / 32, (nA + 8 * kMMApplyRows - 1) / (8 * kMMApplyRows), B);
    NC_CHECK(grid.y <= 65535, "match_score_bwd: too many rows");
    if (dtype == NCNET_F32)
        score_bwd_kernel<float><<<grid, 256, 0, st>>>((const float *)corr, (const unsigned long long *)keys, sums, grad_score, item_weight,
                                                      (float *)grad_corr, B, nA, nB);
    else if (dtype == NCNET_F16)
        score_bwd_kernel<__half><<<grid, 256, 0, st>>>((const __half *)corr, (const unsigned long long *)keys, sums, grad_score, item_weight,
                                                       (__half *)grad_corr, B, nA, nB);
    else
        return fail("match_score_bwd: unsupported dtype %d", dtype);
    NC_LAUNCH_OK("score_bwd_kernel");
    return 0;
}
