// Conv4d on CUDA cores, fp32 FMA (mode NCNET_CONV_FP32): forward, dgrad, wgrad.
// Reference semantics: lib/conv4d.py:11-51 (4D cross-correlation, stride 1, zero "same" padding).
//
// One launch per layer replaces the reference's h*k conv3d launches.  Work decomposition:
//   forward/dgrad: a CTA owns one (b, i, j) and a tile of the (k,l) plane, all output channels of
//     a 16/8/4-channel tile; it walks (cin, d1, d2), staging the halo'd input plane tile and the
//     k*k*CO weights in shared memory; each thread keeps 4 consecutive l positions x CO channels in
//     registers (FMA : LDS = 320 : 22 for k=5, CO=16).
//   wgrad: a CTA owns (group, d1, d2, a tile of input channels) and a slice of the (b,i,j) planes;
//     every thread keeps 4 out-channels x k taps in registers over the whole slice; slices are
//     reduced deterministically by the unpack kernel (no float atomics).
#include "common.cuh"
#include "conv4d_common.cuh"

namespace ncnet {

// ---------------------------------------------------------------------------------------------
// weight re-layout
// ---------------------------------------------------------------------------------------------
// packed forward/dgrad layout: [G][nct][Cin_e][k][k][k*k][CO]   (zero for channels >= Cout_e)
__global__ void pack_weights_kernel(const float *__restrict__ w, float *__restrict__ wp, PackDesc d) {
    const long long total = (long long)d.G_e * d.nct * d.Cin_e * d.k * d.k * d.k * d.k * d.CO;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        long long r = idx;
        const int oo = (int)(r % d.CO); r /= d.CO;
        const int t4 = (int)(r % d.k); r /= d.k;
        const int t3 = (int)(r % d.k); r /= d.k;
        const int t2 = (int)(r % d.k); r /= d.k;
        const int t1 = (int)(r % d.k); r /= d.k;
        const int ci = (int)(r % d.Cin_e); r /= d.Cin_e;
        const int ct = (int)(r % d.nct); r /= d.nct;
        const int g = (int)r;
        const int oe = ct * d.CO + oo;
        float v = 0.f;
        if (oe < d.Cout_e) v = w[packed_source_index(d, g, oe, ci, t1, t2, t3, t4)];
        wp[idx] = v;
    }
}

// wgrad: partial sums part[S][G][Cin][k][k][k*k][CO16*nct] -> reference layout, both groups folded
__global__ void unpack_wgrad_kernel(const float *__restrict__ part, float *__restrict__ gw, int S, int G, int Cin,
                                    int Cout, int k, int cop, int w_layout, const float *__restrict__ out_scale) {
    const int total = Cout * Cin * k * k * k * k;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    int r = idx;
    int o, c, a, b2, c3, d4;
    if (w_layout == NCNET_W_OIKKKK) {
        d4 = r % k; r /= k; c3 = r % k; r /= k; b2 = r % k; r /= k; a = r % k; r /= k; c = r % Cin; r /= Cin; o = r;
    } else {
        d4 = r % k; r /= k; c3 = r % k; r /= k; b2 = r % k; r /= k; c = r % Cin; r /= Cin; o = r % Cout; r /= Cout; a = r;
    }
    const size_t per_group = (size_t)Cin * k * k * k * k * cop;
    float acc = 0.f;
    for (int s = 0; s < S; ++s) {
        const float *p0 = part + ((size_t)s * G + 0) * per_group;
        acc += p0[(((((size_t)c * k + a) * k + b2) * k + c3) * k + d4) * cop + o];
        if (G == 2) {   // group 1 ran with the (A,B) tap pairs exchanged
            const float *p1 = part + ((size_t)s * G + 1) * per_group;
            acc += p1[(((((size_t)c * k + c3) * k + d4) * k + a) * k + b2) * cop + o];
        }
    }
    gw[idx] = out_scale ? acc * __ldg(out_scale) : acc;
}

// ---------------------------------------------------------------------------------------------
// forward / dgrad kernel, CO output channels per CTA tile
// ---------------------------------------------------------------------------------------------
struct ConvArgs {
    const void *x; void *y; const float *wp; const float *bias; const void *mask;
    int B, Cin, Cout, G, Gx;          // effective conv: x has Gx*Cin channels, y has G*Cout
    int d1, d2, d3, d4;
    int nct;                          // output-channel tiles per group
    int tiles_k, tiles_l, TY, TX;     // (k,l)-plane tiling; TX threads x 4 (or 8) positions, TY rows
    int relu;
};

template <int KS, int CO, typename T>
__global__ void __launch_bounds__(256)
conv4d_fwd_kernel(ConvArgs p) {
    constexpr int PAD = KS / 2;
    constexpr int PT = 4;                       // positions per thread along l
    constexpr int NIN = PT + 4;                 // input values per row read (two float4; KS<=5)
    extern __shared__ __align__(16) float smem[];
    const int TX = p.TX, TY = p.TY;
    const int pitch = PT * TX + 4;
    const int in_rows = TY + KS - 1;
    float *s_in = smem;                         // [in_rows][pitch]
    float *s_w = smem + in_rows * pitch;        // [KS*KS][CO]

    int bx = blockIdx.x;
    const int tl = bx % p.tiles_l; bx /= p.tiles_l;
    const int tk = bx % p.tiles_k; bx /= p.tiles_k;
    const int j = bx % p.d2; bx /= p.d2;
    const int i = bx;
    const int g = blockIdx.y / p.nct, ct = blockIdx.y % p.nct;
    const int b = blockIdx.z;
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int tid = ty * TX + tx, nthr = TX * TY;
    const int k0 = tk * TY, l0 = tl * PT * TX;
    const size_t plane = (size_t)p.d3 * p.d4;
    const size_t vol = (size_t)p.d1 * p.d2 * plane;
    const T *xg = (const T *)p.x + ((size_t)b * p.Gx * p.Cin + (size_t)(p.Gx == 1 ? 0 : g) * p.Cin) * vol;
    const float *wg = p.wp + ((size_t)g * p.nct + ct) * p.Cin * (KS * KS * KS * KS * CO);

    float acc[PT][CO];
#pragma unroll
    for (int q = 0; q < PT; ++q)
#pragma unroll
        for (int o = 0; o < CO; ++o) acc[q][o] = 0.f;

    for (int c = 0; c < p.Cin; ++c) {
        for (int t1 = 0; t1 < KS; ++t1) {
            const int ii = i + t1 - PAD;
            if (ii < 0 || ii >= p.d1) continue;
            for (int t2 = 0; t2 < KS; ++t2) {
                const int jj = j + t2 - PAD;
                if (jj < 0 || jj >= p.d2) continue;
                const T *xp = xg + (size_t)c * vol + ((size_t)ii * p.d2 + jj) * plane;
                __syncthreads();
                for (int e = tid; e < in_rows * pitch; e += nthr) {
                    const int r = e / pitch, cc = e - r * pitch;
                    const int kk = k0 + r - PAD, ll = l0 + cc - PAD;
                    float v = 0.f;
                    if (kk >= 0 && kk < p.d3 && ll >= 0 && ll < p.d4) v = ldf(xp + (size_t)kk * p.d4 + ll);
                    s_in[e] = v;
                }
                const float *wsrc = wg + (((size_t)c * KS + t1) * KS + t2) * (KS * KS * CO);
                for (int e = tid; e < KS * KS * CO; e += nthr) s_w[e] = __ldg(wsrc + e);
                __syncthreads();
#pragma unroll
                for (int t3 = 0; t3 < KS; ++t3) {
                    float in[NIN];
                    const float4 *row = reinterpret_cast<const float4 *>(s_in + (ty + t3) * pitch + PT * tx);
                    const float4 v0 = row[0], v1 = row[1];
                    in[0] = v0.x; in[1] = v0.y; in[2] = v0.z; in[3] = v0.w;
                    in[4] = v1.x; in[5] = v1.y; in[6] = v1.z; in[7] = v1.w;
#pragma unroll
                    for (int t4 = 0; t4 < KS; ++t4) {
                        const float *wrow = s_w + (t3 * KS + t4) * CO;
#pragma unroll
                        for (int o4 = 0; o4 < CO; o4 += 4) {
                            const float4 w4 = *reinterpret_cast<const float4 *>(wrow + o4);
#pragma unroll
                            for (int q = 0; q < PT; ++q) {
                                acc[q][o4 + 0] = fmaf(in[q + t4], w4.x, acc[q][o4 + 0]);
                                acc[q][o4 + 1] = fmaf(in[q + t4], w4.y, acc[q][o4 + 1]);
                                acc[q][o4 + 2] = fmaf(in[q + t4], w4.z, acc[q][o4 + 2]);
                                acc[q][o4 + 3] = fmaf(in[q + t4], w4.w, acc[q][o4 + 3]);
                            }
                        }
                    }
                }
            }
        }
    }

    const int kk = k0 + ty;
    if (kk >= p.d3) return;
    const size_t pos0 = ((size_t)i * p.d2 + j) * plane + (size_t)kk * p.d4;
#pragma unroll
    for (int o = 0; o < CO; ++o) {
        const int oe = ct * CO + o;
        if (oe >= p.Cout) break;
        const float bv = p.bias ? __ldg(p.bias + oe) : 0.f;
        const size_t chan = ((size_t)b * p.G * p.Cout + (size_t)g * p.Cout + oe) * vol + pos0;
#pragma unroll
        for (int q = 0; q < PT; ++q) {
            const int ll = l0 + PT * tx + q;
            if (ll < p.d4) {
                float v = acc[q][o] + bv;
                if (p.relu) v = fmaxf(v, 0.f);
                if (p.mask && !(ldf((const T *)p.mask + chan + ll) > 0.f)) v = 0.f;
                stf((T *)p.y + chan + ll, v);
            }
        }
    }
}

// Cout == 1 per group ("thin"): the k*k (d3,d4) weights live in registers, 8 positions per thread
template <int KS, typename T>
__global__ void __launch_bounds__(256)
conv4d_fwd_thin_kernel(ConvArgs p) {
    constexpr int PAD = KS / 2;
    constexpr int PT = 8;
    extern __shared__ __align__(16) float smem[];
    const int TX = p.TX, TY = p.TY;
    const int pitch = PT * TX + 4;
    const int in_rows = TY + KS - 1;
    float *s_in = smem;
    float *s_w = smem + in_rows * pitch;        // [KS*KS] (+pad)

    int bx = blockIdx.x;
    const int tl = bx % p.tiles_l; bx /= p.tiles_l;
    const int tk = bx % p.tiles_k; bx /= p.tiles_k;
    const int j = bx % p.d2; bx /= p.d2;
    const int i = bx;
    const int g = blockIdx.y;
    const int b = blockIdx.z;
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int tid = ty * TX + tx, nthr = TX * TY;
    const int k0 = tk * TY, l0 = tl * PT * TX;
    const size_t plane = (size_t)p.d3 * p.d4;
    const size_t vol = (size_t)p.d1 * p.d2 * plane;
    const T *xg = (const T *)p.x + ((size_t)b * p.Gx * p.Cin + (size_t)(p.Gx == 1 ? 0 : g) * p.Cin) * vol;
    const float *wg = p.wp + (size_t)g * p.Cin * (KS * KS * KS * KS);   // CO == 1, nct == 1

    float acc[PT];
#pragma unroll
    for (int q = 0; q < PT; ++q) acc[q] = 0.f;

    for (int c = 0; c < p.Cin; ++c) {
        for (int t1 = 0; t1 < KS; ++t1) {
            const int ii = i + t1 - PAD;
            if (ii < 0 || ii >= p.d1) continue;
            for (int t2 = 0; t2 < KS; ++t2) {
                const int jj = j + t2 - PAD;
                if (jj < 0 || jj >= p.d2) continue;
                const T *xp = xg + (size_t)c * vol + ((size_t)ii * p.d2 + jj) * plane;
                __syncthreads();
                for (int e = tid; e < in_rows * pitch; e += nthr) {
                    const int r = e / pitch, cc = e - r * pitch;
                    const int kk = k0 + r - PAD, ll = l0 + cc - PAD;
                    float v = 0.f;
                    if (kk >= 0 && kk < p.d3 && ll >= 0 && ll < p.d4) v = ldf(xp + (size_t)kk * p.d4 + ll);
                    s_in[e] = v;
                }
                const float *wsrc = wg + (((size_t)c * KS + t1) * KS + t2) * (KS * KS);
                for (int e = tid; e < KS * KS; e += nthr) s_w[e] = __ldg(wsrc + e);
                __syncthreads();
                float w[KS * KS];
#pragma unroll
                for (int e = 0; e < KS * KS; ++e) w[e] = s_w[e];
#pragma unroll
                for (int t3 = 0; t3 < KS; ++t3) {
                    float in[PT + 4];
                    const float4 *row = reinterpret_cast<const float4 *>(s_in + (ty + t3) * pitch + PT * tx);
                    const float4 v0 = row[0], v1 = row[1], v2 = row[2];
                    in[0] = v0.x; in[1] = v0.y; in[2] = v0.z; in[3] = v0.w;
                    in[4] = v1.x; in[5] = v1.y; in[6] = v1.z; in[7] = v1.w;
                    in[8] = v2.x; in[9] = v2.y; in[10] = v2.z; in[11] = v2.w;
#pragma unroll
                    for (int t4 = 0; t4 < KS; ++t4)
#pragma unroll
                        for (int q = 0; q < PT; ++q) acc[q] = fmaf(in[q + t4], w[t3 * KS + t4], acc[q]);
                }
            }
        }
    }
    const int kk = k0 + ty;
    if (kk >= p.d3) return;
    const float bv = p.bias ? __ldg(p.bias) : 0.f;
    const size_t chan = ((size_t)b * p.G + g) * vol + ((size_t)i * p.d2 + j) * plane + (size_t)kk * p.d4;
#pragma unroll
    for (int q = 0; q < PT; ++q) {
        const int ll = l0 + PT * tx + q;
        if (ll < p.d4) {
            float v = acc[q] + bv;
            if (p.relu) v = fmaxf(v, 0.f);
            if (p.mask && !(ldf((const T *)p.mask + chan + ll) > 0.f)) v = 0.f;
            stf((T *)p.y + chan + ll, v);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// wgrad
// ---------------------------------------------------------------------------------------------
struct WgradArgs {
    const void *x; const void *gy; float *part;
    int B, Cin, Cout, G, Gx;
    int d1, d2, d3, d4;
    int S;            // slices of the (b,i,j) plane list
    int nctile;       // input-channel tiles
    int TK;           // rows of the (k,l) plane per stage
    int cop;          // padded Cout (multiple of 4, the packed innermost dim)
    int ksplit;       // threads sharing one accumulator set (rows interleaved)
};

// CT input channels per CTA; each thread: 4 out-channels x KS taps along d4, fixed (c, d3)
template <int KS, int CT, typename T>
__global__ void __launch_bounds__(256)
conv4d_wgrad_kernel(WgradArgs p) {
    constexpr int PAD = KS / 2;
    extern __shared__ __align__(16) float smem[];
    const int OQ = p.cop / 4;
    const int npos = p.TK * p.d4;
    const int xpitch = p.d4 + KS - 1;
    const int xrows = p.TK + KS - 1;
    float4 *s_g = reinterpret_cast<float4 *>(smem);              // [OQ][npos] float4 (4 out channels)
    float *s_x = smem + (size_t)OQ * npos * 4;                    // [CT][xrows][xpitch]

    int by = blockIdx.y;
    const int ctile = by % p.nctile; by /= p.nctile;
    const int t2 = by % KS; by /= KS;
    const int t1 = by % KS; by /= KS;
    const int g = by;
    const int s = blockIdx.x;

    const int tid = threadIdx.x, nthr = blockDim.x;
    // thread -> (ks, t3, c, oq), oq fastest
    int r = tid;
    const int oq = r % OQ; r /= OQ;
    const int cl = r % CT; r /= CT;
    const int t3 = r % KS; r /= KS;
    const int ks = r;
    const bool active = ks < p.ksplit;

    const size_t plane = (size_t)p.d3 * p.d4;
    const size_t vol = (size_t)p.d1 * p.d2 * plane;
    const int nplanes = p.B * p.d1 * p.d2;
    const int per = (nplanes + p.S - 1) / p.S;
    const int pbeg = s * per, pend = min(nplanes, pbeg + per);
    const int ktiles = (p.d3 + p.TK - 1) / p.TK;

    float acc[4][KS];
#pragma unroll
    for (int o = 0; o < 4; ++o)
#pragma unroll
        for (int t = 0; t < KS; ++t) acc[o][t] = 0.f;

    for (int pl = pbeg; pl < pend; ++pl) {
        const int b = pl / (p.d1 * p.d2);
        const int ij = pl - b * (p.d1 * p.d2);
        const int i = ij / p.d2, j = ij - i * p.d2;
        const int ii = i + t1 - PAD, jj = j + t2 - PAD;
        if (ii < 0 || ii >= p.d1 || jj < 0 || jj >= p.d2) continue;
        const T *gyb = (const T *)p.gy + ((size_t)b * p.G * p.Cout + (size_t)g * p.Cout) * vol +
                       ((size_t)i * p.d2 + j) * plane;
        const T *xb = (const T *)p.x + ((size_t)b * p.Gx * p.Cin + (size_t)(p.Gx == 1 ? 0 : g) * p.Cin) * vol +
                      ((size_t)ii * p.d2 + jj) * plane;
        for (int kt = 0; kt < ktiles; ++kt) {
            const int k0 = kt * p.TK;
            const int rows = min(p.TK, p.d3 - k0);
            __syncthreads();
            // grad_y tile -> s_g[oq][pos].{x,y,z,w}
            for (int e = tid; e < p.cop * npos; e += nthr) {
                const int o = e / npos, pos = e - o * npos;
                const int rr = pos / p.d4;
                float v = 0.f;
                if (o < p.Cout && rr < rows) v = ldf(gyb + (size_t)o * vol + (size_t)k0 * p.d4 + pos);
                reinterpret_cast<float *>(s_g)[((size_t)(o >> 2) * npos + pos) * 4 + (o & 3)] = v;
            }
            for (int e = tid; e < CT * xrows * xpitch; e += nthr) {
                const int c = e / (xrows * xpitch);
                const int rem = e - c * (xrows * xpitch);
                const int rr = rem / xpitch, cc = rem - rr * xpitch;
                const int kk = k0 + rr - PAD, ll = cc - PAD;
                const int ch = ctile * CT + c;
                float v = 0.f;
                if (ch < p.Cin && kk >= 0 && kk < p.d3 && ll >= 0 && ll < p.d4)
                    v = ldf(xb + (size_t)ch * vol + (size_t)kk * p.d4 + ll);
                s_x[e] = v;
            }
            __syncthreads();
            if (active) {
                for (int rr = ks; rr < rows; rr += p.ksplit) {
                    const float4 *gp = s_g + (size_t)oq * npos + (size_t)rr * p.d4;
                    const float *xr = s_x + ((size_t)cl * xrows + rr + t3) * xpitch;
                    float win[KS];
#pragma unroll
                    for (int t = 0; t < KS - 1; ++t) win[t + 1] = xr[t];
                    for (int l = 0; l < p.d4; ++l) {
#pragma unroll
                        for (int t = 0; t < KS - 1; ++t) win[t] = win[t + 1];
                        win[KS - 1] = xr[l + KS - 1];
                        const float4 g4 = gp[l];
#pragma unroll
                        for (int t = 0; t < KS; ++t) {
                            acc[0][t] = fmaf(g4.x, win[t], acc[0][t]);
                            acc[1][t] = fmaf(g4.y, win[t], acc[1][t]);
                            acc[2][t] = fmaf(g4.z, win[t], acc[2][t]);
                            acc[3][t] = fmaf(g4.w, win[t], acc[3][t]);
                        }
                    }
                }
            }
        }
    }
    // reduce the ksplit copies through shared memory, then store this slice's partial
    __syncthreads();
    float *red = smem;                                  // [nthr][4*KS]
    if (tid < nthr) {
#pragma unroll
        for (int o = 0; o < 4; ++o)
#pragma unroll
            for (int t = 0; t < KS; ++t) red[(size_t)tid * (4 * KS) + o * KS + t] = active ? acc[o][t] : 0.f;
    }
    __syncthreads();
    const int group_threads = OQ * CT * KS;             // threads with ks == 0
    if (tid < group_threads) {
        const int ch = ctile * CT + cl;
        if (ch < p.Cin) {
            const size_t per_group = (size_t)p.Cin * KS * KS * KS * KS * p.cop;
            float *dst = p.part + ((size_t)s * p.G + g) * per_group +
                         ((((size_t)ch * KS + t1) * KS + t2) * KS + t3) * KS * p.cop;
#pragma unroll
            for (int o = 0; o < 4; ++o)
#pragma unroll
                for (int t = 0; t < KS; ++t) {
                    float v = 0.f;
                    for (int q = 0; q < p.ksplit; ++q) v += red[(size_t)(tid + q * group_threads) * (4 * KS) + o * KS + t];
                    dst[(size_t)t * p.cop + oq * 4 + o] = v;
                }
        }
    }
}

// bias gradient: db[o] = sum over batch, groups and positions of grad_y
template <typename T>
__global__ void __launch_bounds__(256)
bias_grad_kernel(const T *__restrict__ gy, float *__restrict__ db, int Cout, int G, size_t vol, int chunks) {
    const int chan = blockIdx.y;                 // over B*G*Cout
    const int o = chan % Cout;
    const T *src = gy + (size_t)chan * vol;
    const size_t per = (vol + chunks - 1) / chunks;
    const size_t beg = (size_t)blockIdx.x * per, end = min(vol, beg + per);
    float a = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;       // four loads in flight per thread
    size_t e = beg + threadIdx.x;
    for (; e + 3 * blockDim.x < end; e += 4 * blockDim.x) {
        a += ldf(src + e); a1 += ldf(src + e + blockDim.x); a2 += ldf(src + e + 2 * blockDim.x); a3 += ldf(src + e + 3 * blockDim.x);
    }
    for (; e < end; e += blockDim.x) a += ldf(src + e);
    a = warp_sum((a + a1) + (a2 + a3));
    __shared__ float sm[8];
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int w = 0; w < 8; ++w) t += sm[w];
        // partial of (chunk, b, g) for channel o:  part[(chunk*B*G + bg)*Cout + o]
        db[((size_t)blockIdx.x * (gridDim.y / Cout) + chan / Cout) * Cout + o] = t;
    }
}

// db[o] = sum over entries of part[e*W + o], one warp per channel, fixed order
__global__ void bias_reduce_kernel(const float *__restrict__ part, float *__restrict__ db, int Cout, int entries, int W,
                                   const float *__restrict__ out_scale) {
    const int o = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (o >= Cout) return;
    float t = 0.f;
    for (int e = lane; e < entries; e += 32) t += part[(size_t)e * W + o];
    t = warp_sum(t);
    if (lane == 0) db[o] = out_scale ? t * __ldg(out_scale) : t;
}

int bias_reduce(const float *part, float *db, int Cout, int entries, int W, const float *out_scale, cudaStream_t st) {
    NC_CHECK(Cout <= 32, "bias gradient: more than 32 output channels");
    bias_reduce_kernel<<<1, 32 * Cout, 0, st>>>(part, db, Cout, entries, W, out_scale);
    NC_LAUNCH_OK("bias_reduce_kernel");
    return 0;
}

template <typename T>
__global__ void relu_mask_bcast_kernel(const T *__restrict__ y, const T *__restrict__ g, T *__restrict__ gz, int G,
                                       long long n, const float *__restrict__ scale) {
    const int b = blockIdx.y;
    const float sc = scale ? __ldg(scale) : 1.f;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x) {
        const float gv = ldf(g + (size_t)b * n + e) * sc;
        for (int q = 0; q < G; ++q) {
            const size_t off = ((size_t)b * G + q) * n + e;
            stf(gz + off, ldf(y + off) > 0.f ? gv : 0.f);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
static int pick_co(int cout) { return cout == 1 ? 1 : (cout <= 4 ? 4 : (cout <= 8 ? 8 : 16)); }

size_t conv4d_fp32_scratch_bytes(int Cin, int Cout, int k, int symmetric) {
    // large enough for forward (Cin->Cout) and dgrad (G*Cout->Cin merged or per-group Cout->Cin)
    const int G = symmetric ? 2 : 1;
    auto bytes = [&](int ci, int co, int g) {
        const int CO = pick_co(co);
        const int nct = (co + CO - 1) / CO;
        return (size_t)g * nct * ci * k * k * k * k * CO * sizeof(float);
    };
    size_t a = bytes(Cin, Cout, G);
    size_t b = bytes(Cout, Cin, G);
    size_t c = bytes(G * Cout, Cin, 1);
    size_t m = a > b ? a : b;
    return (m > c ? m : c) + 256;
}

template <int KS, typename T>
static int launch_conv(ConvArgs a, int CO, cudaStream_t st) {
    // plane tiling: TX threads x PT positions along l, TY rows along k, <= 256 threads
    const int PT = (CO == 1) ? 8 : 4;
    const int max_tx = 8;
    int tiles_l = (a.d4 + PT * max_tx - 1) / (PT * max_tx);
    int TX = ((a.d4 + tiles_l - 1) / tiles_l + PT - 1) / PT;
    int max_ty = 256 / TX;
    int tiles_k = (a.d3 + max_ty - 1) / max_ty;
    int TY = (a.d3 + tiles_k - 1) / tiles_k;
    a.tiles_l = tiles_l; a.tiles_k = tiles_k; a.TX = TX; a.TY = TY;
    const long long gx = (long long)a.d1 * a.d2 * tiles_k * tiles_l;
    NC_CHECK(gx < (1ll << 31) && a.B <= 65535, "conv4d: grid too large");
    dim3 grid((unsigned)gx, a.G * a.nct, a.B), block(TX, TY);
    const size_t smem = ((size_t)(TY + KS - 1) * (PT * TX + 4) + (size_t)KS * KS * (CO == 1 ? 1 : CO) + 8) * sizeof(float);
    switch (CO) {
        case 16: conv4d_fwd_kernel<KS, 16, T><<<grid, block, smem, st>>>(a); break;
        case 8: conv4d_fwd_kernel<KS, 8, T><<<grid, block, smem, st>>>(a); break;
        case 4: conv4d_fwd_kernel<KS, 4, T><<<grid, block, smem, st>>>(a); break;
        case 1: conv4d_fwd_thin_kernel<KS, T><<<grid, block, smem, st>>>(a); break;
        default: return fail("conv4d: bad channel tile %d", CO);
    }
    NC_LAUNCH_OK("conv4d_fwd_kernel");
    return 0;
}

// runs the effective convolution described by `d` (already validated)
template <typename T>
static int run_conv(const void *x, const float *w, const float *bias, const void *mask, void *y, float *scratch,
                    PackDesc d, int B, int Gx, int d1, int d2, int d3, int d4, int relu, cudaStream_t st) {
    d.CO = pick_co(d.Cout_e);
    d.nct = (d.Cout_e + d.CO - 1) / d.CO;
    const long long total = (long long)d.G_e * d.nct * d.Cin_e * d.k * d.k * d.k * d.k * d.CO;
    const int pb = (int)((total + 255) / 256 < 1184 ? (total + 255) / 256 : 1184);
    pack_weights_kernel<<<pb, 256, 0, st>>>(w, scratch, d);
    NC_LAUNCH_OK("pack_weights_kernel");
    ConvArgs a{};
    a.x = x; a.y = y; a.wp = scratch; a.bias = bias; a.mask = mask;
    a.B = B; a.Cin = d.Cin_e; a.Cout = d.Cout_e; a.G = d.G_e; a.Gx = Gx;
    a.d1 = d1; a.d2 = d2; a.d3 = d3; a.d4 = d4; a.nct = d.nct; a.relu = relu;
    if (d.k == 3) return launch_conv<3, T>(a, d.CO, st);
    if (d.k == 5) return launch_conv<5, T>(a, d.CO, st);
    return fail("conv4d: kernel size %d not supported (3 or 5)", d.k);
}

int conv4d_fp32_fwd(const void *x, const float *w, const float *bias, void *y, void *scratch, int B, int Cin,
                    int Cout, int d1, int d2, int d3, int d4, int k, int w_layout, int symmetric, int shared_input,
                    int relu, int dtype, cudaStream_t st) {
    PackDesc d{};
    d.Cin = Cin; d.Cout = Cout; d.k = k; d.w_layout = w_layout; d.dgrad = 0; d.merge_groups = 0;
    d.G_e = symmetric ? 2 : 1; d.Cin_e = Cin; d.Cout_e = Cout;
    const int Gx = (symmetric && !shared_input) ? 2 : 1;
    if (dtype == NCNET_F32)
        return run_conv<float>(x, w, bias, nullptr, y, (float *)scratch, d, B, Gx, d1, d2, d3, d4, relu, st);
    return run_conv<__half>(x, w, bias, nullptr, y, (float *)scratch, d, B, Gx, d1, d2, d3, d4, relu, st);
}

int conv4d_fp32_dgrad(const void *gy, const float *w, const void *mask, void *gx, void *scratch, int B, int Cin,
                      int Cout, int d1, int d2, int d3, int d4, int k, int w_layout, int symmetric, int shared_input,
                      int dtype, cudaStream_t st) {
    PackDesc d{};
    d.Cin = Cin; d.Cout = Cout; d.k = k; d.w_layout = w_layout; d.dgrad = 1;
    const int G = symmetric ? 2 : 1;
    int Gx;
    if (G == 2 && shared_input) {        // both branches flow into the same Cin channels: one conv, 2*Cout inputs
        d.merge_groups = 1; d.G_e = 1; d.Cin_e = 2 * Cout; d.Cout_e = Cin; Gx = 1;
    } else {
        d.merge_groups = 0; d.G_e = G; d.Cin_e = Cout; d.Cout_e = Cin; Gx = G;
    }
    if (dtype == NCNET_F32)
        return run_conv<float>(gy, w, nullptr, mask, gx, (float *)scratch, d, B, Gx, d1, d2, d3, d4, 0, st);
    return run_conv<__half>(gy, w, nullptr, mask, gx, (float *)scratch, d, B, Gx, d1, d2, d3, d4, 0, st);
}

// ---- wgrad host ------------------------------------------------------------------------------
struct WgradPlan { int CT, nctile, cop, TK, ksplit, threads, S; size_t smem, part_bytes; };

static WgradPlan wgrad_plan(int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int G) {
    WgradPlan w{};
    w.cop = Cout <= 16 ? ((Cout + 3) / 4) * 4 : -1;
    w.CT = Cin >= 4 ? 4 : 1;
    w.nctile = (Cin + w.CT - 1) / w.CT;
    const int OQ = w.cop / 4;
    const int base = OQ * w.CT * k;
    w.ksplit = 256 / base;
    if (w.ksplit < 1) w.ksplit = 1;
    if (w.ksplit > d3) w.ksplit = d3;
    w.threads = ((base * w.ksplit + 31) / 32) * 32;
    // rows per stage so the tiles fit ~96 KB
    int TK = d3;
    auto bytes = [&](int tk) {
        return ((size_t)w.cop * tk * d4 + (size_t)w.CT * (tk + k - 1) * (d4 + k - 1)) * sizeof(float);
    };
    while (TK > 1 && bytes(TK) > 96 * 1024) TK = (TK + 1) / 2;
    w.TK = TK;
    size_t red = (size_t)w.threads * 4 * k * sizeof(float);
    w.smem = bytes(TK) > red ? bytes(TK) : red;
    const long long base_blocks = (long long)G * k * k * w.nctile;
    const long long planes = (long long)B * d1 * d2;
    long long S = (6ll * kNumSMs + base_blocks - 1) / base_blocks;
    if (S > planes) S = planes;
    if (S < 1) S = 1;
    w.S = (int)S;
    w.part_bytes = (size_t)w.S * G * Cin * k * k * k * k * w.cop * sizeof(float);
    return w;
}

size_t conv4d_fp32_wgrad_scratch_bytes(int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int symmetric) {
    if (Cout > 16) return 0;
    return wgrad_plan(B, Cin, Cout, d1, d2, d3, d4, k, symmetric ? 2 : 1).part_bytes + 256;
}

template <int KS, typename T>
static int launch_wgrad(WgradArgs a, const WgradPlan &w, cudaStream_t st) {
    dim3 grid(w.S, a.G * KS * KS * w.nctile);
    cudaError_t e;
    if (w.CT == 4) {
        e = cudaFuncSetAttribute(conv4d_wgrad_kernel<KS, 4, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)w.smem);
        NC_CHECK(e == cudaSuccess, "wgrad: smem attr: %s", cudaGetErrorString(e));
        conv4d_wgrad_kernel<KS, 4, T><<<grid, w.threads, w.smem, st>>>(a);
    } else {
        e = cudaFuncSetAttribute(conv4d_wgrad_kernel<KS, 1, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)w.smem);
        NC_CHECK(e == cudaSuccess, "wgrad: smem attr: %s", cudaGetErrorString(e));
        conv4d_wgrad_kernel<KS, 1, T><<<grid, w.threads, w.smem, st>>>(a);
    }
    NC_LAUNCH_OK("conv4d_wgrad_kernel");
    return 0;
}

// shared tail of both wgrad back ends: slice partials -> reference layout (+ bias gradient)
int wgrad_finish(const float *part, float *gw, float *gb, float *bias_part, const float *out_scale, const void *gy, int S, int G, int B, int Cin,
                 int Cout, int k, int cop, int w_layout, size_t vol, int dtype, cudaStream_t st) {
    const int total = Cout * Cin * k * k * k * k;
    unpack_wgrad_kernel<<<(total + 255) / 256, 256, 0, st>>>(part, gw, S, G, Cin, Cout, k, cop, w_layout, out_scale);
    NC_LAUNCH_OK("unpack_wgrad_kernel");
    if (gb) {
        // enough blocks to fill the GPU at small batches (a few channels x 6 chunks left 140 SMs idle)
        int chunks = (int)((kBiasTargetCtas + (size_t)B * G * Cout - 1) / ((size_t)B * G * Cout));
        const int max_chunks = (int)((vol + 2047) / 2048);
        if (chunks > max_chunks) chunks = max_chunks;
        if (chunks < 1) chunks = 1;
        NC_CHECK(B * G * Cout <= 65535, "wgrad: too many channels for bias grid");
        if (dtype == NCNET_F32)
            bias_grad_kernel<float><<<dim3(chunks, B * G * Cout), 256, 0, st>>>((const float *)gy, bias_part, Cout, G, vol, chunks);
        else
            bias_grad_kernel<__half><<<dim3(chunks, B * G * Cout), 256, 0, st>>>((const __half *)gy, bias_part, Cout, G, vol, chunks);
        NC_LAUNCH_OK("bias_grad_kernel");
        return bias_reduce(bias_part, gb, Cout, chunks * B * G, Cout, out_scale, st);
    }
    return 0;
}

template <typename T>
static int run_wgrad(const void *x, const void *gy, float *gw, float *gb, void *scratch, float *bias_part, const float *out_scale, int B, int Cin,
                     int Cout, int d1, int d2, int d3, int d4, int k, int w_layout, int symmetric, int shared_input,
                     cudaStream_t st) {
    const int G = symmetric ? 2 : 1;
    NC_CHECK(Cout <= 16, "conv4d_wgrad: Cout=%d > 16 not supported", Cout);
    const WgradPlan w = wgrad_plan(B, Cin, Cout, d1, d2, d3, d4, k, G);
    WgradArgs a{};
    a.x = x; a.gy = gy; a.part = (float *)scratch;
    a.B = B; a.Cin = Cin; a.Cout = Cout; a.G = G; a.Gx = (G == 2 && !shared_input) ? 2 : 1;
    a.d1 = d1; a.d2 = d2; a.d3 = d3; a.d4 = d4; a.S = w.S; a.nctile = w.nctile; a.TK = w.TK; a.cop = w.cop;
    a.ksplit = w.ksplit;
    // slices whose every plane is skipped still must publish zeros
    cudaError_t e = cudaMemsetAsync(scratch, 0, w.part_bytes, st);
    NC_CHECK(e == cudaSuccess, "wgrad: memset: %s", cudaGetErrorString(e));
    int rc = (k == 3) ? launch_wgrad<3, T>(a, w, st) : (k == 5 ? launch_wgrad<5, T>(a, w, st)
                                                              : fail("conv4d_wgrad: kernel size %d not supported", k));
    if (rc) return rc;
    return wgrad_finish((const float *)scratch, gw, gb, bias_part, out_scale, gy, w.S, G, B, Cin, Cout, k, w.cop, w_layout,
                        (size_t)d1 * d2 * d3 * d4, sizeof(T) == 4 ? NCNET_F32 : NCNET_F16, st);
}

int conv4d_fp32_wgrad(const void *x, const void *gy, float *gw, float *gb, void *scratch, float *bias_part, const float *out_scale, int B, int Cin,
                      int Cout, int d1, int d2, int d3, int d4, int k, int w_layout, int symmetric, int shared_input,
                      int dtype, cudaStream_t st) {
    if (dtype == NCNET_F32)
        return run_wgrad<float>(x, gy, gw, gb, scratch, bias_part, out_scale, B, Cin, Cout, d1, d2, d3, d4, k, w_layout, symmetric,
                                shared_input, st);
    return run_wgrad<__half>(x, gy, gw, gb, scratch, bias_part, out_scale, B, Cin, Cout, d1, d2, d3, d4, k, w_layout, symmetric,
                             shared_input, st);
}

// Power-of-two gradient scale for the fp16 backward pass.  scale3 = {S, 1/S, scratch}: S = 2^e with max|g| * S in
// [8, 16), so the largest gradient sits 12 binades below the fp16 overflow threshold and ordinary ones far above its
// subnormal range (gradients of a mean over ~10^6 cells are ~1e-7 and would otherwise be quantised to multiples of
// 6e-8).  Scaling by a power of two is exact and the backward pass is linear in g, so (1/S) * backward(S * g) is the
// same function of g; only the fp16 rounding inside moves to where it is harmless.
template <typename T>
__global__ void __launch_bounds__(256)
grad_absmax_kernel(const T *__restrict__ g, long long n, float *__restrict__ scale3) {
    float m = 0.f;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x) {
        const float v = fabsf(ldf(g + e));
        m = (v == v && v <= 3.0e38f) ? fmaxf(m, v) : m;                 // ignore NaN / inf
    }
    m = warp_max(m);
    __shared__ float sm[8];
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) m = fmaxf(m, sm[w]);
        atomicMax(reinterpret_cast<unsigned *>(scale3 + 2), __float_as_uint(m));   // non-negative floats order as integers
    }
}

__global__ void grad_scale_finish_kernel(float *__restrict__ scale3) {
    const float m = scale3[2];
    int e = 0;
    float s = 1.f;
    if (m > 0.f) {
        frexpf(m, &e);                       // m = f * 2^e, f in [0.5, 1)
        int k = 4 - e;                       // max * 2^k in [8, 16)
        k = k > 100 ? 100 : (k < -100 ? -100 : k);
        s = ldexpf(1.f, k);
    }
    scale3[0] = s;
    scale3[1] = 1.f / s;
}

int grad_scale_pow2(const void *g, long long n, int dtype, float *scale3, cudaStream_t st) {
    cudaError_t e = cudaMemsetAsync(scale3, 0, 3 * sizeof(float), st);
    NC_CHECK(e == cudaSuccess, "grad_scale: memset: %s", cudaGetErrorString(e));
    const int blocks = (int)((n + 2047) / 2048 < 4 * kNumSMs ? (n + 2047) / 2048 : 4 * kNumSMs);
    if (dtype == NCNET_F32) grad_absmax_kernel<float><<<blocks, 256, 0, st>>>((const float *)g, n, scale3);
    else grad_absmax_kernel<__half><<<blocks, 256, 0, st>>>((const __half *)g, n, scale3);
    NC_LAUNCH_OK("grad_absmax_kernel");
    grad_scale_finish_kernel<<<1, 1, 0, st>>>(scale3);
    NC_LAUNCH_OK("grad_scale_finish_kernel");
    return 0;
}

int relu_mask_bcast(const void *y, const void *g, void *gz, int B, int G, long long n, int dtype, const float *scale,
                    cudaStream_t st) {
    int bx = (int)((n + 255) / 256 < 2048 ? (n + 255) / 256 : 2048);
    if (dtype == NCNET_F32)
        relu_mask_bcast_kernel<float><<<dim3(bx, B), 256, 0, st>>>((const float *)y, (const float *)g, (float *)gz, G, n, scale);
    else
        relu_mask_bcast_kernel<__half><<<dim3(bx, B), 256, 0, st>>>((const __half *)y, (const __half *)g, (__half *)gz, G, n, scale);
    NC_LAUNCH_OK("relu_mask_bcast_kernel");
    return 0;
}

}  // namespace ncnet
