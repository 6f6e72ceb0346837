// Shared between the Conv4d back ends: how an "effective" convolution (forward, A<->B swapped
// branch, dgrad) reads the reference's weight tensor.
#pragma once
#include "common.cuh"

namespace ncnet {

struct PackDesc {
    int Cin, Cout, k;      // the layer as the reference defines it
    int w_layout;          // NCNET_W_OIKKKK / NCNET_W_KOIKKK
    int dgrad;             // 1: effective conv is the input-gradient conv (taps flipped, channels transposed)
    int merge_groups;      // dgrad + shared input: both branches merged into one conv with 2*Cout inputs
    int G_e, Cin_e, Cout_e;  // effective grouped conv
    int CO, nct;           // output-channel tile and tiles per group (back-end specific)
};

__host__ __device__ inline size_t ref_weight_index(const PackDesc &d, int o, int c, int a, int b, int c3, int d4) {
    const size_t k = d.k;
    if (d.w_layout == NCNET_W_OIKKKK) return (((((size_t)o * d.Cin + c) * k + a) * k + b) * k + c3) * k + d4;
    return (((((size_t)a * d.Cout + o) * d.Cin + c) * k + b) * k + c3) * k + d4;   // [k,Cout,Cin,k,k,k]
}

// element of the reference weight that effective-conv tap E_g[oe][ci][t1..t4] equals
__host__ __device__ inline size_t packed_source_index(const PackDesc &d, int g, int oe, int ci, int t1, int t2, int t3,
                                                      int t4) {
    int grp = g, o, c;
    if (!d.dgrad) {
        o = oe; c = ci;
    } else {
        if (d.merge_groups) { grp = ci / d.Cout; o = ci % d.Cout; } else { o = ci; }
        c = oe;
        t1 = d.k - 1 - t1; t2 = d.k - 1 - t2; t3 = d.k - 1 - t3; t4 = d.k - 1 - t4;
    }
    if (grp == 1) {   // branch run on the A<->B transposed volume: W'[o,c,a,b,c',d'] = W[o,c,c',d',a,b]
        int s;
        s = t1; t1 = t3; t3 = s;
        s = t2; t2 = t4; t4 = s;
    }
    return ref_weight_index(d, o, c, t1, t2, t3, t4);
}

// fp32 CUDA-core back end (conv4d_fp32.cu)
size_t conv4d_fp32_scratch_bytes(int Cin, int Cout, int k, int symmetric);
int conv4d_fp32_fwd(const void *x, const float *w, const float *bias, void *y, void *scratch, int B, int Cin, int Cout,
                    int d1, int d2, int d3, int d4, int k, int w_layout, int symmetric, int shared_input, int relu,
                    int dtype, cudaStream_t st);
int conv4d_fp32_dgrad(const void *gy, const float *w, const void *mask, void *gx, void *scratch, int B, int Cin,
                      int Cout, int d1, int d2, int d3, int d4, int k, int w_layout, int symmetric, int shared_input,
                      int dtype, cudaStream_t st);
size_t conv4d_fp32_wgrad_scratch_bytes(int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int symmetric);
// bias_part: bias_partial_bytes(...) bytes for the two-stage (atomic-free, run-to-run reproducible) bias-gradient sum
int conv4d_fp32_wgrad(const void *x, const void *gy, float *gw, float *gb, void *scratch, float *bias_part, const float *out_scale, int B, int Cin,
                      int Cout, int d1, int d2, int d3, int d4, int k, int w_layout, int symmetric, int shared_input,
                      int dtype, cudaStream_t st);

// Bias gradient in two deterministic stages: `chunks` CTAs per channel row write partial sums
// part[entry][W] (entry = chunk-major over (chunk, batch, group); W = channels per entry), then one small
// kernel adds the entries of each output channel in index order.  The region lives at the head of the wgrad scratch.
constexpr int kBiasTargetCtas = 16 * kNumSMs;
inline size_t bias_partial_bytes(int B, int G, int Cout) {
    const size_t rows = (size_t)B * G * (Cout > 2 ? Cout : 2);        // channel rows of either back end
    return (((size_t)kBiasTargetCtas + rows) * 8 * sizeof(float) + 255) & ~(size_t)255;
}
// out_scale (device, may be null): the reduced sums are multiplied by *out_scale (gradient un-scaling)
int bias_reduce(const float *part, float *db, int Cout, int entries, int W, const float *out_scale, cudaStream_t st);
int relu_mask_bcast(const void *y, const void *g, void *gz, int B, int G, long long n, int dtype, const float *scale,
                    cudaStream_t st);
int grad_scale_pow2(const void *g, long long n, int dtype, float *scale3, cudaStream_t st);

}  // namespace ncnet
