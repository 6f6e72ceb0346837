// tcgen05 (fp16 operands, fp32 accumulate in TMEM) Conv4d back end -- conv4d_tc.cu / conv4d_tc_wgrad.cu
#pragma once
#include "common.cuh"

namespace ncnet {
// CB16: channel-blocked, zero-haloed fp16 activations  [blocks of 16 ch][2][d1*d2][d3+2p][d4+2p][8], p = k/2
size_t cb16_bytes(size_t nblk, int d1, int d2, int d3, int d4, int k);
int cb16_pack(const void *x, int dtype, void *cb, int nblk, int C, int d1, int d2, int d3, int d4, int k, int tap_mode,
              cudaStream_t st);
int cb16_unpack(const void *cb, void *x, int dtype, int nblk, int C, int d1, int d2, int d3, int d4, int k, cudaStream_t st);

size_t conv4d_tc_scratch_bytes(int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int symmetric);
// true when the effective convolution (Cin_e -> Cout_e per group) can run on the tensor-core kernel
bool conv4d_tc_supports(int Cin_e, int Cout_e, int k, int d3, int d4);
// all three return 2 when the shape / layout combination is outside the tensor-core envelope
int conv4d_tc_fwd(const void *x, int x_dtype, const float *w, const float *bias, void *y, int y_dtype, void *scratch, int B,
                  int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int w_layout, int symmetric, int shared_input,
                  int relu, cudaStream_t st);
int conv4d_tc_dgrad(const void *gy, int gy_dtype, const float *w, const void *mask, int mask_dtype, void *gx, int gx_dtype,
                    void *scratch, int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int w_layout,
                    int symmetric, int shared_input, cudaStream_t st);
// conv4d_tc_walk16.cu: mode M1 layers with CB16 output on persistent CTA pairs with resident weights; returns 2 outside its envelope
bool conv4d_tc_walk16_ok(int k, int d3, int d4);
int conv4d_tc_walk16_run(const __half *act, const __half *wp, const float *bias, const __half *mask, __half *out, __half *out_pm,
                         int B, int G, int Gx, int Cout, int d1, int d2, int d3, int d4, int k, int relu, cudaStream_t st);
// conv4d_tc_walk1.cu: mode M4 layers (Cin == 1) with CB16 output as a persistent walk along iA; og = 2: both branches of a
// shared input side by side in N
struct PackDesc;
bool conv4d_tc_walk1_ok(int k, int og, int d3, int d4);
int conv4d_tc_walk1_run(const __half *act, const float *w, __half *wp, const PackDesc &d, const float *bias, const __half *mask,
                        __half *out, __half *out_pm, int B, int og, int Gx, int d1, int d2, int d3, int d4, int relu,
                        cudaStream_t st);
size_t conv4d_tc_wgrad_scratch_bytes(int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int symmetric);
int conv4d_tc_wgrad(const void *x, int x_dtype, const void *gy, int gy_dtype, float *gw, float *gb, void *scratch,
                    float *bias_part, const float *out_scale, int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int w_layout,
                    int symmetric, int shared_input, cudaStream_t st);
}  // namespace ncnet
