// tcgen05 (fp16 operands, fp32 accumulate in TMEM) Conv4d back end -- see conv4d_tc.cu
#pragma once
#include "common.cuh"

namespace ncnet {
size_t conv4d_tc_scratch_bytes(int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int symmetric);
// true when the effective convolution (Cin_e -> Cout_e per group) can run on the tensor-core kernel
bool conv4d_tc_supports(int Cin_e, int Cout_e, int k, int d3, int d4);
int conv4d_tc_fwd(const void *x, const float *w, const float *bias, void *y, void *scratch, int B, int Cin, int Cout,
                  int d1, int d2, int d3, int d4, int k, int w_layout, int symmetric, int shared_input, int relu,
                  int dtype, cudaStream_t st);
int conv4d_tc_dgrad(const void *gy, const float *w, const void *mask, void *gx, void *scratch, int B, int Cin,
                    int Cout, int d1, int d2, int d3, int d4, int k, int w_layout, int symmetric, int shared_input,
                    int dtype, cudaStream_t st);
size_t conv4d_tc_wgrad_scratch_bytes(int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int symmetric);
int conv4d_tc_wgrad(const void *x, const void *gy, float *gw, float *gb, void *scratch, int B, int Cin, int Cout,
                    int d1, int d2, int d3, int d4, int k, int w_layout, int symmetric, int shared_input, int dtype,
                    cudaStream_t st);
}  // namespace ncnet
