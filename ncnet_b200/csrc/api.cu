// C-ABI glue: error state, argument validation and back-end dispatch for the Conv4d entry points.
#include "common.cuh"
#include "conv4d_common.cuh"
#include "conv4d_tc.cuh"

namespace ncnet {

char *last_error_buf() {
    static thread_local char buf[512] = "";
    return buf;
}

int fail(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(last_error_buf(), 512, fmt, ap);
    va_end(ap);
    return 1;
}

static int check_conv(const char *who, int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int w_layout,
                      int dtype, int mode) {
    NC_CHECK(B > 0 && Cin > 0 && Cout > 0 && d1 > 0 && d2 > 0 && d3 > 0 && d4 > 0, "%s: bad shape", who);
    NC_CHECK(k % 2 == 1, "%s: even kernel size %d (the reference only supports odd sizes)", who, k);
    NC_CHECK(k == 3 || k == 5, "%s: kernel size %d not supported (3 or 5)", who, k);
    NC_CHECK(w_layout == NCNET_W_OIKKKK || w_layout == NCNET_W_KOIKKK, "%s: bad weight layout %d", who, w_layout);
    NC_CHECK(dtype == NCNET_F32 || dtype == NCNET_F16, "%s: unsupported dtype %d", who, dtype);
    NC_CHECK(mode == NCNET_CONV_FP32 || mode == NCNET_CONV_TC16, "%s: unknown mode %d", who, mode);
    return 0;
}

}  // namespace ncnet

using namespace ncnet;

extern "C" int ncnet_version(void) { return 100; }   // 0.1.0
extern "C" const char *ncnet_last_error(void) { return last_error_buf(); }

extern "C" size_t ncnet_conv4d_scratch_bytes(int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k,
                                             int symmetric, int mode) {
    size_t a = conv4d_fp32_scratch_bytes(Cin, Cout, k, symmetric);
    if (mode == NCNET_CONV_TC16) {
        size_t b = conv4d_tc_scratch_bytes(B, Cin, Cout, d1, d2, d3, d4, k, symmetric);
        return a > b ? a : b;
    }
    return a;
}

extern "C" int ncnet_conv4d_fwd(const void *x, const float *weight, const float *bias, void *y, void *scratch, int B,
                                int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int w_layout, int symmetric,
                                int shared_input, int relu, int dtype, int mode, void *stream) {
    NC_CHECK(x && weight && y && scratch, "conv4d_fwd: null pointer");
    if (check_conv("conv4d_fwd", B, Cin, Cout, d1, d2, d3, d4, k, w_layout, dtype, mode)) return 1;
    cudaStream_t st = (cudaStream_t)stream;
    if (mode == NCNET_CONV_TC16) {
        const int rc = conv4d_tc_fwd(x, weight, bias, y, scratch, B, Cin, Cout, d1, d2, d3, d4, k, w_layout, symmetric,
                                     shared_input, relu, dtype, st);
        if (rc != 2) return rc;      // 2: shape outside the tensor-core kernel's envelope -> CUDA-core kernel
    }
    return conv4d_fp32_fwd(x, weight, bias, y, scratch, B, Cin, Cout, d1, d2, d3, d4, k, w_layout, symmetric,
                           shared_input, relu, dtype, st);
}

extern "C" int ncnet_conv4d_dgrad(const void *grad_y, const float *weight, const void *y_prev_mask, void *grad_x,
                                  void *scratch, int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k,
                                  int w_layout, int symmetric, int shared_input, int dtype, int mode, void *stream) {
    NC_CHECK(grad_y && weight && grad_x && scratch, "conv4d_dgrad: null pointer");
    if (check_conv("conv4d_dgrad", B, Cin, Cout, d1, d2, d3, d4, k, w_layout, dtype, mode)) return 1;
    cudaStream_t st = (cudaStream_t)stream;
    if (mode == NCNET_CONV_TC16) {
        const int rc = conv4d_tc_dgrad(grad_y, weight, y_prev_mask, grad_x, scratch, B, Cin, Cout, d1, d2, d3, d4, k,
                                       w_layout, symmetric, shared_input, dtype, st);
        if (rc != 2) return rc;
    }
    return conv4d_fp32_dgrad(grad_y, weight, y_prev_mask, grad_x, scratch, B, Cin, Cout, d1, d2, d3, d4, k, w_layout,
                             symmetric, shared_input, dtype, st);
}

extern "C" size_t ncnet_conv4d_wgrad_scratch_bytes(int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k,
                                                   int symmetric, int mode) {
    size_t a = conv4d_fp32_wgrad_scratch_bytes(B, Cin, Cout, d1, d2, d3, d4, k, symmetric);
    if (mode == NCNET_CONV_TC16) {
        size_t b = conv4d_tc_wgrad_scratch_bytes(B, Cin, Cout, d1, d2, d3, d4, k, symmetric);
        return a > b ? a : b;
    }
    return a;
}

extern "C" int ncnet_conv4d_wgrad(const void *x, const void *grad_y, float *grad_weight, float *grad_bias,
                                  void *scratch, int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k,
                                  int w_layout, int symmetric, int shared_input, int dtype, int mode, void *stream) {
    NC_CHECK(x && grad_y && grad_weight && scratch, "conv4d_wgrad: null pointer");
    if (check_conv("conv4d_wgrad", B, Cin, Cout, d1, d2, d3, d4, k, w_layout, dtype, mode)) return 1;
    cudaStream_t st = (cudaStream_t)stream;
    if (mode == NCNET_CONV_TC16) {
        const int rc = conv4d_tc_wgrad(x, grad_y, grad_weight, grad_bias, scratch, B, Cin, Cout, d1, d2, d3, d4, k,
                                       w_layout, symmetric, shared_input, dtype, st);
        if (rc != 2) return rc;      // 2: outside the tensor-core kernel's envelope -> CUDA-core kernel
    }
    return conv4d_fp32_wgrad(x, grad_y, grad_weight, grad_bias, scratch, B, Cin, Cout, d1, d2, d3, d4, k, w_layout,
                             symmetric, shared_input, dtype, st);
}

extern "C" int ncnet_relu_mask_bcast(const void *y, const void *grad, void *gz, int B, int G, long long n, int dtype,
                                     void *stream) {
    NC_CHECK(y && grad && gz, "relu_mask_bcast: null pointer");
    NC_CHECK(B > 0 && B <= 65535 && G > 0 && n > 0, "relu_mask_bcast: bad shape");
    NC_CHECK(dtype == NCNET_F32 || dtype == NCNET_F16, "relu_mask_bcast: unsupported dtype %d", dtype);
    return relu_mask_bcast(y, grad, gz, B, G, n, dtype, (cudaStream_t)stream);
}
