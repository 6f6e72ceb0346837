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

static bool planar(int dt) { return dt == NCNET_F32 || dt == NCNET_F16; }

extern "C" int ncnet_conv4d_fwd(const void *x, const float *weight, const float *bias, void *y, void *scratch, int B,
                                int Cin, int Cout, int d1, int d2, int d3, int d4, int k, int w_layout, int symmetric,
                                int shared_input, int relu, int x_dtype, int y_dtype, int mode, void *stream) {
    NC_CHECK(x && weight && y && scratch, "conv4d_fwd: null pointer");
    if (check_conv("conv4d_fwd", B, Cin, Cout, d1, d2, d3, d4, k, w_layout, NCNET_F32, mode)) return 1;
    cudaStream_t st = (cudaStream_t)stream;
    if (mode == NCNET_CONV_TC16) {
        const int rc = conv4d_tc_fwd(x, x_dtype, weight, bias, y, y_dtype, scratch, B, Cin, Cout, d1, d2, d3, d4, k,
                                     w_layout, symmetric, shared_input, relu, st);
        if (rc != 2) return rc;      // 2: outside the tensor-core kernel's envelope -> CUDA-core kernel
    }
    NC_CHECK(planar(x_dtype) && x_dtype == y_dtype,
             "conv4d_fwd: this shape runs on the CUDA-core kernel, which needs planar tensors of one dtype (got %d -> %d)",
             x_dtype, y_dtype);
    return conv4d_fp32_fwd(x, weight, bias, y, scratch, B, Cin, Cout, d1, d2, d3, d4, k, w_layout, symmetric,
                           shared_input, relu, x_dtype, st);
}

extern "C" int ncnet_conv4d_dgrad(const void *grad_y, const float *weight, const void *y_prev_mask, void *grad_x,
                                  void *scratch, int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k,
                                  int w_layout, int symmetric, int shared_input, int gy_dtype, int mask_dtype,
                                  int gx_dtype, int mode, void *stream) {
    NC_CHECK(grad_y && weight && grad_x && scratch, "conv4d_dgrad: null pointer");
    if (check_conv("conv4d_dgrad", B, Cin, Cout, d1, d2, d3, d4, k, w_layout, NCNET_F32, mode)) return 1;
    cudaStream_t st = (cudaStream_t)stream;
    if (mode == NCNET_CONV_TC16) {
        const int rc = conv4d_tc_dgrad(grad_y, gy_dtype, weight, y_prev_mask, mask_dtype, grad_x, gx_dtype, scratch, B,
                                       Cin, Cout, d1, d2, d3, d4, k, w_layout, symmetric, shared_input, st);
        if (rc != 2) return rc;
    }
    NC_CHECK(planar(gy_dtype) && gy_dtype == gx_dtype && (!y_prev_mask || mask_dtype == gx_dtype),
             "conv4d_dgrad: this shape runs on the CUDA-core kernel, which needs planar tensors of one dtype");
    return conv4d_fp32_dgrad(grad_y, weight, y_prev_mask, grad_x, scratch, B, Cin, Cout, d1, d2, d3, d4, k, w_layout,
                             symmetric, shared_input, gy_dtype, st);
}

extern "C" size_t ncnet_conv4d_wgrad_scratch_bytes(int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k,
                                                   int symmetric, int mode) {
    size_t a = conv4d_fp32_wgrad_scratch_bytes(B, Cin, Cout, d1, d2, d3, d4, k, symmetric);
    if (mode == NCNET_CONV_TC16) {
        size_t b = conv4d_tc_wgrad_scratch_bytes(B, Cin, Cout, d1, d2, d3, d4, k, symmetric);
        a = a > b ? a : b;
    }
    if (a == 0) return 0;
    return a + bias_partial_bytes(B, symmetric ? 2 : 1, Cout);       // head of the scratch: bias-gradient partial sums
}

extern "C" int ncnet_conv4d_wgrad(const void *x, const void *grad_y, float *grad_weight, float *grad_bias,
                                  void *scratch, int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k,
                                  int w_layout, int symmetric, int shared_input, int x_dtype, int gy_dtype, int mode,
                                  const float *out_scale, void *stream) {
    NC_CHECK(x && grad_y && grad_weight && scratch, "conv4d_wgrad: null pointer");
    if (check_conv("conv4d_wgrad", B, Cin, Cout, d1, d2, d3, d4, k, w_layout, NCNET_F32, mode)) return 1;
    cudaStream_t st = (cudaStream_t)stream;
    float *bias_part = (float *)scratch;
    scratch = (uint8_t *)scratch + bias_partial_bytes(B, symmetric ? 2 : 1, Cout);
    if (mode == NCNET_CONV_TC16) {
        const int rc = conv4d_tc_wgrad(x, x_dtype, grad_y, gy_dtype, grad_weight, grad_bias, scratch, bias_part, out_scale, B, Cin, Cout,
                                       d1, d2, d3, d4, k, w_layout, symmetric, shared_input, st);
        if (rc != 2) return rc;      // 2: outside the tensor-core kernel's envelope -> CUDA-core kernel
    }
    NC_CHECK(planar(x_dtype) && x_dtype == gy_dtype,
             "conv4d_wgrad: this shape runs on the CUDA-core kernel, which needs planar tensors of one dtype");
    return conv4d_fp32_wgrad(x, grad_y, grad_weight, grad_bias, scratch, bias_part, out_scale, B, Cin, Cout, d1, d2, d3, d4, k,
                             w_layout, symmetric, shared_input, x_dtype, st);
}

extern "C" int ncnet_conv4d_tc_supported(int Cin, int Cout, int k, int d3, int d4) {
    if (Cin < 1 || Cout < 1 || (k != 3 && k != 5) || d3 < 1 || d4 < 1) return 0;
    const bool fwd = conv4d_tc_supports(Cin, Cout, k, d3, d4);
    const bool bwd = conv4d_tc_supports(Cout, Cin, k, d3, d4);
    const bool wg = conv4d_tc_wgrad_scratch_bytes(1, Cin, Cout, 1, 1, d3, d4, k, 0) > 0;
    return fwd && bwd && wg ? 1 : 0;
}

extern "C" int ncnet_conv4d_tc_fwd_supported(int Cin, int Cout, int k, int d3, int d4) {
    if (Cin < 1 || Cout < 1 || (k != 3 && k != 5) || d3 < 1 || d4 < 1) return 0;
    return conv4d_tc_supports(Cin, Cout, k, d3, d4) ? 1 : 0;
}

extern "C" size_t ncnet_cb16_bytes(int nblk, int d1, int d2, int d3, int d4, int k) {
    return cb16_bytes((size_t)nblk, d1, d2, d3, d4, k);
}

extern "C" int ncnet_cb16_pack(const void *planar_t, void *cb16, int nblk, int C, int d1, int d2, int d3, int d4, int k,
                               int dtype, void *stream) {
    NC_CHECK(planar_t && cb16, "cb16_pack: null pointer");
    NC_CHECK(d1 > 0 && d2 > 0 && d3 > 0 && d4 > 0 && (k == 3 || k == 5), "cb16_pack: bad shape");
    return cb16_pack(planar_t, dtype, cb16, nblk, C, d1, d2, d3, d4, k, 0, (cudaStream_t)stream);
}

extern "C" int ncnet_cb16_unpack(const void *cb16, void *planar_t, int nblk, int C, int d1, int d2, int d3, int d4, int k,
                                 int dtype, void *stream) {
    NC_CHECK(planar_t && cb16, "cb16_unpack: null pointer");
    NC_CHECK(d1 > 0 && d2 > 0 && d3 > 0 && d4 > 0 && (k == 3 || k == 5), "cb16_unpack: bad shape");
    return cb16_unpack(cb16, planar_t, dtype, nblk, C, d1, d2, d3, d4, k, (cudaStream_t)stream);
}

extern "C" int ncnet_grad_scale(const void *grad, long long n, int dtype, float *scale3, void *stream) {
    NC_CHECK(grad && scale3 && n > 0, "grad_scale: bad arguments");
    NC_CHECK(dtype == NCNET_F32 || dtype == NCNET_F16, "grad_scale: unsupported dtype %d", dtype);
    return grad_scale_pow2(grad, n, dtype, scale3, (cudaStream_t)stream);
}

extern "C" int ncnet_relu_mask_bcast(const void *y, const void *grad, void *gz, int B, int G, long long n, int dtype,
                                     const float *scale, void *stream) {
    NC_CHECK(y && grad && gz, "relu_mask_bcast: null pointer");
    NC_CHECK(B > 0 && B <= 65535 && G > 0 && n > 0, "relu_mask_bcast: bad shape");
    NC_CHECK(dtype == NCNET_F32 || dtype == NCNET_F16, "relu_mask_bcast: unsupported dtype %d", dtype);
    return relu_mask_bcast(y, grad, gz, B, G, n, dtype, scale, (cudaStream_t)stream);
}
