// tcgen05 Conv4d back end (placeholder until the implicit-GEMM kernel lands): reports an error,
// never silently falls back.
#include "conv4d_tc.cuh"

namespace ncnet {
size_t conv4d_tc_scratch_bytes(int, int, int, int) { return 0; }
int conv4d_tc_fwd(const void *, const float *, const float *, void *, void *, int, int, int, int, int, int, int, int,
                  int, int, int, int, int, cudaStream_t) {
    return fail("conv4d_fwd: NCNET_CONV_TC16 is not built into this library");
}
int conv4d_tc_dgrad(const void *, const float *, const void *, void *, void *, int, int, int, int, int, int, int, int,
                    int, int, int, int, cudaStream_t) {
    return fail("conv4d_dgrad: NCNET_CONV_TC16 is not built into this library");
}
}  // namespace ncnet
