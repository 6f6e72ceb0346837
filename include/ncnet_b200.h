/*
 * ncnet_b200 -- C ABI of the B200-native NCNet matching head.
 *
 * The reference (leejaeyong7/ncnet) has no FFI layer: its head is pure
 * PyTorch.  Each entry point below replaces the body of one reference
 * function/module (cited as file:line relative to the reference checkout);
 * the Python drop-in modules in ncnet_b200/ call them through ctypes.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless named host_*; all tensors are
 *     dense, row-major, in the reference's own layout;
 *   - the library never allocates or frees device memory and keeps no pointer
 *     after returning; scratch is passed in by the caller (sizes from the
 *     *_scratch_bytes queries).  The one exception is
 *     ncnet_mutual_matching_fwd_noalloc, whose scratch is an array built into
 *     the library and whose per-device address the library caches;
 *   - `stream` is a cudaStream_t passed as void*; launches are asynchronous,
 *     nothing synchronises the device;
 *   - return 0 on success, non-zero on error; ncnet_last_error() returns the
 *     message of the calling thread's last failure;
 *   - dtype: NCNET_F32 (0) or NCNET_F16 (1) = storage type of the activation
 *     tensors.  Arithmetic is always fp32 (fp32 accumulation).
 */
#ifndef NCNET_B200_H
#define NCNET_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NCNET_F32 0
#define NCNET_F16 1
/* Conv4d activations only, TC16 mode only: channel-blocked, zero-haloed fp16 ("CB16"),
 * [blocks of 16 channels][2 halves][d1*d2][d3+2p][d4+2p][8], p = k/2 of the layer that reads or writes
 * it; a tensor [B, G*C] (C<=16 per group) has B*G blocks.  Lets a stack of layers hand activations
 * from one tensor-core launch to the next without layout passes (sizes: ncnet_cb16_bytes). */
#define NCNET_CB16 2
/* A CB16 tensor immediately followed (at the next multiple of 256 bytes after its ncnet_cb16_bytes) by its
 * position-major twin: [blocks][d1*d2][(d3+2p)*(d4+2p)][16] fp16, the two 16-byte halves of a position swapped where
 * bit 7 of the in-plane byte offset is set.  ncnet_conv4d_fwd / _dgrad write both when asked for this output type;
 * ncnet_conv4d_wgrad then reads the twin directly instead of converting (training stacks); every other use reads
 * the CB16 part.  Buffer size: 2 * ((ncnet_cb16_bytes + 255) & ~255). */
#define NCNET_CB16PM 3

/* weight layouts accepted by the Conv4d entry points */
#define NCNET_W_OIKKKK 0 /* [Cout,Cin,k,k,k,k]   (lib/conv4d.py:16-18, permute_filters=True) */
#define NCNET_W_KOIKKK 1 /* [k,Cout,Cin,k,k,k]   (the stored, pre-permuted layout, lib/conv4d.py:76-77) */

/* Conv4d compute modes */
#define NCNET_CONV_FP32 0 /* fp32 FMA on CUDA cores */
#define NCNET_CONV_TC16 1 /* fp16 operands, fp32 accumulate, tcgen05 tensor cores */

int ncnet_version(void);
const char *ncnet_last_error(void);

/* featureL2Norm -- lib/model.py:14-17.   x,y: [B,C,HW];  y = x / sqrt(sum_c x^2 + 1e-6) */
int ncnet_feature_l2norm(const void *x, void *y, int B, int C, int HW, int dtype, void *stream);

/* FeatureCorrelation(shape='4D') -- lib/model.py:106-118.
 * fa: [B,C,nA], fb: [B,C,nB] -> out: [B,nA,nB];  out[b,a,t] = sum_c fa[b,c,a]*fb[b,c,t].
 * relu_l2norm != 0 applies the module's normalization=True branch (ReLU, then L2 norm
 * over the singleton channel dim: v/sqrt(v*v+1e-6)).
 * The 3D shape (lib/model.py:97-105) is the same product with fa's two spatial axes swapped
 * and the roles of A and B exchanged; the Python wrapper does that re-indexing. */
int ncnet_corr4d_fwd(const void *fa, const void *fb, void *out, int B, int C, int nA, int nB,
                     int dtype, int relu_l2norm, void *stream);
/* The same product on the tensor cores (tcgen05, fp16 operands, fp32 accumulation in TMEM; <= 1e-3 normalised
 * error).  scratch: ncnet_corr4d_tc_scratch_bytes(...) bytes (the fp16 operand tiles).  l2norm_inputs != 0 fuses
 * featureL2Norm (lib/model.py:14-17, applied by the trunk at lib/model.py:85-86) into the operand packing, for
 * callers that hand over un-normalised trunk outputs. */
size_t ncnet_corr4d_tc_scratch_bytes(int B, int C, int nA, int nB);
int ncnet_corr4d_fwd_tc(const void *fa, const void *fb, void *out, void *scratch, int B, int C, int nA, int nB,
                        int dtype, int relu_l2norm, int l2norm_inputs, void *stream);
/* Correlation with the relocalisation max-pool fused into its epilogue -- lib/model.py:269-272 (FeatureCorrelation
 * followed by maxpool4d, lib/model.py:177-191) for one image pair (the reference's maxpool4d is batch-1 only).
 * fa: [C,hA,wA], fb: [C,hB,wB], sides even, k_size = 2.  value: [hA/2,wA/2,hB/2,wB/2] in `dtype`; max_i..max_l: the
 * window-relative arg-max offsets of the same shape (idx_bytes 8 = int64 as the reference returns them, 1 = uint8).
 * Bit-identical to ncnet_corr4d_fwd_tc followed by ncnet_maxpool4d_fwd, but the un-pooled volume (16x the output) is
 * never written.  scratch: ncnet_corr4d_tc_scratch_bytes(1, C, hA*wA, hB*wB).  relu_l2norm / l2norm_inputs as above. */
int ncnet_corr4d_pool_fwd_tc(const void *fa, const void *fb, void *value, void *max_i, void *max_j, void *max_k,
                             void *max_l, void *scratch, int C, int hA, int wA, int hB, int wB, int k_size,
                             int idx_bytes, int dtype, int relu_l2norm, int l2norm_inputs, void *stream);
/* gradient of the plain (un-normalised) correlation wrt both feature maps (autograd of lib/model.py:113) */
int ncnet_corr4d_bwd(const void *grad_out, const void *fa, const void *fb, void *grad_fa, void *grad_fb,
                     int B, int C, int nA, int nB, int dtype, void *stream);

/* MutualMatching -- lib/model.py:155-175.  x,y: [B,nA,nB] (the volume seen as a matrix,
 * nA=fs1*fs2, nB=fs3*fs4).  keys: scratch/saved state, (B*(nA+nB)) uint64: packed
 * (max value, first arg-max) per source cell then per target cell; it is also what the
 * backward pass consumes.  in_groups=2 reads x as [B,2,nA,nB] and uses x[:,0]+x[:,1]
 * (the sum of the two symmetric NeighConsensus branches, lib/model.py:147). */
size_t ncnet_mutual_matching_keys_bytes(int B, int nA, int nB);
int ncnet_mutual_matching_fwd(const void *x, void *y, uint64_t *keys, int B, int nA, int nB,
                              int in_groups, int dtype, void *stream);
/* The same forward pass for callers that will not run the backward pass (inference): pass 1 tracks the maxima only,
 * which halves its instruction count; `keys` then holds (max value, index 0) and must not be given to _bwd. */
int ncnet_mutual_matching_fwd_infer(const void *x, void *y, uint64_t *keys, int B, int nA, int nB,
                                    int in_groups, int dtype, void *stream);
/* The two passes of ncnet_mutual_matching_fwd on their own, for volumes sharded into slabs along the first
 * source axis (BASELINE.json config 4): every rank runs _max on its rows, the per-target halves of `keys`
 * (entries [nA, nA+nB) of each batch item) are max-reduced across ranks by the caller, then every rank runs
 * _apply.  Keys compare as unsigned 64-bit integers: larger value wins, then the smaller index. */
int ncnet_mutual_matching_max(const void *x, uint64_t *keys, int B, int nA, int nB, int in_groups, int dtype,
                              void *stream);
int ncnet_mutual_matching_apply(const void *x, const uint64_t *keys, void *y, int B, int nA, int nB, int in_groups,
                                int dtype, void *stream);
/* backward: grad_x = d(y)/d(x)^T grad_y incl. the flow through both max reductions.
 * sums: scratch of ncnet_mutual_matching_sums_bytes(B,nA,nB) bytes (the reduced row/column sums followed by
 * per-CTA partial sums: the sums are formed in a fixed order, without floating-point atomics, so two runs give
 * the same bits).  x is the forward input (in_groups as in forward; grad_x is [B,nA,nB]: the gradient of the
 * summed volume). */
size_t ncnet_mutual_matching_sums_bytes(int B, int nA, int nB);
int ncnet_mutual_matching_bwd(const void *x, const void *grad_y, const uint64_t *keys, float *sums,
                              void *grad_x, int B, int nA, int nB, int in_groups, int dtype, void *stream);

/* MutualMatching forward for callers that cannot allocate device memory (no counterpart in the
 * reference: it exists for hosts such as inference plugins or real-time loops whose contract
 * forbids cudaMalloc where the head runs, so they have no way to provide the `keys` table).
 * Same result as ncnet_mutual_matching_fwd; the key table is a 2 MiB array built into the library
 * (held on every GPU that loads it, until the process exits), so B*(nA+nB) must be <= 262144 --
 * larger problems are rejected with an error.  Calls no cudaMalloc/cudaFree or allocating routine.
 * The first call on each device looks up the table's device address (cudaGetSymbolAddress); make
 * that first call with the device current and before starting any stream capture.  All calls on a
 * device share the one table: it still holds the previous call's maxima on entry (the part a call
 * uses is cleared on its stream first), and two calls whose kernels overlap on the GPU -- different
 * streams without an ordering between them -- race on it.  The cached address belongs to the CUDA
 * context it was looked up in: after cudaDeviceReset (or any teardown of the primary context) it is
 * stale and this entry point must not be called again in that process.  The arg-max keys are not
 * returned, so this variant serves inference; the backward pass needs ncnet_mutual_matching_fwd. */
int ncnet_mutual_matching_fwd_noalloc(const void *x, void *y, int B, int nA, int nB, int in_groups,
                                      int dtype, void *stream);

/* Weak-supervision matching score -- train.py:123-134 with normalization='softmax':
 *   score = ( mean_{b,t} max_a softmax_a(corr[b,a,t]) + mean_{b,a} max_t softmax_t(corr[b,a,t]) ) / 2
 * (the reference adds the two score vectors, which needs nA == nB; for nA != nB the two means are taken
 * separately).  corr: [B,nA,nB].  keys: B*(nA+nB) uint64 and sums: ncnet_mutual_matching_sums_bytes(B,nA,nB) bytes
 * (its first B*(nA+nB) floats are the row/column sums of exponentials) are outputs that the backward pass
 * consumes; score: one float on the device.  _bwd: grad_corr = grad_score[0] * d(score)/d(corr).
 * item_weight (device, B floats, may be NULL): the mean over the batch becomes sum_b item_weight[b] * score_b.  With
 * the positive pairs and the rolled negative pairs of train.py:135-139 stacked into one batch of 2B and weights
 * (-1/B, ..., +1/B, ...) the result is the weak loss itself (score_neg - score_pos) from ONE pass of the head. */
int ncnet_match_score_fwd(const void *corr, uint64_t *keys, float *sums, const float *item_weight, float *score, int B,
                          int nA, int nB, int dtype, void *stream);
int ncnet_match_score_bwd(const void *corr, const uint64_t *keys, const float *sums, const float *item_weight,
                          const float *grad_score, void *grad_corr, int B, int nA, int nB, int dtype, void *stream);

/* conv4d / Conv4d.forward -- lib/conv4d.py:11-51,81-82  (+ fused ReLU, lib/model.py:137-138,
 * + the symmetric second branch of NeighConsensus.forward, lib/model.py:147).
 *
 * x: [B, Gin*Cin, d1,d2,d3,d4]   y: [B, G*Cout, d1,d2,d3,d4]   weight: fp32, layout w_layout,
 * bias: fp32 [Cout] or NULL.  symmetric=0: G=1.  symmetric=1: G=2 -- group 0 uses the
 * weights as stored, group 1 uses them with the (A,B) tap pairs exchanged
 * (W'[o,c,a,b,c',d'] = W[o,c,c',d',a,b]), which equals running the layer on the A<->B
 * transposed volume and transposing back.  shared_input!=0: both groups read the same
 * Cin channels (Gin=1, first layer); else group g reads channels [g*Cin,(g+1)*Cin).
 * relu!=0 fuses max(0,.) after the bias.  mode: NCNET_CONV_FP32 / NCNET_CONV_TC16.
 * scratch: ncnet_conv4d_scratch_bytes(...) bytes (re-laid-out weights; in TC16 mode also the
 * channel-blocked fp16 copy of the input).  One query covers fwd and dgrad of the layer. */
size_t ncnet_conv4d_scratch_bytes(int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k,
                                  int symmetric, int mode);
int ncnet_conv4d_fwd(const void *x, const float *weight, const float *bias, void *y, void *scratch,
                     int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k,
                     int w_layout, int symmetric, int shared_input, int relu, int x_dtype, int y_dtype,
                     int mode, void *stream);
/* gradient wrt the layer input (autograd of lib/conv4d.py:39-48).  grad_y: [B,G*Cout,...];
 * grad_x: [B,Gin*Cin,...].  y_prev_mask (may be NULL): a tensor shaped like grad_x; where it is
 * <= 0 the result is zeroed (the ReLU of the previous layer folded into this epilogue); it must be
 * CB16 exactly when grad_x is.  With shared_input the two groups' gradients are summed into the
 * Cin channels. */
int ncnet_conv4d_dgrad(const void *grad_y, const float *weight, const void *y_prev_mask, void *grad_x,
                       void *scratch, int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k,
                       int w_layout, int symmetric, int shared_input, int gy_dtype, int mask_dtype,
                       int gx_dtype, int mode, void *stream);
/* gradient wrt weight (same layout as `weight`, fp32, OVERWRITTEN) and bias (fp32 [Cout], may be
 * NULL).  With symmetric=1 the two groups' contributions are folded back into the one weight.
 * out_scale (device float, may be NULL): both results are multiplied by *out_scale -- the 1/S of
 * ncnet_grad_scale when grad_y carries a scaled gradient. */
size_t ncnet_conv4d_wgrad_scratch_bytes(int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k,
                                        int symmetric, int mode);
int ncnet_conv4d_wgrad(const void *x, const void *grad_y, float *grad_weight, float *grad_bias,
                       void *scratch, int B, int Cin, int Cout, int d1, int d2, int d3, int d4, int k,
                       int w_layout, int symmetric, int shared_input, int x_dtype, int gy_dtype, int mode,
                       const float *out_scale, void *stream);

/* 1 when a layer with Cin -> Cout channels per group (kernel k, last two volume dims d3,d4) runs
 * entirely on the tensor-core kernels in TC16 mode (forward, dgrad and wgrad), i.e. can take and
 * produce CB16 activations; otherwise TC16 mode routes it to the CUDA-core kernels (planar only). */
int ncnet_conv4d_tc_supported(int Cin, int Cout, int k, int d3, int d4);
/* Forward only (inference): 1 when ncnet_conv4d_fwd can take and produce NCNET_CB16 for this layer shape, even if the
 * tensor-core dgrad / wgrad envelope (whole planes staged in shared memory) does not cover it. */
int ncnet_conv4d_tc_fwd_supported(int Cin, int Cout, int k, int d3, int d4);

/* CB16 helpers: byte size of a tensor of `nblk` 16-channel blocks; conversion from / to the planar
 * reference layout [nblk, C, d1,d2,d3,d4] (C <= 16, dtype NCNET_F32/F16). */
size_t ncnet_cb16_bytes(int nblk, int d1, int d2, int d3, int d4, int k);
int ncnet_cb16_pack(const void *planar, void *cb16, int nblk, int C, int d1, int d2, int d3, int d4, int k,
                    int dtype, void *stream);
int ncnet_cb16_unpack(const void *cb16, void *planar, int nblk, int C, int d1, int d2, int d3, int d4, int k,
                      int dtype, void *stream);

/* ReLU backward on the last layer's two branch outputs (lib/model.py:137-138,147):
 * gz[b,g,p] = (y[b,g,p] > 0) ? scale * grad[b,p] : 0 for g < G.   y,gz: [B,G,n]; grad: [B,n];
 * scale: device float or NULL (= 1). */
int ncnet_relu_mask_bcast(const void *y, const void *grad, void *gz, int B, int G, long long n,
                          int dtype, const float *scale, void *stream);
/* Power-of-two loss scale for a backward pass that stores gradients in fp16 (TC16 mode; no counterpart in the
 * reference, whose backward is fp32).  grad: n values; scale3: 3 device floats, on return {S, 1/S, max|grad|} with
 * S = 2^e such that max|grad|*S is in [8,16) (S = 1 for an all-zero gradient).  Scaling by a power of two is exact
 * and the backward pass is linear in grad, so un-scaling the weight gradients with 1/S gives the same function; the
 * fp16 roundings in between then act on values around 1 instead of on ~1e-7 values in fp16's subnormal range. */
int ncnet_grad_scale(const void *grad, long long n, int dtype, float *scale3, void *stream);

/* maxpool4d -- lib/model.py:177-191.  x: [H1,H2,H3,H4] (batch 1, channel 0), each divisible by k.
 * value: [H1/k,...] same dtype; offsets: four tensors of that shape, window-relative arg-max
 * coordinates, (i,j,k,l) row-major window scan, ties -> first.  idx_bytes: 8 (int64, what the
 * reference returns) or 1 (uint8). */
int ncnet_maxpool4d_fwd(const void *x, void *value, void *max_i, void *max_j, void *max_k, void *max_l,
                        int H1, int H2, int H3, int H4, int k, int idx_bytes, int dtype, void *stream);

/* corr_to_matches -- lib/point_tnf.py:12-80.  corr: [B,nA,nB] with nA=fs1*fs2, nB=fs3*fs4.
 * invert=0: one match per target cell (arg-max over the source axis), N=nB;
 * invert=1: one per source cell, N=nA.  do_softmax: soft-max along the reduced axis first,
 * arg-max taken on the soft-max values, ties -> first index.  delta_*: the four maxpool4d
 * offset tensors [fs1,fs2,fs3,fs4] (batch 1 only, may be NULL), idx_bytes as above, k_size the
 * pooling factor.  lut_*: fp32 coordinate tables of lengths fs2*k, fs1*k, fs4*k, fs3*k built
 * by the caller exactly like lib/point_tnf.py:16-27.  Outputs [B,N]: xA,yA,xB,yB,score fp32;
 * iA,jA,iB,jB int64.  scratch: ncnet_corr_to_matches_scratch_bytes(...) bytes (0 for most shapes: only the
 * invert=0 reduction over a volume with few batch items and many source cells splits its rows over CTAs and
 * combines them through scratch; NULL is accepted when the query returns 0). */
size_t ncnet_corr_to_matches_scratch_bytes(int B, int fs1, int fs2, int fs3, int fs4, int invert, int dtype);
int ncnet_corr_to_matches(const void *corr, int B, int fs1, int fs2, int fs3, int fs4,
                          int do_softmax, int invert, int k_size,
                          const void *delta_iA, const void *delta_jA, const void *delta_iB,
                          const void *delta_jB, int idx_bytes,
                          const float *lut_xA, const float *lut_yA, const float *lut_xB, const float *lut_yB,
                          float *xA, float *yA, float *xB, float *yB, float *score,
                          int64_t *iA, int64_t *jA, int64_t *iB, int64_t *jB, void *scratch, int dtype, void *stream);

/* InLoc match post-processing -- eval_inloc.py:151-203, on the device instead of numpy on the host.
 * Input: the (xA,yA,xB,yB,score) columns of corr_to_matches for one direction (n1 matches) and optionally the
 * other (n2 matches) -- the reference concatenates them in this order.  n2 == 0 (second-direction pointers and
 * scratch may be NULL) is the reference's single-direction branch: nothing is removed, the order is kept, the
 * coordinates are recentred.  Otherwise duplicate (xA,yA,xB,yB) columns are removed keeping the highest score (the reference sorts by descending score
 * and lets np.unique keep the first occurrence); the survivors come out in np.unique's order (lexicographic in
 * xA,yA,xB,yB), recentred as  v*(n-1)/n + 0.5/n  with n = recentre_* (the reference's fs*k_size per axis; 0 = leave
 * the coordinate alone), as rows [xA,yA,xB,yB,score] of out[(n1+n2),5]; unused rows are zero; *count (device int) =
 * number of rows written.  scratch: ncnet_matches_unique_scratch_bytes(n1+n2) bytes.  The duplicate search compares
 * all pairs (at most 2^20 matches). */
size_t ncnet_matches_unique_scratch_bytes(int n);
int ncnet_matches_sort_unique(const float *xA, const float *yA, const float *xB, const float *yB, const float *score,
                              int n1, const float *xA2, const float *yA2, const float *xB2, const float *yB2,
                              const float *score2, int n2, int recentre_yA, int recentre_xA, int recentre_yB,
                              int recentre_xB, float *out, int *count, void *scratch, void *stream);

/* PointsToUnitCoords / PointsToPixelCoords -- lib/point_tnf.py:151-167 (normalize_axis / unnormalize_axis, :6-10).
 * points,out: [B,2,N] (row 0 = x scaled by the image width, row 1 = y by the height); im_size: [B,im_stride] fp32
 * rows (h, w, ...).  to_unit != 0: pixel -> [-1,1]; 0: the inverse. */
int ncnet_points_axis_transform(const float *points, const float *im_size, int im_stride, float *out, int B, int N,
                                int to_unit, void *stream);
/* bilinearInterpPointTnf -- lib/point_tnf.py:96-148.  matches xA,yA,xB,yB: [B,fs*fs] (corr_to_matches, B->A
 * direction: xB,yB is the regular target grid); grid: the fs values np.linspace(-1,1,fs) as fp32;
 * target_points_norm,out: [B,2,N].  Like the reference the neighbour look-ups address the flattened match arrays
 * without a batch offset (the evaluation uses batch 1). */
int ncnet_bilinear_point_tnf(const float *xA, const float *yA, const float *xB, const float *yB, const float *grid,
                             int fs, const float *target_points_norm, float *out, int B, int N, void *stream);
/* pck -- lib/eval_util.py:12-25.  source_points, warped_points: [B,2,N] (invalid source points are -1 and trail the
 * valid ones); L_pck: [B]; out[b] = fraction of the valid points with distance <= L_pck[b]*alpha. */
int ncnet_pck(const float *source_points, const float *warped_points, const float *L_pck, float alpha, float *out,
              int B, int N, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* NCNET_B200_H */
