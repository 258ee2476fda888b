/*
 * pl_int8.h -- C-ABI of the B200 (sm_100a) int8 conv / GEMM hot path.
 *
 * This is the drop-in boundary: the "new lite/backends/cuda/math int8 entry
 * points" that Paddle-Lite's kCUDA kernel classes call.  Plain C: device
 * pointers, sizes, POD descriptors and a cudaStream_t; no C++/torch types.
 * Every function only ENQUEUES work on `stream` (never synchronises), is
 * re-entrant, keeps no global state, and returns 0 on success, a positive
 * cudaError_t, or a negative PL_ERR_* code.  There is no CPU fallback.
 *
 * Reference interfaces replaced (paths in chenjiaoAngel/Paddle-Lite):
 *   pl_conv2d_i8_nhwc    lite/backends/cuda/math/cudnn_conv.cc:320-585
 *                        (CudnnConv2DInt8<P>::{init,create,run}: cuDNN int8
 *                        NHWC conv + a separate bias/act/requant kernel,
 *                        lite/backends/cuda/math/activation.cu:91-446)
 *   pl_dwconv2d_i8_nhwc  lite/backends/arm/math/conv3x3s{1,2}_depthwise_int8.cc,
 *                        conv5x5s{1,2}_depthwise_int8.cc (no kCUDA kernel exists)
 *   pl_gemm_i8           lite/backends/arm/math/gemm_s8.cc, gemv_arm_int8.cc:25-143;
 *                        shape contract of lite/backends/cuda/math/gemm.cc:25-121
 *   pl_quantize_f32_i8 / pl_dequantize_i8_f32
 *                        lite/kernels/cuda/calib_compute.cu:27-95,
 *                        lite/backends/arm/math/type_trans.cc:34-187
 *   pl_nchw_to_nhwc / pl_nhwc_to_nchw
 *                        lite/backends/cuda/math/transpose.cu:31-184
 *   pl_pool2d_i8_nhwc    lite/kernels/cuda/pool_compute.cu:28-447 (fp32 there)
 *   pl_add_i8            lite/backends/cuda/math/elementwise.cu:127-165
 *   fold helpers         lite/backends/cuda/math/cudnn_conv.cc:349-372,
 *                        lite/kernels/arm/conv_gemmlike.cc:207-264
 *
 * Layouts: activations NHWC int8 with an explicit pixel pitch `ld*` in
 * elements (>= channels; multiples of 16 for the tensor-core paths so that
 * TMA can address them; a channel slice of a wider buffer is a pointer offset
 * + the wide pitch, which makes concat/split free).  Conv filters arrive in
 * the reference's kCUDA int8 layout OHWI [K,R,S,C/groups]
 * (cudnn_conv.cc:329-337) and are re-packed once by pl_pack_conv_weight_i8.
 *
 * Epilogue ("ARM convention"): v = fmaf((float)acc, scale[n],
 * bias[n]); activation; int8 output = clamp(round-half-away(v), -127, 127);
 * PL_ROUND_CUDA_LEGACY = clamp [-128,127] + round-half-even
 * (lite/backends/cuda/math/utils.h:61-66).  `scale`/`bias` are the FOLDED
 * vectors (see pl_fold_scales), bias may be NULL.
 *
 * Stream ordering: every kernel is launched with programmatic stream serialization and waits
 * (griddepcontrol.wait) right before its first access to ACTIVATIONS or the residual; packed filters and
 * the scale / bias vectors are read before that wait, as constants.  That is safe for data produced by
 * this library's own pack / fold kernels and by host uploads on the same stream (none of them triggers
 * launch_dependents, so the dependent grid cannot start before they have completed and flushed), and for
 * anything that completed before the launch.  A caller that writes filters or scales with its OWN kernel
 * that calls cudaTriggerProgrammaticLaunchCompletion / griddepcontrol.launch_dependents must put an event or
 * a stream synchronisation between that kernel and the first launch that reads them.
 */
#ifndef PL_INT8_H_
#define PL_INT8_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* pl_stream_t; /* == cudaStream_t */

/* lite_api::ActivationType (lite/api/paddle_place.h:100-114) */
enum { PL_ACT_NONE = 0, PL_ACT_RELU = 1, PL_ACT_RELU6 = 2, PL_ACT_LEAKY = 4 };
enum { PL_OUT_INT8 = 0, PL_OUT_FP32 = 1, PL_OUT_INT32 = 2 };
enum { PL_ROUND_ARM = 0, PL_ROUND_CUDA_LEGACY = 1 };
enum { PL_POOL_MAX = 0, PL_POOL_AVG = 1 };

enum {
  PL_OK = 0,
  PL_ERR_ARG = -1,         /* bad / inconsistent argument */
  PL_ERR_ALIGN = -2,       /* pointer or pitch not 16-byte aligned */
  PL_ERR_UNSUPPORTED = -3, /* shape outside what the sm_100a kernels cover */
  PL_ERR_DRIVER = -4,      /* tensor-map encode / driver entry point failed */
  PL_ERR_DEVICE = -5,      /* device is not sm_100 */
  PL_ERR_LIMIT = -6        /* request exceeds a fixed in-library limit (pl_gemm_i8_peers_static) */
};

/* Fused epilogue shared by conv / depthwise / gemm. */
typedef struct pl_epilogue {
  const float* scale; /* [n_out] folded per-output-channel scale (device) */
  const float* bias;  /* [n_out] folded bias (device) or NULL */
  int act;            /* PL_ACT_* */
  float act_arg;      /* relu6: six (already / out_scale for int8 out); leaky: alpha */
  int out_dtype;      /* PL_OUT_* ; PL_OUT_INT32 stores the raw accumulator */
  int round_mode;     /* PL_ROUND_* */
  /* optional fused residual (ConvParam::residualData, op_params.h:446-502):
   * v = fmaf((float)res[m,n], res_scale, v) before the activation. */
  const int8_t* residual;
  int ld_res;
  float res_scale;
} pl_epilogue;

typedef struct pl_conv_desc {
  int N, H, W, C; /* input NHWC, C = logical channels */
  int ldx;        /* input pixel pitch (elements) */
  int K;          /* output channels */
  int ldy;        /* output pixel pitch (elements of the output dtype) */
  int R, S;       /* filter height, width */
  int stride_h, stride_w;
  int pad_t, pad_b, pad_l, pad_r; /* ConvParam::paddings order */
  int dil_h, dil_w;
  int groups;
} pl_conv_desc;

/* Scheduling hint, no reference counterpart (the reference's kernels own no tile order): the conv / depthwise /
 * gemm / pool launches that the CALLING THREAD issues after this call walk their output tiles last-to-first
 * (reverse != 0) or first-to-last (0, the default).  Results are identical either way.  A runtime that alternates
 * the order along a chain of layers lets every layer start on the bytes its producer wrote last, which are still in
 * the 126 MB L2 (ResNet-50 b256: -1 % step time, nets.DeviceNet.bind).  Thread-local; captured CUDA graphs keep the
 * order they were captured with. */
void pl_set_tile_order(int reverse);

/* ---- host-side helpers (no device work) ---------------------------------- */
int pl_conv_out_size(int in, int k, int dil, int pad0, int pad1, int stride);
/* K-block (bytes of the reduction dim staged per pipeline stage) chosen for C. */
int pl_conv_block_k(int C);
/* bytes of the packed filter produced by pl_pack_conv_weight_i8 */
size_t pl_conv_packed_weight_bytes(int K, int R, int S, int C);
/* s[c] = w_scale[c]*in_scale (fp32 out) or *in_scale/out_scale (int8 out);
 * bias[c] /= out_scale for int8 out; relu6 six /= out_scale for int8 out.
 * HOST pointers.  n_w_scale is 1 or n_out. */
int pl_fold_scales(const float* w_scale, int n_w_scale, int n_out, float in_scale,
                   float out_scale, int out_dtype, const float* bias,
                   float* scale_folded, float* bias_folded, int act,
                   float act_arg, float* act_arg_folded);

/* ---- one-off transforms (PrepareForRun) ------------------------------------ */
/* OHWI [K,R,S,C] int8 (device) -> packed [K][R*S][Cb*BK] zero padded (device) */
int pl_pack_conv_weight_i8(const int8_t* w_ohwi, int8_t* w_packed, int K, int R,
                           int S, int C, pl_stream_t stream);
/* [rows, cols] -> [cols, rows] int8 (fc W [K,N] -> [N,K]); pitches in elements */
int pl_transpose_i8(const int8_t* src, int ld_src, int8_t* dst, int ld_dst,
                    int rows, int cols, pl_stream_t stream);

/* ---- tensor-core paths (tcgen05 kind::i8, TMEM int32, TMA) ------------------ */
/* dense conv (groups == 1): implicit GEMM, M = N*P*Q, N = K, K = R*S*C.
 * x, w_packed 16-byte aligned; ldx % 16 == 0. */
int pl_conv2d_i8_nhwc(const int8_t* x, const int8_t* w_packed, void* y,
                      const pl_conv_desc* d, const pl_epilogue* e,
                      pl_stream_t stream);
/* C[M,N] = epilogue(A[M,K] . Bt[N,K]^T).  A, Bt K-contiguous ("K-major"),
 * lda % 16 == 0, ldb % 16 == 0. */
int pl_gemm_i8(const int8_t* A, int lda, const int8_t* Bt, int ldb, void* C,
               int ldc, int M, int N, int K, const pl_epilogue* e,
               pl_stream_t stream);
/* Column-parallel (N-split) GEMM fused with the all-gather of its output columns (BASELINE config 5;
 * no counterpart in the reference, which has no multi-GPU path): same as pl_gemm_i8, and every
 * output tile is ALSO written to n_peers other buffers C_peers[i] (same shape/pitch as C: normally the
 * same column block of the full [M, N_total] result on every other GPU, reached through NVLink peer
 * mappings).  The stores are issued tile by tile by the kernel's TMA store warp, so the transfer
 * overlaps the mainloop; the caller only synchronises the ranks afterwards.  int8 output only.
 * maps_dev: 64-byte aligned device scratch of n_peers * 128 bytes. */
int pl_gemm_i8_peers(const int8_t* A, int lda, const int8_t* Bt, int ldb, void* C, void* const* C_peers,
                     int n_peers, void* maps_dev, int ldc, int M, int N, int K, const pl_epilogue* e,
                     pl_stream_t stream);

/* One-off set-up for repeated / captured launches: encodes the peer maps for (C, ldc, M, N, K, e) and uploads
 * them synchronously into maps_dev.  Afterwards pl_gemm_i8_peers may be called with C_peers == NULL (and the
 * same n_peers, maps_dev, shape and epilogue): the call is then a single kernel launch -- nothing is encoded
 * or copied, so it can be captured in a CUDA graph. */
int pl_gemm_i8_peers_prepare(void* C, void* const* C_peers, int n_peers, void* maps_dev, int ldc, int M, int N,
                             int K, const pl_epilogue* e);
/* Allocation-free variant of pl_gemm_i8_peers.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving
 * loops, harnesses that forbid it).  Same arithmetic, same outputs, same peer stores; the scratch that holds
 * the peer tensor maps is not passed in but lives in the library image: a file-scope
 * `static __device__ __align__(256)` array of PL_STATIC_PEERS_MAX * 128 bytes (1 KB: the 7 peers of an 8-GPU box fit; pl_gemm_i8_peers takes up to 15) that every GPU which
 * loads the library holds until the process exits.  Its device address is obtained once per device (with
 * that device current) through cudaGetSymbolAddress.  The call itself allocates nothing: no cudaMalloc /
 * cudaFree / allocating library routine (one small cudaMemcpyAsync uploads the maps, as in
 * pl_gemm_i8_peers).  n_peers > PL_STATIC_PEERS_MAX returns PL_ERR_LIMIT (and says so on stderr).
 * All calls on one GPU share the table, which keeps the last call's maps (nothing in it has to start at
 * zero: every call rewrites the entries it uses).  Calls on the same stream are ordered; calls that
 * OVERLAP on the GPU (different streams, or different host threads driving the same device) race on it. */
#define PL_STATIC_PEERS_MAX 8
int pl_gemm_i8_peers_static(const int8_t* A, int lda, const int8_t* Bt, int ldb, void* C, void* const* C_peers,
                            int n_peers, int ldc, int M, int N, int K, const pl_epilogue* e,
                            pl_stream_t stream);

/* ---- bandwidth-bound paths (16-byte int8 vectors, no tensor cores) --------- */
/* depthwise conv: groups == C == K, filter [C,R,S,1] as given by the reference */
int pl_dwconv2d_i8_nhwc(const int8_t* x, const int8_t* w_crs, void* y,
                        const pl_conv_desc* d, const pl_epilogue* e,
                        pl_stream_t stream);
int pl_quantize_f32_i8(const float* x, int8_t* y, int64_t n, float scale,
                       int round_mode, pl_stream_t stream);
int pl_dequantize_i8_f32(const int8_t* x, float* y, int64_t n, float scale,
                         pl_stream_t stream);
/* fused entry of the int8 region: fp32 NCHW -> int8 NHWC (pitch ldy, pad
 * channels written as 0) with q = sat(round(x * (1/scale))) */
int pl_quantize_nchw_f32_to_nhwc_i8(const float* x, int8_t* y, int N, int C,
                                    int H, int W, int ldy, float scale,
                                    int round_mode, pl_stream_t stream);
/* pack transform for thin-channel first layers (C <= 4 per 32-bit word): explicit
 * im2col A[N*P*Q, ldA] with A[m][(r*S+s)*cbytes + j] = x[n, ih, iw, j] (0 outside the
 * image and in the row pad), so that the conv runs as pl_gemm_i8 with fat K rows
 * instead of R*S thin TMA im2col loads.  cbytes % 4 == 0, ldx % 4 == 0, ldA % 16 == 0. */
int pl_im2col_i8_nhwc(const int8_t* x, int ldx, int8_t* A, int ldA, int N, int H,
                      int W, int cbytes, int R, int S, int sh, int sw, int pt,
                      int pl, int dh, int dw, int P, int Q, pl_stream_t stream);
/* thin first layers, direct: input with C <= 4 stored as ONE 32-bit word per pixel
 * (ldx == 4, e.g. from pl_quantize_nchw_f32_to_nhwc_i8 with ldy = 4).  Extra warps of the
 * tensor-core kernel gather the R x S window of every output pixel on chip, so no im2col
 * matrix is written to HBM.  dilation 1; windows: R in {3,5,7} with S <= 8 (see
 * pl_thin_conv_supported).  w_thin comes from pl_pack_thin_conv_weight_i8 (OHWI in). */
int pl_thin_conv_supported(int R, int S, int C, int dil_h, int dil_w);
size_t pl_thin_conv_packed_weight_bytes(int K, int R, int S);
int pl_pack_thin_conv_weight_i8(const int8_t* w_ohwi, int8_t* w_thin, int K, int R, int S,
                                int C, pl_stream_t stream);
int pl_thin_conv2d_i8_nhwc(const int8_t* x, const int8_t* w_thin, void* y,
                           const pl_conv_desc* d, const pl_epilogue* e, pl_stream_t stream);
/* thin first layers with stride 2, space-to-depth form (the fast path for 7x7/s2/p3 and 3x3/s2/p1
 * stems): the input lives as 16-byte "s2d pixels" = 2x2 original pixels x 4 channels (byte order
 * dy, dx, c), [N][HP][WS] with zero borders, written by pl_quantize_nchw_f32_to_s2d_i8 into a buffer
 * of pl_s2d_geometry(...).bytes that the caller zero-filled once.  A stride-2 R x S conv is then a
 * stride-1 TA x TB-tap conv over 16-byte pixels; every output row's A operand is TA contiguous
 * spans of the buffer (1-D bulk copies) that the tensor core reads in place through overlapping
 * un-swizzled descriptors -- no im2col and no gather.  Same arithmetic as pl_conv2d_i8_nhwc on the
 * original NHWC tensor (the replaced reference path is the same ConvComputeInt8,
 * lite/kernels/cuda/conv_compute.cc:55-106).  Supported: stride 2x2, dilation 1, C <= 4, even H/W,
 * pad_t == pad_l, R == S, Q <= 128 and at most 4 taps per axis after the 2x2 fold. */
typedef struct pl_s2d_geometry {
  int supported;
  int HP, WS;        /* padded s2d rows / 16-byte pixels per row */
  int pad_s;         /* border rows / pixels in front (top and left) */
  int shift;         /* original tap r sits at s2d tap a, sub-row dy with r = 2a + dy - shift */
  int TA, TB;        /* tap rows, taps per row (even) */
  int P, Q;          /* output size */
  size_t bytes;      /* buffer size incl. the read-ahead slack after the last row */
} pl_s2d_geometry;
pl_s2d_geometry pl_s2d_conv_geometry(int N, int H, int W, int C, int R, int S, int stride_h, int stride_w,
                                     int pad_t, int pad_b, int pad_l, int pad_r, int dil_h, int dil_w);
int pl_quantize_nchw_f32_to_s2d_i8(const float* x, int8_t* y_s2d, int N, int C, int H, int W,
                                   const pl_s2d_geometry* g, float scale, int round_mode, pl_stream_t stream);
/* Narrow feeds of the same entry (the fp32 NCHW tensor is 154 MB per 256 images: PCIe-bound end to end).
 * f16: q = calib((float)x).  u8: the image normalisation the reference applies on the host,
 * x = ((float)u8 - mean[c]) * norm[c] (lite/utils/cv/image2tensor.cc, `(pixel - mean) * scale`), then the
 * same calib step (lite/kernels/cuda/calib_compute.cu:27-95); mean / norm are HOST arrays of C floats.
 * x must be aligned to 4 elements (8 / 4 bytes). */
int pl_quantize_nchw_f16_to_s2d_i8(const void* x_f16, int8_t* y_s2d, int N, int C, int H, int W,
                                   const pl_s2d_geometry* g, float scale, int round_mode, pl_stream_t stream);
int pl_quantize_nchw_u8_to_s2d_i8(const uint8_t* x_u8, int8_t* y_s2d, int N, int C, int H, int W,
                                  const pl_s2d_geometry* g, const float* mean, const float* norm, float scale,
                                  int round_mode, pl_stream_t stream);
/* the same entry for an int8 NHWC tensor that already exists (C <= 4, pixel pitch ldx): the reference's op chain
 * calib -> layout -> conv ends here when the conv is a stride-2 stem.  Replaces the per-Run widening of the pixel
 * pitch in front of cudnn's int8 conv (lite/backends/cuda/math/cudnn_conv.cc:329-337 needs C % 4 == 0). */
int pl_nhwc_i8_to_s2d_i8(const int8_t* x, int ldx, int8_t* y_s2d, int N, int C, int H, int W,
                         const pl_s2d_geometry* g, pl_stream_t stream);
/* channel pitch change of an int8 NHWC tensor, pad channels zero ([pixels][ld_src] -> [pixels][ld_dst],
 * ld_dst % 4 == 0): how a kernel class stages an activation with C % 16 != 0 at a TMA-legal pitch
 * (lite/kernels/cuda/conv_compute.cc has no such case: cudnn int8 rejects those shapes) */
int pl_repitch_i8_nhwc(const int8_t* x, int ld_src, int8_t* y, int ld_dst, int64_t pixels, int C, pl_stream_t stream);
size_t pl_s2d_conv_packed_weight_bytes(int K, const pl_s2d_geometry* g);
int pl_pack_s2d_conv_weight_i8(const int8_t* w_ohwi, int8_t* w_s2d, int K, int R, int S, int C,
                               const pl_s2d_geometry* g, pl_stream_t stream);
/* d describes the ORIGINAL conv (H, W, C, R, S, strides, pads; ldx is ignored) */
int pl_s2d_conv2d_i8(const int8_t* x_s2d, const int8_t* w_s2d, void* y, const pl_conv_desc* d,
                     const pl_epilogue* e, pl_stream_t stream);
/* plain layout transforms; elem_bytes 1 or 4; ld = NHWC pixel pitch (elements) */
int pl_nchw_to_nhwc(const void* x, void* y, int N, int C, int H, int W, int ld,
                    int elem_bytes, pl_stream_t stream);
int pl_nhwc_to_nchw(const void* x, void* y, int N, int C, int H, int W, int ld,
                    int elem_bytes, pl_stream_t stream);
/* int8 NHWC pooling.  rescale = in_scale/out_scale (int8 out) or in_scale
 * (fp32 out).  max: v = (float)max * rescale; avg: v = ((float)sum/(float)cnt)
 * * rescale, cnt = clipped window (exclusive) or kh*kw. */
int pl_pool2d_i8_nhwc(const int8_t* x, int ldx, void* y, int ldy, int N, int H,
                      int W, int C, int kh, int kw, int sh, int sw, int pt,
                      int pl, int P, int Q, int pool_type, int exclusive,
                      float rescale, int out_dtype, int round_mode,
                      pl_stream_t stream);
/* adaptive pooling (PoolParam::adaptive: ksize is the OUTPUT size P x Q): window
 * [floor(i * in / out), ceil((i + 1) * in / out)) per axis, avg divides by the window size
 * (lite/kernels/cuda/pool_compute.cu:69-106, 150-187); values as pl_pool2d_i8_nhwc. */
int pl_adaptive_pool2d_i8_nhwc(const int8_t* x, int ldx, void* y, int ldy, int N, int H, int W, int C, int P,
                               int Q, int pool_type, float rescale, int out_dtype, int round_mode,
                               pl_stream_t stream);
/* y = q(act(fmaf((float)a, sa, (float)b * sb))) -- the ResNet residual add */
int pl_add_i8(const int8_t* a, const int8_t* b, void* y, int64_t n, float sa,
              float sb, int act, float act_arg, int out_dtype, int round_mode,
              pl_stream_t stream);

/* ---- channel movement between int8 NHWC ops ---------------
 * Pure byte moves (the reference's NCHW memcpy loops become per-pixel moves), so a
 * ShuffleNet-style network stays int8-NHWC-resident between convs.  `pixels` = N*H*W,
 * every pitch in elements.  Chunks of 16 / 4 / 1 bytes are picked from the alignment of
 * pointers, pitches and channel boundaries.  At most 8 parts.
 *   pl_concat_i8_nhwc           lite/backends/arm/math/concat.h:29-60 (axis = channels)
 *   pl_split_i8_nhwc            lite/backends/arm/math/split.cc:54-83
 *   pl_shuffle_channel_i8_nhwc  lite/backends/arm/math/shuffle_channel.cc:24-76:
 *                               out[j * group + i] = in[i * (C / group) + j]; x != y
 *   pl_relu_i8                  lite/kernels/cuda/relu_compute.cu:23-33 on the integers,
 *                               y = min(max(x, 0), hi); hi = 127 is plain relu, a smaller
 *                               hi is relu6 at the caller's quantised 6 / scale */
int pl_concat_i8_nhwc(int n_in, const int8_t* const* xs, const int* cs, const int* ldxs,
                      int8_t* y, int ldy, int64_t pixels, pl_stream_t stream);
int pl_split_i8_nhwc(const int8_t* x, int ldx, int n_out, int8_t* const* ys, const int* cs,
                     const int* ldys, int64_t pixels, pl_stream_t stream);
int pl_shuffle_channel_i8_nhwc(const int8_t* x, int ldx, int8_t* y, int ldy, int C, int group,
                               int64_t pixels, pl_stream_t stream);
int pl_relu_i8(const int8_t* x, int8_t* y, int64_t n, int hi, pl_stream_t stream);
/* concat(a, b) + shuffle_channel(group 2) in one pass (the tail of a ShuffleNet-v2 unit:
 * concat.h:29-60 followed by shuffle_channel.cc:24-76): y[2j] = a[j], y[2j+1] = b[j], j < c_half */
int pl_concat_shuffle2_i8_nhwc(const int8_t* a, int lda, const int8_t* b, int ldb, int8_t* y,
                               int ldy, int c_half, int64_t pixels, pl_stream_t stream);

/* ---- fp32 ops a quantised model keeps between / after its int8 convs -----------
 * NHWC fp32, pitches in elements.
 *   pl_channel_gather_f32_nhwc   dst[p, j] = srcs[table[j][0]][p * ld + table[j][1]]: any chain of concat
 *                                (lite/backends/arm/math/concat.h:29-60), shuffle_channel (shuffle_channel.cc:24-76)
 *                                and split (split.cc:54-83) composed into one channel map, at most 16 sources;
 *                                out_dtype PL_OUT_INT8 also applies the calib that follows such a chain in a
 *                                quantised model (lite/kernels/cuda/calib_compute.cu:27-95: x * (1 / scale), half
 *                                away, [-127, 127]).  table_dev: C x 2 int32 on the device.
 *   pl_conv2d_transpose_f32_nhwc lite/kernels/arm/conv_transpose_compute.cc:1-230 (golden deconv_basic,
 *                                lite/tests/utils/naive_math_impl.h:540-642); w as the reference stores it
 *                                [Cin, Cout, kh, kw]; dilation 1, groups 1; optional fused relu and a quantised
 *                                (int8) output for a quantised consumer
 *   pl_add_f32                   lite/backends/cuda/math/elementwise.cu (tensor + tensor, optional relu)
 *   pl_softmax_f32_nhwc          lite/kernels/cuda/softmax_compute.cu, over the channels of every pixel
 *   pl_scale_f32                 lite/kernels/cuda/scale_compute.cc */
int pl_channel_gather_f32_nhwc(int n_src, const float* const* srcs, const int* src_ld, const int32_t* table_dev,
                               void* dst, int ldd, int C, int64_t pixels, int out_dtype, float scale, int round_mode,
                               pl_stream_t stream);
int pl_conv2d_transpose_f32_nhwc(const float* x, const float* w_iohw, const float* bias, void* y, int N, int H, int W,
                                 int Cin, int Cout, int kh, int kw, int stride_h, int stride_w, int pad_t, int pad_l,
                                 int P, int Q, int ldy, int act, float act_arg, int out_dtype, float out_scale,
                                 int round_mode, pl_stream_t stream);
int pl_add_f32(const float* a, const float* b, float* y, int64_t n, int act, pl_stream_t stream);
int pl_softmax_f32_nhwc(const float* x, float* y, int64_t pixels, int C, int ldx, int ldy, pl_stream_t stream);
int pl_scale_f32(const float* x, float* y, int64_t n, float scale, float bias, int bias_after_scale,
                 pl_stream_t stream);

/* N-split GEMM fused with its all-gather through NVSwitch multicast (NVLS): every staged int8 output tile is copied
 * with multimem.st to `C_mc`, the MULTICAST address of this rank's column block; the switch writes it into the
 * same block of every rank bound to the multicast object (this one included).  One copy of the block leaves the
 * GPU instead of one per peer.  `C` is the rank's own (unicast) view of the block: used for the alignment rules
 * only, never written by this call.  int8 output, no fused residual.  Replaces, like pl_gemm_i8_peers, the
 * all-gather a column-parallel fc would run after lite/kernels/arm/fc_compute.cc:117-164. */
int pl_gemm_i8_mcast(const int8_t* A, int lda, const int8_t* Bt, int ldb, void* C, void* C_mc, int ldc, int M, int N,
                     int K, const pl_epilogue* e, pl_stream_t stream);

/* rows x row_bytes block of device memory -> the NVSwitch multicast address of the same block (coalesced
 * multimem.st): the all-gather step for N-split results that have no staged epilogue (fp32 out): every rank
 * broadcasts its own column block after a plain pl_gemm_i8.  16-byte alignment of both bases, pitches and the row. */
int pl_mcast_copy_2d(const void* src, int64_t ld_src_bytes, void* dst_mc, int64_t ld_dst_bytes, int64_t rows,
                     int64_t row_bytes, pl_stream_t stream);

/* ---- introspection ---------------------------------------------------------- */
/* number of kernels this library has launched in this process (bench.py's
 * gpu_launches claim) */
int64_t pl_launch_count(void);
/* developer knobs (PL_* environment variables) are read once per process and cached; this drops the cache so
 * that a changed environment is seen by the next launch (tests that A/B two code paths call it) */
void pl_refresh_knobs(void);
/* the tile plan of the calling thread's most recent tensor-core launch (conv / gemm), for tests and benches that
 * must know which code path ran: plan[0..9] = {mode, BN, BK, stages | k-blocks per barrier group << 8, cta_pairs, tile_pairs, band_rows, grid,
 * filter_resident, epilogue variant}; returns the number of ints written (<= n) */
int pl_last_igemm_plan(int* plan, int n);
const char* pl_version(void);
/* tensor-pipe micro-benchmark: issues `iters` back-to-back 128xBNx32
 * tcgen05.mma kind::i8 per CTA on `ctas` CTAs from resident smem; returns 0
 * and leaves a checksum in out[0] so the work cannot be elided. */
int pl_i8_mma_peak(int ctas, int iters, int bn, int32_t* out, pl_stream_t stream);
/* same, with the operand layout of a 32 / 64 / 128-byte K block (one iteration = row_bytes/32
 * MMAs): measures how the smem swizzle mode and N change the cost of one tcgen05.mma */
int pl_i8_mma_probe(int ctas, int iters, int bn, int row_bytes, int32_t* out, pl_stream_t stream);
/* the same for CTA pairs (clusters of two CTAs, tcgen05.mma.cta_group::2 with M = 256, B split over the two
 * CTAs' shared memories): `ctas` must be even; one MMA counts 2 x 128 x bn x 32 MACs */
int pl_i8_mma_probe_pairs(int ctas, int iters, int bn, int row_bytes, int32_t* out, pl_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* PL_INT8_H_ */
