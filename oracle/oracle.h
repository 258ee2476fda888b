/*
 * oracle.h -- CPU restatement of Paddle-Lite's int8 conv / GEMM hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it.  The product path (paddle-lite_b200/)
 * never links or calls this library and has no CPU fallback.
 *
 * Parity status:
 *   conv2d / depthwise_conv2d / fc contraction : pinned against the reference's
 *       own naive conv_basic<int8_t,int> / basic_gemm<int8_t,int>
 *       (lite/tests/utils/naive_math_impl.h:245-453) compiled from where it
 *       lies into oracle/_ref/ (see oracle/Makefile, oracle/ref_shim.cc) and
 *       against KAT-1 {36,72,108,144} of
 *       lite/kernels/cuda/conv_compute_test.cc:232-298.
 *   epilogue (scale, bias, act, round, clamp)  : restated from
 *       lite/backends/arm/math/conv_block_utils.h:3185-3226 (cvt_kernel),
 *       lite/backends/arm/math/gemv_arm_int8.cc:93-143 (write_gemv_out),
 *       lite/backends/arm/math/saturate.h:130-147; the reference compiles that
 *       text only for ARM, so the epilogue is pinned by restatement + the
 *       calib goldens (lite/kernels/arm/calib_compute_test.cc:37-69,
 *       lite/kernels/cuda/calib_compute_cuda_test.cc:35-46).
 *       Since round 2 the last step -- saturate_cast<int8_t>(roundf(v)) -- is the
 *       reference's own object code: saturate.h is compiled into oracle/_ref
 *       (ref_saturate_cast_i8_f32) and orc_round_sat_i8 is checked against it on
 *       every tie and every float neighbour of a tie (tests/test_oracle.py).
 *   pool2d int8 : window / exclusive / average semantics pinned against the
 *       reference's fp32 golden pool_compute_ref
 *       (lite/kernels/cuda/pool_compute_test.cc:150-249, compiled from where it
 *       lies into oracle/_ref/libref_extra.so) on the dequantised integers: max
 *       and avg are identical in fp32, the int8 output is its requantisation.
 *   mul / matmul int8 : accumulators pinned against the kernels-test baseline
 *       matrix_mul + transpose (lite/tests/kernels/matmul_compute_test.cc:23-46)
 *       for all four transpose combinations and the all-ones vector of
 *       lite/kernels/cuda/matmul_compute_test.cc:33-95; the int8 epilogue is the
 *       fc one.  No reference int8 mul/matmul kernel exists on any target.
 *   elementwise_add int8, fused residual : no reference kernel exists on any
 *       target -- "parity unpinned", defined here by analogy.
 *
 * All tensors are host pointers.  Activations are NHWC, filters OHWI
 * ([K, R, S, C/groups]) -- the kCUDA int8 layout of
 * lite/kernels/cuda/conv_compute.cc:55-106.
 */
#ifndef PL_ORACLE_H_
#define PL_ORACLE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* lite_api::ActivationType values (lite/api/paddle_place.h:100-114) */
enum { ORC_ACT_NONE = 0, ORC_ACT_RELU = 1, ORC_ACT_RELU6 = 2, ORC_ACT_LEAKY = 4 };
/* output element type */
enum { ORC_OUT_INT8 = 0, ORC_OUT_FP32 = 1, ORC_OUT_INT32 = 2 };
/* rounding / clamping convention (SURVEY.md section 8 header) */
enum { ORC_ROUND_ARM = 0, ORC_ROUND_CUDA_LEGACY = 1 };
/* pooling */
enum { ORC_POOL_MAX = 0, ORC_POOL_AVG = 1 };

/* ---- scale folding: lite/kernels/arm/conv_gemmlike.cc:207-264,
 *      lite/kernels/arm/fc_compute.cc:117-164,
 *      lite/backends/cuda/math/cudnn_conv.cc:349-372 ------------------------- */
int orc_fold_scales(const float* w_scale, int n_w_scale, int n_out,
                    float in_scale, float out_scale, int out_dtype,
                    const float* bias /* nullable, [n_out] */,
                    float* scale_folded /* [n_out] */,
                    float* bias_folded /* [n_out], zeros when bias==NULL */,
                    int act, float act_arg, float* act_arg_folded);

/* ---- scalar epilogue on one accumulator ---------------------------------- */
float orc_epilogue_f32(int32_t acc, float scale, float bias, int act,
                       float act_arg, int round_mode, int vec4_path);
int8_t orc_round_sat_i8(float v, int round_mode);

/* ---- exact int32 contractions --------------------------------------------- */
int orc_conv2d_i8_acc(const int8_t* x, const int8_t* w, int32_t* acc, int N,
                      int H, int W, int C, int K, int R, int S, int sh, int sw,
                      int pt, int pl, int dh, int dw, int groups, int P, int Q);
/* B given as [K,N] row-major (b_is_nk==0, the fc W layout) or [N,K] (==1). */
int orc_gemm_i8_acc(const int8_t* A, const int8_t* B, int32_t* acc, int M,
                    int N, int K, int trans_a, int b_is_nk);

/* ---- epilogue over an [M, n_out] accumulator matrix ----------------------- */
int orc_epilogue(const int32_t* acc, int64_t M, int n_out, const float* scale,
                 const float* bias, int act, float act_arg, int out_dtype,
                 int round_mode, void* y);

int orc_epilogue_res(const int32_t* acc, int64_t M, int n_out, const float* scale,
                     const float* bias, const int8_t* res, int ld_res,
                     float res_scale, int act, float act_arg, int out_dtype,
                     int round_mode, void* y);

/* ---- whole ops (contraction + folded epilogue) ----------------------------- */
int orc_conv2d_i8_nhwc(const int8_t* x, const int8_t* w, const float* bias_f,
                       const float* scale_f, void* y, int N, int H, int W,
                       int C, int K, int R, int S, int sh, int sw, int pt,
                       int pb, int pl, int pr, int dh, int dw, int groups,
                       int act, float act_arg, int out_dtype, int round_mode);
int orc_gemm_i8(const int8_t* A, const int8_t* B, const float* bias_f,
                const float* scale_f, void* y, int M, int N, int K,
                int trans_a, int b_is_nk, int act, float act_arg,
                int out_dtype, int round_mode);

/* ---- calib: lite/backends/arm/math/type_trans.cc:34-187,268-...;
 *      lite/kernels/cuda/calib_compute.cu:27-95 ------------------------------ */
int orc_quantize_f32_i8(const float* x, int8_t* y, int64_t n, float scale,
                        int round_mode);
int orc_dequantize_i8_f32(const int8_t* x, float* y, int64_t n, float scale);

/* ---- layout: lite/kernels/cuda/layout_compute.cc:41-94 --------------------- */
int orc_nchw_to_nhwc(const void* x, void* y, int N, int C, int H, int W,
                     int elem_bytes);
int orc_nhwc_to_nchw(const void* x, void* y, int N, int C, int H, int W,
                     int elem_bytes);

/* ---- int8 pooling / residual add (parity unpinned, see header) ------------- */
int orc_pool2d_i8_nhwc(const int8_t* x, void* y, int N, int H, int W, int C,
                       int kh, int kw, int sh, int sw, int pt, int pl, int P,
                       int Q, int pool_type, int exclusive, float rescale,
                       int out_dtype, int round_mode);
int orc_add_i8(const int8_t* a, const int8_t* b, void* y, int64_t n, float sa,
               float sb, int act, float act_arg, int out_dtype, int round_mode);

/* ---- channel movement between int8 ops, NCHW like the reference:
 * pure byte moves, so the int8 versions are the reference's memcpy loops on 1-byte elements */
int orc_concat_nchw(int n_in, const void* const* xs, const int* cs, void* y,
                    int N, int64_t inner_bytes);
int orc_split_nchw(const void* x, int n_out, void* const* ys, const int* cs,
                   int N, int64_t inner_bytes);
int orc_shuffle_channel_nchw(const void* x, void* y, int group, int N, int C,
                             int64_t spatial_bytes);
int orc_relu_i8(const int8_t* x, int8_t* y, int64_t n, int hi);

int orc_conv_out_size(int in, int k, int dil, int pad0, int pad1, int stride);
int orc_num_threads(void);

#ifdef __cplusplus
}
#endif
#endif /* PL_ORACLE_H_ */
