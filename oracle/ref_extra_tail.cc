// ref_extra_tail.cc -- C-ABI tail of oracle/_ref/libref_extra.so (see ref_extra_prelude.h).
}  // namespace lite
}  // namespace paddle

extern "C" {
// fp32 NCHW pooling golden of the reference's kCUDA pool test
void ref_pool_compute(const float* x, float* y, int n, int c, int h, int w, int oh, int ow, int kh, int kw,
                      int sh, int sw, int pad_t, int pad_b, int pad_l, int pad_r, int is_avg, int exclusive,
                      int global_pooling) {
  using namespace paddle::lite;
  Tensor tx, ty;
  tx.dims_.d = {n, c, h, w};
  ty.dims_.d = {n, c, oh, ow};
  tx.ptr = const_cast<float*>(x);
  ty.ptr = y;
  operators::PoolParam p;
  p.x = &tx;
  p.output = &ty;
  p.pooling_type = is_avg ? "avg" : "max";
  p.ksize = {kh, kw};
  p.strides = {sh, sw};
  p.paddings = std::make_shared<std::vector<int>>(std::vector<int>{pad_t, pad_b, pad_l, pad_r});
  p.exclusive = exclusive != 0;
  p.global_pooling = global_pooling != 0;
  pool_compute_ref(p);
}
// out[m, n] = sum_k x[m, k] * y[k, n] * alpha  (the kernels-test baseline)
void ref_matrix_mul(int m, int k, int n, float alpha, const float* x, const float* y, float* out) {
  paddle::lite::matrix_mul(m, k, n, alpha, x, y, out);
}
void ref_transpose(int m, int n, const float* src, float* dst) { paddle::lite::transpose(m, n, src, dst); }
}
