// extern "C" wrappers of the building-block ops (include/b2v.h, "Building-block ops"); tests/ drive every
// kernel through these against the oracle.
#include "common.cuh"
#include "ops.h"

using namespace b2v;

extern "C" int b2v_op_gemm(const void* a_dev, const void* w_dev, const float* bias_dev, const float* rowbias_dev,
                           int64_t rows_per_group, const void* residual_dev, void* out_dev, int64_t M, int N, int Kc,
                           int taps, int64_t tap_row_stride, int64_t rows_per_batch, int flags, int act_dtype,
                           void* stream) {
  B2V_CHECK(a_dev && w_dev && out_dev, "null argument");
  B2V_TRY(b2v_device_check());
  GemmArgs g;
  g.adt = act_dtype;
  g.a = a_dev; g.w = w_dev; g.bias = bias_dev; g.rowbias = rowbias_dev; g.rows_per_group = rows_per_group;
  g.residual = residual_dev; g.out = out_dev; g.M = M; g.N = N; g.Kc = Kc; g.taps = taps;
  g.tap_row_stride = tap_row_stride; g.rows_per_batch = rows_per_batch; g.flags = flags;
  return gemm_tiled(g, (cudaStream_t)stream);
}

extern "C" int b2v_row_stats_slots(int64_t M, int N, int64_t rows_per_batch) { return row_stats_slots(M, N, rows_per_batch); }

extern "C" int b2v_op_gemm_rowstats(const void* a_dev, const void* w_dev, const float* bias_dev, const void* residual_dev,
                                    void* out_dev, int64_t M, int N, int Kc, int act_dtype, float* row_stats_dev,
                                    void* stream) {
  B2V_CHECK(a_dev && w_dev && out_dev && row_stats_dev, "null argument");
  B2V_TRY(b2v_device_check());
  B2V_CHECK(row_stats_slots(M, N, 0) > 0, "this GEMM shape has no row-statistics epilogue");
  GemmArgs g;
  g.adt = act_dtype;
  g.a = a_dev; g.w = w_dev; g.bias = bias_dev; g.residual = residual_dev; g.out = out_dev; g.M = M; g.N = N; g.Kc = Kc;
  g.row_stats = row_stats_dev;
  return gemm_tiled(g, (cudaStream_t)stream);
}

extern "C" int b2v_op_gemm_lnin(const void* a_raw_dev, const void* w_gamma_dev, const float* bias_folded_dev,
                                const float* csum_dev, const float* ln_stats_dev, int ln_slots, float eps, void* out_dev,
                                int64_t M, int N, int Kc, int flags, int act_dtype, void* stream) {
  B2V_CHECK(a_raw_dev && w_gamma_dev && bias_folded_dev && csum_dev && ln_stats_dev && out_dev, "null argument");
  B2V_TRY(b2v_device_check());
  GemmArgs g;
  g.adt = act_dtype;
  g.a = a_raw_dev; g.w = w_gamma_dev; g.bias = bias_folded_dev; g.out = out_dev; g.M = M; g.N = N; g.Kc = Kc;
  g.flags = flags;
  g.ln_stats = ln_stats_dev; g.ln_csum = csum_dev; g.ln_slots = ln_slots; g.ln_eps = eps;
  return gemm_tiled(g, (cudaStream_t)stream);
}

extern "C" int b2v_gn_stats_ok(int64_t M, int N, int64_t rows_per_batch, int groups) {
  return gn_stats_ok(M, N, rows_per_batch, groups) ? 1 : 0;
}

extern "C" int b2v_op_gemm_gnstats(const void* a_dev, const void* w_dev, const float* bias_dev, const float* rowbias_dev,
                                   int64_t rows_per_group, const void* residual_dev, void* out_dev, int64_t M, int N,
                                   int Kc, int taps, int64_t tap_row_stride, int64_t rows_per_batch, int act_dtype,
                                   float* gn_stats_dev, int gn_groups, void* stream) {
  B2V_CHECK(a_dev && w_dev && out_dev && gn_stats_dev, "null argument");
  B2V_TRY(b2v_device_check());
  B2V_CHECK(gn_stats_ok(M, N, rows_per_batch, gn_groups), "this GEMM shape does not emit GroupNorm statistics (b2v_gn_stats_ok)");
  GemmArgs g;
  g.adt = act_dtype;
  g.a = a_dev; g.w = w_dev; g.bias = bias_dev; g.rowbias = rowbias_dev; g.rows_per_group = rows_per_group;
  g.residual = residual_dev; g.out = out_dev; g.M = M; g.N = N; g.Kc = Kc; g.taps = taps;
  g.tap_row_stride = tap_row_stride; g.rows_per_batch = rows_per_batch;
  g.gn_stats = gn_stats_dev; g.gn_groups = gn_groups;
  return gemm_tiled(g, (cudaStream_t)stream);
}

extern "C" int b2v_op_conv3x3_gnstats(const void* x_dev, const void* w_dev, const float* bias_dev,
                                      const float* rowbias_dev, int64_t rows_per_group, const void* residual_dev,
                                      void* out_dev, int N, int H, int W, int Cin, int Cout, int act_dtype,
                                      float* gn_stats_dev, int gn_groups, void* stream) {
  B2V_CHECK(x_dev && w_dev && out_dev && gn_stats_dev, "null argument");
  B2V_TRY(b2v_device_check());
  B2V_CHECK(gn_stats_ok((int64_t)N * H * W, Cout, 0, gn_groups), "this conv shape does not emit GroupNorm statistics");
  ConvArgs c;
  c.adt = act_dtype;
  c.x = x_dev; c.w = w_dev; c.bias = bias_dev; c.rowbias = rowbias_dev; c.rows_per_group = rows_per_group;
  c.residual = residual_dev; c.out = out_dev; c.N = N; c.H = H; c.W = W; c.Cin = Cin; c.Cout = Cout; c.stride = 1;
  c.gn_stats = gn_stats_dev; c.gn_groups = gn_groups;
  return conv3x3_im2col(c, (cudaStream_t)stream);
}

extern "C" int b2v_op_group_norm_pre(const void* x_dev, int C, const float* gamma_dev, const float* beta_dev, void* y_dev,
                                     int64_t n_items, int64_t rows_per_item, int groups, float eps, int silu,
                                     void* scratch_dev, const float* pre_dev, int act_dtype, void* stream) {
  B2V_CHECK(x_dev && gamma_dev && beta_dev && y_dev && scratch_dev && pre_dev, "null argument");
  B2V_TRY(b2v_device_check());
  return group_norm(act_dtype, x_dev, C, nullptr, 0, gamma_dev, beta_dev, y_dev, n_items, rows_per_item, groups, eps, silu,
                    scratch_dev, (cudaStream_t)stream, pre_dev);
}

extern "C" size_t b2v_op_conv3x3_scratch_bytes(int N, int H, int W, int Cin, int stride, int mode) {
  if (mode == 0) return 0;
  const int64_t P = (H - 1) / stride + 1, Q = (W - 1) / stride + 1;
  return (size_t)N * P * Q * 9 * Cin * 2;
}

extern "C" int b2v_op_conv3x3(const void* x_dev, const void* w_dev, const float* bias_dev, const float* rowbias_dev,
                              int64_t rows_per_group, const void* residual_dev, void* out_dev, int N, int H, int W,
                              int Cin, int Cout, int stride, int mode, void* scratch_dev, int act_dtype,
                              void* stream) {
  B2V_CHECK(x_dev && w_dev && out_dev, "null argument");
  B2V_TRY(b2v_device_check());
  cudaStream_t s = (cudaStream_t)stream;
  if (mode == 0) {
    ConvArgs c;
    c.adt = act_dtype;
    c.x = x_dev; c.w = w_dev; c.bias = bias_dev; c.rowbias = rowbias_dev; c.rows_per_group = rows_per_group;
    c.residual = residual_dev; c.out = out_dev; c.N = N; c.H = H; c.W = W; c.Cin = Cin; c.Cout = Cout; c.stride = stride;
    return conv3x3_im2col(c, s);
  }
  B2V_CHECK(scratch_dev, "explicit im2col needs scratch");
  B2V_TRY(im2col3x3(x_dev, scratch_dev, N, H, W, Cin, stride, s));
  const int64_t P = (H - 1) / stride + 1, Q = (W - 1) / stride + 1;
  GemmArgs g;
  g.adt = act_dtype;
  g.a = scratch_dev; g.w = w_dev; g.bias = bias_dev; g.rowbias = rowbias_dev; g.rows_per_group = rows_per_group;
  g.residual = residual_dev; g.out = out_dev; g.M = (int64_t)N * P * Q; g.N = Cout; g.Kc = 9 * Cin; g.taps = 1;
  return gemm_tiled(g, s);
}

extern "C" int b2v_op_group_norm(const void* x_dev, int C1, const void* x2_dev, int C2, const float* gamma_dev,
                                 const float* beta_dev, void* y_dev, int64_t n_items, int64_t rows_per_item,
                                 int groups, float eps, int silu, void* stats_dev, int act_dtype, void* stream) {
  B2V_CHECK(x_dev && gamma_dev && beta_dev && y_dev && stats_dev, "null argument");
  B2V_TRY(b2v_device_check());
  return group_norm(act_dtype, x_dev, C1, x2_dev, C2, gamma_dev, beta_dev, y_dev, n_items, rows_per_item, groups, eps, silu,
                    stats_dev, (cudaStream_t)stream);
}

extern "C" size_t b2v_op_group_norm_scratch_bytes(int64_t n_items, int64_t rows_per_item, int groups) {
  return group_norm_scratch_bytes(n_items, rows_per_item, groups);
}

extern "C" int b2v_op_layer_norm(const void* x_dev, const float* gamma_dev, const float* beta_dev, void* y_dev,
                                 int64_t rows, int C, float eps, int act_dtype, void* stream) {
  B2V_CHECK(x_dev && gamma_dev && beta_dev && y_dev, "null argument");
  B2V_TRY(b2v_device_check());
  return layer_norm(act_dtype, x_dev, gamma_dev, beta_dev, y_dev, rows, C, eps, (cudaStream_t)stream);
}

extern "C" int b2v_op_attention(const void* q_dev, int64_t ldq, const void* k_dev, const void* v_dev, int64_t ldkv,
                                void* out_dev, int64_t ldo, int nb, int heads, int Sq, int Skv, int kv_group,
                                float scale, int act_dtype, void* stream) {
  B2V_TRY(b2v_device_check());
  AttnArgs a;
  a.adt = act_dtype;
  a.q = q_dev; a.k = k_dev; a.v = v_dev; a.out = out_dev; a.ldq = ldq; a.ldkv = ldkv; a.ldo = ldo;
  a.nb = nb; a.heads = heads; a.Sq = Sq; a.Skv = Skv; a.kv_group = kv_group; a.scale = scale;
  return attention(a, (cudaStream_t)stream);
}

extern "C" int b2v_op_temporal_attention(const void* qkv_dev, int64_t ld, int q_off, int k_off, int v_off,
                                         void* out_dev, int64_t ldo, int B, int F, int64_t HW, int heads, float scale,
                                         int act_dtype, void* stream) {
  B2V_TRY(b2v_device_check());
  return temporal_attention(act_dtype, qkv_dev, ld, q_off, k_off, v_off, out_dev, ldo, B, F, HW, heads, scale,
                            (cudaStream_t)stream);
}
