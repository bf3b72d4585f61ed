// Internal (C++) launch API of the kernels; capi.cu exposes the tested subset through include/b2v.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "tmap.h"

namespace b2v {

// norm.cu
// adt = element type of the activations (B2V_BF16 or B2V_F16); upsample / concat / im2col move 2-byte elements blindly
int group_norm(int adt, const void* x1, int C1, const void* x2, int C2, const float* gamma, const float* beta, void* y,
               int64_t n_items, int64_t rows_per_item, int groups, float eps, int silu, void* scratch,
               cudaStream_t s, const float* pre = nullptr);
// pre: {sum, sumsq} per 32-row block and group, [n_items * rows_per_item / 32][groups][2] fp32, as emitted by the GEMM
// epilogue that produced x1 (GemmArgs::gn_stats): the statistics pass is skipped (x2 must be null)
size_t group_norm_scratch_bytes(int64_t n_items, int64_t rows_per_item, int groups);
int layer_norm(int adt, const void* x, const float* gamma, const float* beta, void* y, int64_t rows, int C, float eps,
               cudaStream_t s);
int upsample2x(const void* x, void* y, int N, int H, int W, int C, cudaStream_t s);
int concat_channels(const void* a, int Ca, const void* b, int Cb, void* y, int64_t rows, cudaStream_t s);
int im2col3x3(const void* x, void* col, int N, int H, int W, int C, int stride, cudaStream_t s);

// attention.cu
struct AttnArgs {
  int adt = B2V_BF16;       // element type of q / k / v / out: B2V_BF16 or B2V_F16
  const void* q = nullptr;  // row (b, s) at q + (b*Sq + s)*ldq, head h at column h*64
  const void* k = nullptr;  // bf16, row (b/kv_group, s) at k + ((b/kv_group)*Skv + s)*ldkv
  const void* v = nullptr;
  void* out = nullptr;
  int64_t ldq = 0, ldkv = 0, ldo = 0;
  int nb = 0, heads = 0, Sq = 0, Skv = 0, kv_group = 1;
  float scale = 0.125f;
};
int attention(const AttnArgs& a, cudaStream_t s);
int temporal_attention(int adt, const void* qkv, int64_t ld, int q_off, int k_off, int v_off, void* out, int64_t ldo, int B,
                       int F, int64_t HW, int heads, float scale, cudaStream_t s);

// small.cu : SIMT kernels for the thin ends of the network
int set_scalar(float* dst_dev, float v, cudaStream_t s);
// timestep_dev: device float (set by set_scalar outside the replayed graph)
int time_embedding(const float* timestep_dev, int B, int dim0, const float* w1, const float* b1, const float* w2, const float* b2,
                   int temb_dim, float* silu_emb_out /*[B,temb_dim] = SiLU(emb)*/, cudaStream_t s);
// out[b, n] = bias[n] + sum_k W[n,k] * in[b,k]  (fp32, tiny M)
int small_linear(const float* in, const float* w, const float* bias, float* out, int B, int N, int K, cudaStream_t s);
// NCFHW (io dtype) -> channels-last rows [B*F*HW, Cpad] of the activation dtype, zero padded channels (input of the
// tensor-core conv_in)
int ncfhw_to_rows(const void* x, int io_dtype, void* y, int adt, int B, int Cin, int F, int64_t HW, int Cpad,
                  cudaStream_t s);
// dtype conversion of the text embeddings to rows of the activation dtype
int convert_to_act(const void* x, int dtype, void* y, int adt, int64_t n, cudaStream_t s);

}  // namespace b2v
