// Host-side TMA descriptor (CUtensorMap) construction and the argument structs of the GEMM entry points.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b2v.h"

namespace b2v {

// bf16 tensor, 128-byte swizzle, zero fill out of bounds.
int tmap_tiled_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                    const uint32_t* box);
// same, 128-byte swizzle or none (row pitch = box[0] elements; used by the epilogue's TMA stores)
int tmap_tiled_bf16_sw(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                       const uint32_t* box, bool swizzle128);
int tmap_im2col_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                     const int* lower, const int* upper, uint32_t channels_per_pixel, uint32_t pixels_per_column,
                     const uint32_t* elem_strides);

struct GemmArgs {
  int adt = B2V_BF16;       // element type of a / w / residual / out: B2V_BF16 or B2V_F16
  const void* a = nullptr;  // [M, lda]
  int64_t lda = 0;          // 0 -> Kc
  const void* w = nullptr;  // bf16 [N, taps*Kc]
  const float* bias = nullptr;
  const float* rowbias = nullptr;
  int64_t rows_per_group = 0;
  const void* residual = nullptr;  // bf16 [M, ldo]
  void* out = nullptr;             // bf16 [M, ldo]
  int64_t ldo = 0;                 // 0 -> N (N/2 for GEGLU)
  int64_t M = 0;
  int N = 0;
  int Kc = 0;
  int taps = 1;
  int64_t tap_row_stride = 0;
  int64_t rows_per_batch = 0;  // 0 -> M
  int flags = 0;
  // GroupNorm statistics of the OUTPUT, emitted by the epilogue (see gn_stats_ok): per 32-row block and group the fp32
  // {sum, sum of squares} of the values just produced, [M/32][gn_groups][2]. Ignored (left untouched) when the shape
  // does not qualify — ask gn_stats_ok first.
  float* gn_stats = nullptr;
  int gn_groups = 32;
  // LayerNorm fused across two linear layers (see gemm.cu FUSE_RSTAT / FUSE_LNIN):
  //  producer: row_stats[M][row_stats_slots()] {sum, sumsq} partials of every output row
  //  consumer: `a` is the RAW activation, ln_stats its partials (ln_slots per row), `w` = W diag(gamma),
  //            ln_csum[N] = row sums of that (rounded) w, `bias` = b + W beta:  out = rstd (acc - mean csum) + bias
  float* row_stats = nullptr;
  const float* ln_stats = nullptr;
  const float* ln_csum = nullptr;
  int ln_slots = 0;
  float ln_eps = 1e-5f;
};
int row_stats_slots(int64_t M, int N, int64_t rows_per_batch);
// true when gemm_tiled / conv3x3_im2col called with these sizes will fill GemmArgs::gn_stats / ConvArgs::gn_stats
bool gn_stats_ok(int64_t M, int N, int64_t rows_per_batch, int groups);

struct ConvArgs {
  int adt = B2V_BF16;       // element type of x / w / residual / out: B2V_BF16 or B2V_F16
  const void* x = nullptr;  // NHWC
  const void* w = nullptr;  // bf16 [Cout][3][3][Cin]
  const float* bias = nullptr;
  const float* rowbias = nullptr;
  int64_t rows_per_group = 0;
  const void* residual = nullptr;
  void* out = nullptr;  // bf16 [N,P,Q,Cout]
  int N = 0, H = 0, W = 0, Cin = 0, Cout = 0, stride = 1;
  // conv_out: write the first ncfhw_cout channels as [B, ncfhw_cout, F, H, W] of ncfhw_dtype instead of bf16 rows
  // (N = B*F images; the weight is zero-padded to Cout = 64 rows)
  void* ncfhw_out = nullptr;
  int ncfhw_dtype = 0, ncfhw_cout = 0, ncfhw_F = 1;
  float* gn_stats = nullptr;  // as GemmArgs::gn_stats, M = N*P*Q rows
  int gn_groups = 32;
};

int pick_bn(int N, bool geglu, int64_t m_units = 0, int cg = 1);
int gemm_tiled(const GemmArgs& g, cudaStream_t s);
int conv3x3_im2col(const ConvArgs& c, cudaStream_t s);

}  // namespace b2v
