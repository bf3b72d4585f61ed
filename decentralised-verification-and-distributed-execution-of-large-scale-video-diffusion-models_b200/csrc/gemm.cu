// tcgen05 / TMEM / TMA GEMM for every dense contraction of the UNet3D
// ([UPSTREAM diffusers] UNet3DConditionModel as called at Distribution/strategies/chunk_only.py:71):
//   * linear layers (q/k/v/out projections, GEGLU feed-forward, proj_in/out)        taps = 1
//   * TemporalConvLayer Conv3d (3,1,1), zero padded in time                           taps = 3
//   * ResnetBlock2D / Down / Upsample Conv2d 3x3 as an implicit GEMM (TMA im2col)   taps = 9
// out[M,N] = sum_tap A_tap[M,Kc] · W[N, tap*Kc:(tap+1)*Kc]^T  (+bias, +row-group bias, +residual | GEGLU)
//
// Kernel shape (persistent, warp specialised, one CTA per SM, 320 threads):
//   warp 0   : TMA producer — A tile {128 rows x 64 ch} + W tile {BN rows x 64 ch} per k-block into a
//              128B-swizzled shared-memory ring (STAGES deep), completion on `full` mbarriers
//   warp 1   : MMA issuer  — one elected lane issues 4 x tcgen05.mma (128 x BN x 16) per k-block into a
//              double-buffered TMEM accumulator; tcgen05.commit releases ring slots / publishes the tile
//   warps 2-9: epilogue    — tcgen05.ld the accumulator (lane = output row; two warps per TMEM lane quarter, one per
//              column half), software-pipelined over 16-column chunks, fused bias / time-embedding row bias /
//              residual (prefetched) / GEGLU with a branch-free erf, bf16 pack, 32-byte row-contiguous global stores
#include <stdlib.h>

#include <mutex>

#include "common.cuh"
#include "ptx.cuh"
#include "tmap.h"

namespace b2v {
namespace {

using namespace ptx;

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int A_BYTES = BM * BK * 2;  // 16 KiB
constexpr int NUM_THREADS = 320;  // TMA warp + MMA warp + 8 epilogue warps
constexpr int TMEM_COLS = 512;

constexpr int EPI_WARPS = 8;
constexpr int SMEM_LIMIT = 227 * 1024;

// Shared-memory plan of one instantiation: [operand ring: STAGES x (A 16 KiB + B)] [epilogue staging: one output panel
// per epilogue warp] [barriers]. An epilogue warp owns 32 rows x GROUP_COLS columns of the tile and hands them to the
// TMA store engine in panels of PANEL_COLS columns: 64 (128-byte rows, 128B swizzle, conflict-free writes) when the
// group width allows it, else the whole group width unswizzled (BN = 160 -> 80 columns, BN = 64 -> 32).
// CG = CTAs per tile: 2 = CTA pair (tcgen05 cta_group::2): the pair computes a 256-row tile, each CTA stages its own 128
// rows of A and HALF of the B tile (BN/2 rows) per k-block.
template <int BN, bool GEGLU, bool RES, int CG>
struct TileCfg {
  static constexpr int B_BYTES = (BN / CG) * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int ACC_STRIDE = (BN > 128) ? 256 : 128;  // TMEM columns between the two accumulators
  static constexpr int OUT_COLS = GEGLU ? BN / 2 : BN;
  static constexpr int GROUP_COLS = OUT_COLS / 2;
  static constexpr int PANEL_COLS = (GROUP_COLS % 64 == 0) ? 64 : GROUP_COLS;
  static constexpr bool PANEL_SWIZZLE = PANEL_COLS == 64;
  static constexpr int PANEL_BYTES = PANEL_COLS * 2 * 32;
  static constexpr int NPANELS = GROUP_COLS / PANEL_COLS;
  static constexpr int NBUF = (RES && NPANELS == 1) ? 2 : 1;  // residual of the next tile prefetched into the other buffer
  static constexpr int EPI_BYTES = EPI_WARPS * NBUF * PANEL_BYTES;
  static constexpr int FIXED = EPI_BYTES + 512 /*barriers*/ + 1024 /*alignment slack*/;
  static constexpr int STAGES_FIT = (SMEM_LIMIT - FIXED) / STAGE_BYTES;
  static constexpr int STAGES = STAGES_FIT > 8 ? 8 : STAGES_FIT;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + FIXED;
  static_assert(STAGES >= 4, "operand ring too shallow");
  static_assert(PANEL_BYTES % 1024 == 0, "panel buffers must keep the 1 KiB swizzle alignment");
};

struct GemmParams {
  int64_t M;               // output rows in total
  int N;                   // GEMM N (before GEGLU halving)
  int Kc;                  // channels per tap
  int taps;                // 1 | 3 | 9
  int64_t rows_per_batch;  // tiled mode: rows of one z-slab (tiles never straddle slabs)
  int64_t tap_row_stride;  // tiled mode, 3 taps: rows per frame
  int m_units_per_batch;  // units of CG x 128 rows per slab
  int num_m_units;
  int num_n_tiles;
  // im2col geometry (output P x Q per image, conv stride)
  int P, Q, conv_stride;
  // epilogue
  const float* bias;        // [N] (GEGLU: packed like W)
  const float* rowbias;     // [M / rows_per_group, N] or null
  int64_t rows_per_group;
  const void* residual;  // [M, ldo] or null (2-byte activations)
  void* out;
  int64_t ldo;
  int gelu;
  // EPI_NCFHW (conv_out): out_raw[((b*nc_cout + co)*nc_F + f)*nc_HW + pix] of dtype nc_dtype, row = (b*nc_F + f)*nc_HW + pix
  void* out_raw;
  int nc_dtype, nc_cout, nc_F;
  int64_t nc_HW;
  // GroupNorm statistics of the output (GNG > 0 instantiations): [M/32][gn_groups]{sum, sumsq}
  float* gn_stats;
  int gn_groups;
  // LayerNorm fused across two GEMMs (FUSE_RSTAT / FUSE_LNIN instantiations):
  //  producer: row_stats[M][2 * num_n_tiles]{sum, sumsq} of each output row over the columns of one epilogue warp
  //  consumer: A is the RAW activation; ln_stats[M][ln_slots] are its row partials, ln_csum[N] = sum_k W'[n][k] with
  //            W' = W diag(gamma) (what `w` holds), bias = b + W beta:  out = rstd (acc - mean csum) + bias
  float* row_stats;
  const float* ln_stats;
  const float* ln_csum;
  int ln_slots;
  float ln_eps;
};
constexpr int FUSE_RSTAT = 64, FUSE_LNIN = 128;  // FUSE template argument: 0 | GNG (2, 4, 8, 16) | FUSE_RSTAT | FUSE_LNIN

constexpr int EPI_PLAIN = 0, EPI_GEGLU_T = 1, EPI_NCFHW = 2;

// gelu_erf(x) = 0.5 x (1 + erf(x / sqrt 2)) with erfc(z) = 2^-Q(z), Q a degree-5 polynomial without constant term
// fitted (minimax, tools/fit_gelu.py) on z in [0, 5]: |erfc error| <= 6.3e-7 in fp32, |gelu error| <= 1.1e-6 — far
// below the bf16 rounding of the output. ONE MUFU + 10 FMA-pipe instructions, no branches (erff() is ~40 with two
// branches; the GEGLU epilogue of the L0 feed-forward was bound by it). The negative side uses 1 + erf = erfc(|z|)
// directly, so there is no cancellation for x << 0. Coefficients are pre-scaled to take |x| instead of z = |x|/sqrt 2.
__device__ __forceinline__ float gelu_fast(float x) {
  const float ax = fabsf(x);
  float q = fmaf(5.204590389e-04f, ax, -7.397513200e-03f);
  q = fmaf(q, ax, 5.256124131e-02f);
  q = fmaf(q, ax, 4.592546802e-01f);
  q = fmaf(q, ax, 1.151091355e+00f);
  float pe;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(pe) : "f"(-q * ax));  // erfc(|x| / sqrt 2)
  const float phi2 = x < 0.f ? pe : 2.0f - pe;                    // 1 + erf(x / sqrt 2)
  return 0.5f * x * phi2;
}

// The same function for a PAIR of gates on packed fp32 (FFMA2 / FMUL2: half the issue slots — the GEGLU epilogue issues
// about as many warp-instructions per tile as the K = 320 mainloop has tensor cycles), written without the sign select:
//   gelu(x) = relu(x) - 0.5 |x| erfc(|x| / sqrt 2)      (x > 0: x (1 - erfc/2) = x (1 + erf)/2;  x < 0: x erfc/2)
// Coefficients are those of gelu_fast, negated so the exponent needs no extra sign flip.
__device__ __forceinline__ uint64_t gelu_fast2(uint64_t x2) {
  float x0, x1;
  unpack2(x2, x0, x1);
  const uint64_t ax2 = pack2(fabsf(x0), fabsf(x1));
  uint64_t q = fma2(pack2(-5.204590389e-04f, -5.204590389e-04f), ax2, pack2(7.397513200e-03f, 7.397513200e-03f));
  q = fma2(q, ax2, pack2(-5.256124131e-02f, -5.256124131e-02f));
  q = fma2(q, ax2, pack2(-4.592546802e-01f, -4.592546802e-01f));
  q = fma2(q, ax2, pack2(-1.151091355e+00f, -1.151091355e+00f));
  float e0, e1;
  unpack2(mul2(q, ax2), e0, e1);
  float p0, p1;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p0) : "f"(e0));  // erfc(|x| / sqrt 2)
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p1) : "f"(e1));
  const uint64_t nh2 = mul2(ax2, pack2(-0.5f, -0.5f));
  return fma2(nh2, pack2(p0, p1), pack2(fmaxf(x0, 0.f), fmaxf(x1, 0.f)));
}

// Epilogue warp role. One warp handles 32 rows (its TMEM lane quarter) x half of the tile's output columns (group 0 / 1)
// in 16-column chunks: tcgen05.ld of chunk c+1 is in flight while chunk c is computed. Results are staged in the warp's
// own shared-memory panel and leave as TMA stores (full 128-byte lines instead of 16-byte scattered stores: the L0
// GEMMs were bound by the LSU/L2 request rate of the latter). With RES the residual panel ARRIVES in the same buffer by
// TMA load and is updated in place; when a tile is one panel wide (all but BN = 256) two buffers alternate so the
// residual of the NEXT tile is already in flight while this one is processed.
// GNG > 0: the columns of this warp (half a tile: 80 / 64 / 128) are GNG whole GroupNorm groups of the OUTPUT tensor. Every lane
// accumulates {sum, sum of squares} of its row per group while the values pass through its registers, the warp folds the
// 32 rows with a fixed-order shuffle tree (deterministic) and writes 2 GNG floats per 32-row block: the consumer's
// GroupNorm needs no statistics pass over the activation (norm.cu: group_norm with `pre`).
template <int BN, int MODE, bool GEGLU, bool RES, int CG, bool F16, int FUSE>
__device__ __forceinline__ void epilogue_loop(const GemmParams& p, const CUtensorMap* tmO, const CUtensorMap* tmR,
                                              uint8_t* stage0, uint32_t rbar0, uint32_t tmem_base, uint32_t tfull0,
                                              uint32_t tempty0, int warp, int lane, int num_tiles) {
  using Cfg = TileCfg<BN, GEGLU, RES, CG>;
  constexpr int GNG = FUSE & 31;
  constexpr bool RSTAT = FUSE == FUSE_RSTAT, LNIN = FUSE == FUSE_LNIN;
  constexpr int OUT_COLS = Cfg::OUT_COLS;
  constexpr int NCH = OUT_COLS / 32;  // 16-column chunks per group
  // work units: tile `u` of this CTA (CG = 1) or of this CTA pair (CG = 2: rank r owns rows [128 r, +128) of the unit)
  const int rank = CG == 2 ? (int)(blockIdx.x & 1) : 0;
  const int unit0 = CG == 2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int ustride = CG == 2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const uint32_t tempty_leader = CG == 2 ? mapa_shared(tempty0, 0) : tempty0;
  constexpr int CH_PER_PANEL = Cfg::PANEL_COLS / 16;
  constexpr int NBUF = Cfg::NBUF;
  const int quarter = warp & 3;
  const int group = (warp - 2) >> 2;
  const int c_base = group * NCH * 16;
  const int lane_off = lane * (Cfg::PANEL_COLS * 2);
  const uint32_t stage_u32 = smem_u32(stage0);

  // TMA coordinates of this warp's part of a tile: (first column, first row inside the slab, slab)
  auto tile_col = [&](int tile) { return (tile % p.num_n_tiles) * OUT_COLS + c_base; };
  auto tile_row0 = [&](int tile) {
    const int m_unit = tile / p.num_n_tiles;
    return ((MODE == 0 ? (m_unit % p.m_units_per_batch) : m_unit) * CG + rank) * BM + quarter * 32;
  };
  auto tile_slab = [&](int tile) { return MODE == 0 ? (tile / p.num_n_tiles) / p.m_units_per_batch : 0; };
  auto load_residual = [&](int tile, int panel, int buf) {  // one lane
    const uint32_t bar = rbar0 + 8 * buf;
    mbar_expect_tx(bar, Cfg::PANEL_BYTES);
    tma_load_3d(smem_u32(stage0 + buf * Cfg::PANEL_BYTES), tmR, bar, tile_col(tile) + panel * Cfg::PANEL_COLS,
                tile_row0(tile), tile_slab(tile));
  };

  uint32_t q = 0;  // panels processed by this warp so far: buffer q % NBUF, residual barrier parity (q / NBUF) & 1
  if (RES && NBUF == 2 && lane == 0 && unit0 < num_tiles) load_residual(unit0, 0, 0);
  int acc = 0;
  uint32_t acc_phase = 0;
  for (int tile = unit0; tile < num_tiles; tile += ustride) {
    const int n_tile = tile % p.num_n_tiles;
    const int row0 = tile_row0(tile), slab = tile_slab(tile);
    const int ocol0 = tile_col(tile);
    const int wcol0 = n_tile * BN + c_base;  // column in the (packed) weight / bias space
    const float* rb = nullptr;
    if (!GEGLU && p.rowbias) {
      const int64_t grow = (MODE == 0 ? (int64_t)slab * p.rows_per_batch : 0) + row0 + lane;
      const int64_t gclamped = grow < p.M ? grow : p.M - 1;  // rows past the end are clipped by the store
      rb = p.rowbias + (gclamped / p.rows_per_group) * (int64_t)p.N + wcol0;
    }
    const uint32_t tacc = tmem_base + acc * Cfg::ACC_STRIDE + ((uint32_t)(quarter * 32) << 16);
    // this lane's output row (LayerNorm fusion: its statistics in, or its partial sums out)
    const int64_t lrow = (MODE == 0 ? (int64_t)slab * p.rows_per_batch : 0) + row0 + lane;
    float ln_rstd = 1.f, ln_nrm = 0.f;  // LNIN: out = rstd * acc + (bias' - rstd * mean * csum)
    if constexpr (LNIN) {
      const float2* st = reinterpret_cast<const float2*>(p.ln_stats) + (lrow < p.M ? lrow : p.M - 1) * p.ln_slots;
      float sm = 0.f, sq = 0.f;
      for (int i = 0; i < p.ln_slots; ++i) {  // fixed order
        const float2 t = __ldg(st + i);
        sm += t.x;
        sq += t.y;
      }
      const float inv_k = 1.0f / (float)p.Kc;
      const float mean = sm * inv_k;
      const float var = fmaxf(fmaf(-mean, mean, sq * inv_k), 0.f);
      ln_rstd = rsqrtf(var + p.ln_eps);
      ln_nrm = -ln_rstd * mean;
    }
    float rsum = 0.f, rsq = 0.f;  // RSTAT
    float gsum[GNG > 0 ? GNG : 1], gsq[GNG > 0 ? GNG : 1];
#pragma unroll
    for (int i = 0; i < (GNG > 0 ? GNG : 1); ++i) gsum[i] = gsq[i] = 0.f;
    uint32_t v[2][16];
    uint32_t g[GEGLU ? 2 : 1][16];
    // bias (+ time-embedding row bias) of a chunk is fetched ONE CHUNK AHEAD, together with the chunk's TMEM load: the
    // adds used to sit right behind their own loads (ncu r01b: 22 % of the GEGLU kernel's stall samples were FADDs
    // waiting on the scoreboard of those LDGs)
    float bv[2][16];
    float bg[GEGLU ? 2 : 1][16];
    // (GEGLU tiles only: on the plain epilogues, which are bound by their stores, the extra live registers cost 4-7 %)
    // (a GEGLU projection without bias adds zeros: a runtime predicate here costs ~10 register moves per column pair)
    constexpr bool has_b = GEGLU;
    auto fetch_bias = [&](int ch, int slot) {
      const int c = ch * 16;
      if (p.bias) {
        const float4* b4 = reinterpret_cast<const float4*>(p.bias + wcol0 + c);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float4 b = __ldg(b4 + j);
          bv[slot][4 * j] = b.x; bv[slot][4 * j + 1] = b.y; bv[slot][4 * j + 2] = b.z; bv[slot][4 * j + 3] = b.w;
          if constexpr (LNIN) {
            const float4 cs = __ldg(reinterpret_cast<const float4*>(p.ln_csum + wcol0 + c) + j);
            bv[slot][4 * j] = fmaf(ln_nrm, cs.x, b.x); bv[slot][4 * j + 1] = fmaf(ln_nrm, cs.y, b.y);
            bv[slot][4 * j + 2] = fmaf(ln_nrm, cs.z, b.z); bv[slot][4 * j + 3] = fmaf(ln_nrm, cs.w, b.w);
          }
        }
        if constexpr (GEGLU) {
          const float4* g4 = reinterpret_cast<const float4*>(p.bias + wcol0 + BN / 2 + c);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 b = __ldg(g4 + j);
            bg[slot][4 * j] = b.x; bg[slot][4 * j + 1] = b.y; bg[slot][4 * j + 2] = b.z; bg[slot][4 * j + 3] = b.w;
            if constexpr (LNIN) {
              const float4 cs = __ldg(reinterpret_cast<const float4*>(p.ln_csum + wcol0 + BN / 2 + c) + j);
              bg[slot][4 * j] = fmaf(ln_nrm, cs.x, b.x); bg[slot][4 * j + 1] = fmaf(ln_nrm, cs.y, b.y);
              bg[slot][4 * j + 2] = fmaf(ln_nrm, cs.z, b.z); bg[slot][4 * j + 3] = fmaf(ln_nrm, cs.w, b.w);
            }
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) { bv[slot][j] = 0.f; if constexpr (GEGLU) bg[slot][j] = 0.f; }
      }
    };
    if (has_b) fetch_bias(0, 0);
#pragma unroll
    for (int ch = 0; ch < NCH; ++ch) {
      const int cur = ch & 1, nxt = cur ^ 1;
      const int c = ch * 16;
      const int k = ch % CH_PER_PANEL, panel = ch / CH_PER_PANEL;
      const int buf = q % NBUF;
      const uint32_t srow = stage_u32 + buf * Cfg::PANEL_BYTES + lane_off;  // shared-space address: STS / LDS, not generic
      if (k == 0) {
        // ---- panel start: the buffer(s) about to be (over)written must have been read out by their TMA store ----
        if (lane == 0) {
          bulk_wait_read0();
          if (RES) {
            if (NBUF == 2) {
              if (tile + ustride < num_tiles) load_residual(tile + ustride, 0, buf ^ 1);
            } else {
              load_residual(tile, panel, 0);
            }
          }
        }
        __syncwarp();
        if (ch == 0) {
          mbar_wait(tfull0 + 8 * acc, acc_phase);
          tc_fence_after();
          tmem_ld16(tacc + c_base, v[0]);
          if constexpr (GEGLU) tmem_ld16(tacc + BN / 2 + c_base, g[0]);
        }
        if (RES) mbar_wait(rbar0 + 8 * buf, (q / NBUF) & 1);
      }
      tmem_wait_ld();
      if (ch + 1 < NCH) {
        tmem_ld16(tacc + c_base + c + 16, v[nxt]);
        if constexpr (GEGLU) tmem_ld16(tacc + BN / 2 + c_base + c + 16, g[nxt]);
        if (has_b) fetch_bias(ch + 1, nxt);
      }
      float f[16];
      if constexpr (!GEGLU) {
#pragma unroll
        for (int j = 0; j < 16; ++j) f[j] = __uint_as_float(v[cur][j]);
      }
      if constexpr (!GEGLU) {
        if constexpr (LNIN) {  // bias' and csum always exist on this path
          const float4* b4 = reinterpret_cast<const float4*>(p.bias + wcol0 + c);
          const float4* c4 = reinterpret_cast<const float4*>(p.ln_csum + wcol0 + c);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 b = __ldg(b4 + j), cs = __ldg(c4 + j);
            f[4 * j] = fmaf(f[4 * j], ln_rstd, fmaf(ln_nrm, cs.x, b.x));
            f[4 * j + 1] = fmaf(f[4 * j + 1], ln_rstd, fmaf(ln_nrm, cs.y, b.y));
            f[4 * j + 2] = fmaf(f[4 * j + 2], ln_rstd, fmaf(ln_nrm, cs.z, b.z));
            f[4 * j + 3] = fmaf(f[4 * j + 3], ln_rstd, fmaf(ln_nrm, cs.w, b.w));
          }
        } else if (p.bias) {
          const float4* b4 = reinterpret_cast<const float4*>(p.bias + wcol0 + c);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 b = __ldg(b4 + j);
            f[4 * j] += b.x; f[4 * j + 1] += b.y; f[4 * j + 2] += b.z; f[4 * j + 3] += b.w;
          }
        }
        if (rb) {
          const float4* r4 = reinterpret_cast<const float4*>(rb + c);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 b = __ldg(r4 + j);
            f[4 * j] += b.x; f[4 * j + 1] += b.y; f[4 * j + 2] += b.z; f[4 * j + 3] += b.w;
          }
        }
      }
      if constexpr (GEGLU) {
        // value * gelu(gate) on column pairs, packed fp32 throughout
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          uint64_t val2 = pack2u(v[cur][2 * j], v[cur][2 * j + 1]);
          uint64_t gate2 = pack2u(g[cur][2 * j], g[cur][2 * j + 1]);
          if constexpr (LNIN) {
            const uint64_t r2 = pack2(ln_rstd, ln_rstd);
            val2 = fma2(val2, r2, pack2(bv[cur][2 * j], bv[cur][2 * j + 1]));
            gate2 = fma2(gate2, r2, pack2(bg[cur][2 * j], bg[cur][2 * j + 1]));
          } else {
            val2 = add2(val2, pack2(bv[cur][2 * j], bv[cur][2 * j + 1]));
            gate2 = add2(gate2, pack2(bg[cur][2 * j], bg[cur][2 * j + 1]));
          }
          unpack2(mul2(val2, gelu_fast2(gate2)), f[2 * j], f[2 * j + 1]);
        }
      } else {
        if (p.gelu) {
#pragma unroll
          for (int j = 0; j < 16; ++j) f[j] = gelu_fast(f[j]);
        }
      }
      // ---- the 16 columns are two 16-byte pieces of this lane's panel row (128B-swizzled when the panel is 64 wide) ----
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int piece = 2 * k + j;
        const int pos = Cfg::PANEL_SWIZZLE ? (piece ^ (lane & 7)) : piece;
        const uint32_t dst = srow + (pos << 4);
        if (RES) {
          const uint4 r = lds128(dst);
          const float2 a = unpack_act2<F16>(r.x), b = unpack_act2<F16>(r.y), cc = unpack_act2<F16>(r.z), d = unpack_act2<F16>(r.w);
          f[8 * j] += a.x; f[8 * j + 1] += a.y; f[8 * j + 2] += b.x; f[8 * j + 3] += b.y;
          f[8 * j + 4] += cc.x; f[8 * j + 5] += cc.y; f[8 * j + 6] += d.x; f[8 * j + 7] += d.y;
        }
        uint4 o;
        o.x = pack_act2<F16>(f[8 * j], f[8 * j + 1]);
        o.y = pack_act2<F16>(f[8 * j + 2], f[8 * j + 3]);
        o.z = pack_act2<F16>(f[8 * j + 4], f[8 * j + 5]);
        o.w = pack_act2<F16>(f[8 * j + 6], f[8 * j + 7]);
        sts128(dst, o);
      }
      if constexpr (RSTAT) {
#pragma unroll
        for (int j = 0; j < 16; ++j) { rsum += f[j]; rsq = fmaf(f[j], f[j], rsq); }
      }
      if constexpr (GNG > 0) {
        constexpr int CPG = (OUT_COLS / 2) / GNG;  // channels per group
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int gi = (c + j) / CPG;  // compile-time: the chunk loop is fully unrolled
          gsum[gi] += f[j];
          gsq[gi] = fmaf(f[j], f[j], gsq[gi]);
        }
      }
      if (k == CH_PER_PANEL - 1) {
        // ---- panel complete: ONE TMA store of 32 rows x PANEL_COLS (rows outside the tensor / slab are clipped) ----
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          tma_store_3d(tmO, smem_u32(stage0 + buf * Cfg::PANEL_BYTES), ocol0 + panel * Cfg::PANEL_COLS, row0, slab);
          bulk_commit();
        }
        ++q;
      }
    }
    tc_fence_before();
    __syncwarp();
    if (lane == 0) {  // the accumulator is free again: tell the MMA warp (of the leader CTA for a pair)
      if (CG == 2) mbar_arrive_cluster(tempty_leader + 8 * acc);
      else mbar_arrive(tempty0 + 8 * acc);
    }
    if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    if constexpr (RSTAT) {
      // this row's {sum, sumsq} over this warp's columns: slot (n_tile, column half) of the row's partial list
      if (lrow < p.M)
        reinterpret_cast<float2*>(p.row_stats)[lrow * (2 * p.num_n_tiles) + 2 * n_tile + group] = make_float2(rsum, rsq);
    }
    if constexpr (GNG > 0) {
      // fold the 32 rows: reduce-scatter over the lane bits 4, 3, ... (a fixed tree), then all-reduce over the rest.
      // Afterwards lane l holds value index l >> SH (even index = sum, odd = sum of squares of group index / 2).
      constexpr int NV = 2 * GNG;
      float rv[NV];
#pragma unroll
      for (int i = 0; i < GNG; ++i) { rv[2 * i] = gsum[i]; rv[2 * i + 1] = gsq[i]; }
      int n = NV;
#pragma unroll
      for (int off = 16; off >= 1; off >>= 1) {
        if (n > 1) {
          n >>= 1;
          const bool up = (lane & off) != 0;
#pragma unroll
          for (int i = 0; i < NV / 2; ++i) {
            if (i < n) {
              const float send = up ? rv[i] : rv[i + n];
              const float keep = up ? rv[i + n] : rv[i];
              rv[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
            }
          }
        } else {
          rv[0] += __shfl_xor_sync(0xffffffffu, rv[0], off);
        }
      }
      constexpr int SH = NV == 32 ? 0 : (NV == 16 ? 1 : (NV == 8 ? 2 : 3));
      if ((lane & ((1 << SH) - 1)) == 0) {
        constexpr int CPG = (OUT_COLS / 2) / GNG;
        const int idx = lane >> SH;
        const int64_t grow = (MODE == 0 ? (int64_t)slab * p.rows_per_batch : 0) + row0;  // first row of this warp
        const int g0 = wcol0 / CPG;
        p.gn_stats[(((grow >> 5) * p.gn_groups) + g0 + (idx >> 1)) * 2 + (idx & 1)] = rv[0];
      }
    }
  }
  if (lane == 0) bulk_wait0();  // this warp's TMA stores have drained before its shared memory goes away
}

// conv_out epilogue: only the first nc_cout (<= 16) accumulator columns are real; thread = pixel writes one element per
// output channel plane (consecutive lanes -> consecutive pixels: coalesced), fp32 bias add, one rounding to the io dtype.
template <int ACC_STRIDE>
__device__ __forceinline__ void epilogue_loop_ncfhw(const GemmParams& p, uint32_t tmem_base, uint32_t tfull0,
                                                    uint32_t tempty0, int warp, int lane, int num_tiles) {
  const int quarter = warp & 3;
  const int group = (warp - 2) >> 2;
  int acc = 0;
  uint32_t acc_phase = 0;
  for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
    const int m_tile = tile / p.num_n_tiles;
    const int64_t grow = (int64_t)m_tile * BM + quarter * 32 + lane;
    mbar_wait(tfull0 + 8 * acc, acc_phase);
    tc_fence_after();
    if (group == 0) {
      uint32_t v[16];
      tmem_ld16(tmem_base + acc * ACC_STRIDE + ((uint32_t)(quarter * 32) << 16), v);
      tmem_wait_ld();
      if (grow < p.M) {
        const int64_t n = grow / p.nc_HW, pix = grow % p.nc_HW;
        const int64_t b = n / p.nc_F, f = n % p.nc_F;
#pragma unroll
        for (int co = 0; co < 16; ++co) {
          if (co < p.nc_cout) {
            const float val = __uint_as_float(v[co]) + (p.bias ? __ldg(p.bias + co) : 0.f);
            const int64_t idx = ((b * p.nc_cout + co) * p.nc_F + f) * p.nc_HW + pix;
            if (p.nc_dtype == B2V_F32) reinterpret_cast<float*>(p.out_raw)[idx] = val;
            else if (p.nc_dtype == B2V_F16) reinterpret_cast<__half*>(p.out_raw)[idx] = __float2half_rn(val);
            else reinterpret_cast<__nv_bfloat16*>(p.out_raw)[idx] = __float2bfloat16_rn(val);
          }
        }
      }
    }
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(tempty0 + 8 * acc);
    if (++acc == 2) { acc = 0; acc_phase ^= 1; }
  }
}

// MODE 0: A through a 3-D tiled tensor map (channels, rows, slabs); MODE 1: A through a 4-D im2col map.
template <int BN, int MODE, int EPI, bool RES, int CG, bool F16, int FUSE = 0>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmR, const GemmParams p) {
  using Cfg = TileCfg<BN, EPI == EPI_GEGLU_T, RES, CG>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * A_BYTES;
  uint8_t* sEpi = smem + STAGES * Cfg::STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sEpi + Cfg::EPI_BYTES);
  // bars: [0,STAGES) full, [STAGES,2*STAGES) empty, tmem_full[2], tmem_empty[2], the TMEM base slot, residual[8 warps][2]
  const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + STAGES);
  const uint32_t tfull0 = smem_u32(bars + 2 * STAGES), tempty0 = smem_u32(bars + 2 * STAGES + 2);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
  const uint32_t rbar0 = smem_u32(bars + 2 * STAGES + 5);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kb_per_tap = p.Kc / BK;
  const int num_kb = p.taps * kb_per_tap;
  const int num_tiles = p.num_m_units * p.num_n_tiles;  // work units (a unit is CG x 128 rows x BN columns)
  // CTA pair: rank 0 (leader) issues the MMAs; both CTAs load operands, both run the epilogue on their own 128 rows
  const int rank = CG == 2 ? (int)(blockIdx.x & 1) : 0;
  const int unit0 = CG == 2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int ustride = CG == 2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    prefetch_tmap(&tmO);
    if (RES) prefetch_tmap(&tmR);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(full0 + 8 * i, 1);
      mbar_init(empty0 + 8 * i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(tfull0 + 8 * i, 1);
      mbar_init(tempty0 + 8 * i, EPI_WARPS * CG);
    }
    for (int i = 0; i < 2 * EPI_WARPS; ++i) mbar_init(rbar0 + 8 * i, 1);
    fence_barrier_init();
  }
  if (CG == 2) cluster_sync_all();  // the peer's barriers exist before anything can signal them
  if (warp == 1) {
    if (CG == 2) { tmem2_alloc(smem_u32(tmem_slot), TMEM_COLS); tmem2_relinquish(); }
    else { tmem_alloc(smem_u32(tmem_slot), TMEM_COLS); tmem_relinquish(); }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // everything above overlapped the previous kernel's tail; from here on its results are read / overwritten

  if (warp == 0) {
    // ===================== TMA producer (both CTAs of a pair) =====================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = unit0; tile < num_tiles; tile += ustride) {
        const int n_tile = tile % p.num_n_tiles, m_unit = tile / p.num_n_tiles;
        int c1 = 0, c2 = 0, c3 = 0;
        if (MODE == 0) {
          c2 = m_unit / p.m_units_per_batch;                                  // slab
          c1 = ((m_unit % p.m_units_per_batch) * CG + rank) * BM;             // first row inside the slab
        } else {
          const int64_t m0 = ((int64_t)m_unit * CG + rank) * BM;
          const int pq = p.P * p.Q;
          c3 = (int)(m0 / pq);
          const int rem = (int)(m0 % pq);
          c2 = (rem / p.Q) * p.conv_stride - 1;                // h of the window's top-left (padding 1)
          c1 = (rem % p.Q) * p.conv_stride - 1;                // w
        }
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(empty0 + 8 * stage, phase ^ 1);
          const uint32_t fb = full0 + 8 * stage;
          // a pair counts the bytes of BOTH CTAs on the leader's barrier; only the leader arrives on it
          if (rank == 0) mbar_expect_tx(fb, CG * Cfg::STAGE_BYTES);
          const int tap = kb / kb_per_tap, kc = (kb % kb_per_tap) * BK;
          const uint32_t dstA = smem_u32(sA + stage * A_BYTES);
          const uint32_t dstB = smem_u32(sB + stage * Cfg::B_BYTES);
          const int brow = n_tile * BN + rank * (BN / CG);
          if (MODE == 0) {
            const int64_t r = (int64_t)c1 + (int64_t)(tap - (p.taps >> 1)) * p.tap_row_stride;
            if (CG == 2) tma2_load_3d(dstA, &tmA, fb, kc, (int)r, c2);
            else tma_load_3d(dstA, &tmA, fb, kc, (int)r, c2);
          } else {
            if (CG == 2) tma2_load_im2col_4d(dstA, &tmA, fb, kc, c1, c2, c3, (uint16_t)(tap % 3), (uint16_t)(tap / 3));
            else tma_load_im2col_4d(dstA, &tmA, fb, kc, c1, c2, c3, (uint16_t)(tap % 3), (uint16_t)(tap / 3));
          }
          if (CG == 2) tma2_load_2d(dstB, &tmB, fb, kb * BK, brow);
          else tma_load_2d(dstB, &tmB, fb, kb * BK, brow);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (the leader CTA of a pair) =====================
    if (rank == 0) {
      constexpr uint32_t idesc = umma_idesc_16(BM * CG, BN, F16, F16);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = unit0; tile < num_tiles; tile += ustride) {
        mbar_wait(tempty0 + 8 * acc, acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * Cfg::ACC_STRIDE;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(full0 + 8 * stage, phase);
          tc_fence_after();
          if (elect_one()) {
            const uint64_t adesc = umma_desc_sw128(smem_u32(sA + stage * A_BYTES));
            const uint64_t bdesc = umma_desc_sw128(smem_u32(sB + stage * Cfg::B_BYTES));
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              // advance 16 bf16 = 32 bytes inside the swizzle atom: +2 in the 16-byte address field
              if (CG == 2) umma2_bf16_ss(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb | k) != 0);
              else umma_bf16_ss(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb | k) != 0);
            }
            if (CG == 2) {
              umma2_commit(empty0 + 8 * stage);                       // frees the slot in BOTH CTAs
              if (kb == num_kb - 1) umma2_commit(tfull0 + 8 * acc);   // both CTAs' epilogues may read their half
            } else {
              umma_commit(empty0 + 8 * stage);                    // slot reusable when these MMAs retire
              if (kb == num_kb - 1) umma_commit(tfull0 + 8 * acc);  // accumulator complete
            }
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue (warps 2..9 -> TMEM lane quarters 2,3,0,1,2,3,0,1; column halves 0,0,0,0,1,1,1,1) ====
    if constexpr (EPI == EPI_NCFHW)
      epilogue_loop_ncfhw<Cfg::ACC_STRIDE>(p, tmem_base, tfull0, tempty0, warp, lane, num_tiles);
    else
      epilogue_loop<BN, MODE, EPI == EPI_GEGLU_T, RES, CG, F16, FUSE>(p, &tmO, &tmR,
                                                                sEpi + (warp - 2) * Cfg::NBUF * Cfg::PANEL_BYTES,
                                                                rbar0 + 16 * (warp - 2), tmem_base, tfull0, tempty0,
                                                                warp, lane, num_tiles);
  }

  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();  // the peer's shared memory / barriers stay valid until both CTAs are done
  if (warp == 1) {
    tc_fence_after();
    if (CG == 2) tmem2_dealloc(tmem_base, TMEM_COLS);
    else tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

template <int BN, int MODE, int EPI, bool RES, int CG, bool F16, int FUSE = 0>
int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, int64_t out_rows_per_slab,
                int64_t out_slabs, cudaStream_t s) {
  using Cfg = TileCfg<BN, EPI == EPI_GEGLU_T, RES, CG>;
  CUtensorMap tmO = tmA, tmR = tmA;  // EPI_NCFHW stores with plain instructions
  if (EPI != EPI_NCFHW) {
    const int n_out = EPI == EPI_GEGLU_T ? p.N / 2 : p.N;
    const uint64_t dims[3] = {(uint64_t)n_out, (uint64_t)out_rows_per_slab, (uint64_t)out_slabs};
    const uint64_t strides[2] = {(uint64_t)p.ldo * 2, (uint64_t)out_rows_per_slab * (uint64_t)p.ldo * 2};
    const uint32_t box[3] = {(uint32_t)Cfg::PANEL_COLS, 32, 1};
    B2V_TRY(tmap_tiled_bf16_sw(&tmO, p.out, 3, dims, strides, box, Cfg::PANEL_SWIZZLE));
    if (RES) B2V_TRY(tmap_tiled_bf16_sw(&tmR, p.residual, 3, dims, strides, box, Cfg::PANEL_SWIZZLE));
  }
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [] {
    attr_err = cudaFuncSetAttribute(gemm_kernel<BN, MODE, EPI, RES, CG, F16, FUSE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    Cfg::SMEM_BYTES);
  });
  if (attr_err != cudaSuccess) return fail(B2V_ERR_CUDA, "gemm smem attr: %s", cudaGetErrorString(attr_err));
  const int units = p.num_m_units * p.num_n_tiles;
  if (CG == 1) {
    const int grid = units < sm_count() ? units : sm_count();
    B2V_CUDA_OK(launch_pdl(gemm_kernel<BN, MODE, EPI, RES, CG, F16, FUSE>, dim3((unsigned)grid), dim3(NUM_THREADS), Cfg::SMEM_BYTES, s,
                           tmA, tmB, tmO, tmR, p));
  } else {
    // one CTA pair (cluster of 2 = the two SMs of a TPC) per work unit, persistent over the units
    const int pairs = units < sm_count() / 2 ? units : sm_count() / 2;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(2 * pairs));
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
    cfg.stream = s;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 2;
    B2V_CUDA_OK(cudaLaunchKernelEx(&cfg, gemm_kernel<BN, MODE, EPI, RES, CG, F16, FUSE>, tmA, tmB, tmO, tmR, p));
  }
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

template <int BN, int MODE, int CG, bool F16, int FUSE = 0>
int launch_plain(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, int64_t orows, int64_t oslabs,
                 cudaStream_t s) {
  if (p.residual) return launch_gemm<BN, MODE, EPI_PLAIN, true, CG, F16, FUSE>(tmA, tmB, p, orows, oslabs, s);
  return launch_gemm<BN, MODE, EPI_PLAIN, false, CG, F16, FUSE>(tmA, tmB, p, orows, oslabs, s);
}

template <int MODE, int CG, bool F16>
int dispatch_gemm_t(int bn, bool geglu, const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p,
                    int64_t orows, int64_t oslabs, cudaStream_t s) {
  if (geglu) {
    if (bn == 256) {
      if constexpr (MODE == 0) {
        if (p.ln_stats) return launch_gemm<256, MODE, EPI_GEGLU_T, false, CG, F16, FUSE_LNIN>(tmA, tmB, p, orows, oslabs, s);
      }
      return launch_gemm<256, MODE, EPI_GEGLU_T, false, CG, F16>(tmA, tmB, p, orows, oslabs, s);
    }
    return fail(B2V_ERR_ARG, "GEGLU epilogue needs N %% 256 == 0");
  }
  if constexpr (MODE == 0) {  // LayerNorm fused across GEMMs: linear layers only
    if (p.ln_stats) {
      if (p.residual) return fail(B2V_ERR_ARG, "a GEMM that normalises its input takes no residual");
      switch (bn) {
        case 256: return launch_gemm<256, MODE, EPI_PLAIN, false, CG, F16, FUSE_LNIN>(tmA, tmB, p, orows, oslabs, s);
        case 160: return launch_gemm<160, MODE, EPI_PLAIN, false, CG, F16, FUSE_LNIN>(tmA, tmB, p, orows, oslabs, s);
        case 128: return launch_gemm<128, MODE, EPI_PLAIN, false, CG, F16, FUSE_LNIN>(tmA, tmB, p, orows, oslabs, s);
      }
      return fail(B2V_ERR_ARG, "no LayerNorm-input tile for N");
    }
    if (p.row_stats) {
      switch (bn) {
        case 256: return launch_plain<256, MODE, CG, F16, FUSE_RSTAT>(tmA, tmB, p, orows, oslabs, s);
        case 160: return launch_plain<160, MODE, CG, F16, FUSE_RSTAT>(tmA, tmB, p, orows, oslabs, s);
        case 128: return launch_plain<128, MODE, CG, F16, FUSE_RSTAT>(tmA, tmB, p, orows, oslabs, s);
      }
      return fail(B2V_ERR_ARG, "no row-statistics tile for N");
    }
  }
  if constexpr (CG == 2) {  // statistics-emitting epilogues exist for CTA pairs only (gn_stats_ok)
    if (p.gn_stats) {
      const int gng = (bn / 2) / (p.N / p.gn_groups);  // whole groups per epilogue warp (gn_stats_shape_ok)
#define B2V_GNS(BN_, G_) if (bn == BN_ && gng == G_) return launch_plain<BN_, MODE, CG, F16, G_>(tmA, tmB, p, orows, oslabs, s)
      B2V_GNS(160, 8); B2V_GNS(160, 4); B2V_GNS(160, 2);      // UNet: 320 / 640 / 1280 channels
      B2V_GNS(256, 16); B2V_GNS(256, 8); B2V_GNS(128, 16);    // SD VAE decoder: 256 / 512 / 128 channels
#undef B2V_GNS
      return fail(B2V_ERR_ARG, "gn_stats: no epilogue for tile %d with %d groups per warp", bn, gng);
    }
  }
  switch (bn) {
    case 256: return launch_plain<256, MODE, CG, F16>(tmA, tmB, p, orows, oslabs, s);
    case 160: return launch_plain<160, MODE, CG, F16>(tmA, tmB, p, orows, oslabs, s);
    case 128: return launch_plain<128, MODE, CG, F16>(tmA, tmB, p, orows, oslabs, s);
    case 64: return launch_plain<64, MODE, CG, F16>(tmA, tmB, p, orows, oslabs, s);
  }
  return fail(B2V_ERR_ARG, "no tile for N");
}

// adt: activation / weight element type of this call (B2V_BF16 or B2V_F16)
template <int MODE, int CG>
int dispatch_gemm(int adt, int bn, bool geglu, const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p,
                  int64_t orows, int64_t oslabs, cudaStream_t s) {
  if (adt == B2V_F16) return dispatch_gemm_t<MODE, CG, true>(bn, geglu, tmA, tmB, p, orows, oslabs, s);
  return dispatch_gemm_t<MODE, CG, false>(bn, geglu, tmA, tmB, p, orows, oslabs, s);
}

// CTA pairs (256-row units, each CTA stages its own A rows and half of B) win or tie on every shape of the census since
// the remote accumulator hand-back became a plain arrive (profiles/r01g_gemm_tile_ab.md: K = 320 GEGLU -8 %, N = 960 -6 %,
// small-M convs -18 %); the only requirement is two 128-row tiles inside a slab. B2V_GEMM_CG=1|2 forces the choice (A/B).
static int pick_cg(int64_t m_tiles_per_slab) {
  static int forced = -1;
  if (forced < 0) {
    const char* e = getenv("B2V_GEMM_CG");
    forced = e ? atoi(e) : 0;
  }
  if (m_tiles_per_slab < 2) return 1;
  if (forced == 1 || forced == 2) return forced;
  return 2;
}

}  // namespace

// Tile width. Candidates are the instantiated widths that divide N; the cost of a candidate is
// waves x (BN + 64): waves = ceil(tiles / resident tile slots) is the wave quantisation (it decides the small-M levels:
// M = 4608, N = 1280 runs 2 waves at BN = 160 against 2 much longer ones at 256), and BN + 64 is the per-tile time — the
// B columns plus a fixed share for the A tile / epilogue hand-over, which is what makes 160 beat 128 at N = 640 / 1920
// (4 instead of 5 A re-reads; measured -8 % to -23 %, profiles/r01g_gemm_tile_ab.md). m_units = row units (128 x CG rows)
// of the whole problem; 0 = unknown (widest tile that divides N). GEGLU tiles are 256 wide by construction of the weight
// interleave.
int pick_bn(int N, bool geglu, int64_t m_units, int cg) {
  if (geglu) return (N % 256 == 0) ? 256 : 0;
  static int forced = -1;  // B2V_GEMM_BN=64|128|160|256 forces the tile width where it divides N (A/B knob of tools/)
  if (forced < 0) {
    const char* e = getenv("B2V_GEMM_BN");
    forced = e ? atoi(e) : 0;
  }
  if ((forced == 64 || forced == 128 || forced == 160 || forced == 256) && N % forced == 0) return forced;
  const int cand[3] = {256, 160, 128};
  const int64_t slots = sm_count() / (cg < 1 ? 1 : cg);
  int best = 0;
  int64_t best_cost = 0;
  for (int i = 0; i < 3; ++i) {
    const int bn = cand[i];
    if (N % bn != 0) continue;
    if (m_units <= 0) return bn;
    const int64_t tiles = m_units * (N / bn);
    const int64_t waves = (tiles + slots - 1) / slots;
    const int64_t cost = waves * (bn + 64);
    if (best == 0 || cost < best_cost) { best = bn; best_cost = cost; }
  }
  if (best) return best;
  return (N % 64 == 0) ? 64 : 0;
}

// The epilogue emits GroupNorm statistics when an epilogue warp's columns (half a tile) are whole groups and that
// (tile, groups-per-warp) pair is instantiated, the tile width is the cost model's own choice, the rows form whole CTA-pair units
// (no clipped rows in any 32-row block) and there are at least two 128-row tiles per slab (CTA pairs).
static bool gn_stats_shape_ok(int64_t M, int N, int64_t rpb, int groups, int bn, int cg) {
  if (groups <= 0 || N % groups != 0 || bn <= 0 || N % bn != 0) return false;
  const int cpg = N / groups, half = bn / 2;
  if (cpg <= 0 || half % cpg != 0) return false;
  const int gng = half / cpg;  // instantiated: 160 x {8, 4, 2}, 256 x {16, 8}, 128 x {16}
  const bool inst = (bn == 160 && (gng == 8 || gng == 4 || gng == 2)) || (bn == 256 && (gng == 16 || gng == 8)) ||
                    (bn == 128 && gng == 16);
  return inst && cg == 2 && rpb % (BM * 2) == 0 && M % rpb == 0;
}
// number of {sum, sumsq} partials per row a GEMM with GemmArgs::row_stats writes (2 per N tile: one per epilogue warp
// column half); 0 when the shape has no row-statistics epilogue (64-wide tiles)
int row_stats_slots(int64_t M, int N, int64_t rows_per_batch) {
  const int64_t rpb = rows_per_batch > 0 ? rows_per_batch : M;
  if (M <= 0 || M % rpb != 0) return 0;
  const int64_t slabs = M / rpb, m_tiles_per_slab = (rpb + BM - 1) / BM;
  const int cg = pick_cg(m_tiles_per_slab);
  const int bn = pick_bn(N, false, slabs * ((m_tiles_per_slab + cg - 1) / cg), cg);
  return (bn == 256 || bn == 160 || bn == 128) ? 2 * (N / bn) : 0;
}
bool gn_stats_ok(int64_t M, int N, int64_t rows_per_batch, int groups) {
  const int64_t rpb = rows_per_batch > 0 ? rows_per_batch : M;
  if (M <= 0 || M % rpb != 0) return false;
  const int64_t slabs = M / rpb, m_tiles_per_slab = (rpb + BM - 1) / BM;
  const int cg = pick_cg(m_tiles_per_slab);
  const int bn = pick_bn(N, false, slabs * ((m_tiles_per_slab + cg - 1) / cg), cg);
  return gn_stats_shape_ok(M, N, rpb, groups, bn, cg);
}

static bool trace_on() {
  static int on = -1;
  if (on < 0) on = getenv("B2V_TRACE") != nullptr;
  return on == 1;
}

// Generic entry used by the UNet host code and by the exported test ops.
int gemm_tiled(const GemmArgs& g, cudaStream_t s) {
  if (trace_on())
    fprintf(stderr, "B2VTRACE gemm M=%lld N=%d K=%d taps=%d flags=%d res=%d rpb=%lld trs=%lld\n", (long long)g.M, g.N,
            g.Kc, g.taps, g.flags, g.residual != nullptr, (long long)g.rows_per_batch, (long long)g.tap_row_stride);
  const bool geglu = (g.flags & B2V_EPI_GEGLU) != 0;
  B2V_CHECK(g.Kc % BK == 0 && g.Kc > 0, "GEMM Kc=%d must be a multiple of 64", g.Kc);
  B2V_CHECK(act_ok(g.adt), "GEMM: activation dtype must be bf16 or fp16");
  B2V_CHECK(g.taps == 1 || g.taps == 3, "tiled GEMM supports 1 or 3 taps");
  B2V_CHECK(!(geglu && (g.residual || g.rowbias)), "the GEGLU epilogue takes no residual / row bias");
  const int64_t rpb = g.rows_per_batch > 0 ? g.rows_per_batch : g.M;
  B2V_CHECK(g.M > 0 && g.M % rpb == 0, "M must be slabs x rows_per_batch");
  const int64_t lda = g.lda > 0 ? g.lda : g.Kc;
  const int64_t slabs = g.M / rpb;
  const int64_t m_tiles_per_slab = (rpb + BM - 1) / BM;
  const int cg = pick_cg(m_tiles_per_slab);
  const int bn = pick_bn(g.N, geglu, slabs * ((m_tiles_per_slab + cg - 1) / cg), cg);
  B2V_CHECK(bn != 0, "GEMM N=%d has no supported tile (needs a multiple of 64; GEGLU of 256)", g.N);
  CUtensorMap tmA, tmB;
  {
    const uint64_t dims[3] = {(uint64_t)g.Kc, (uint64_t)rpb, (uint64_t)slabs};
    const uint64_t strides[2] = {(uint64_t)lda * 2, (uint64_t)rpb * (uint64_t)lda * 2};
    const uint32_t box[3] = {BK, BM, 1};
    B2V_TRY(tmap_tiled_bf16(&tmA, g.a, 3, dims, strides, box));
  }
  {
    const uint64_t K = (uint64_t)g.taps * g.Kc;
    const uint64_t dims[2] = {K, (uint64_t)g.N};
    const uint64_t strides[1] = {K * 2};
    const uint32_t box[2] = {BK, (uint32_t)(bn / cg)};  // a CTA pair loads half of the B tile each
    B2V_TRY(tmap_tiled_bf16(&tmB, g.w, 2, dims, strides, box));
  }
  GemmParams p{};
  p.M = g.M; p.N = g.N; p.Kc = g.Kc; p.taps = g.taps;
  p.rows_per_batch = rpb;
  p.tap_row_stride = g.tap_row_stride;
  p.m_units_per_batch = (int)((rpb + BM * cg - 1) / (BM * cg));
  p.num_m_units = (int)(slabs * p.m_units_per_batch);
  p.num_n_tiles = g.N / bn;
  p.bias = g.bias; p.rowbias = g.rowbias; p.rows_per_group = g.rows_per_group > 0 ? g.rows_per_group : 1;
  p.residual = g.residual;
  p.out = g.out;
  p.ldo = g.ldo > 0 ? g.ldo : (geglu ? g.N / 2 : g.N);
  p.gelu = (g.flags & B2V_EPI_GELU) ? 1 : 0;
  if (g.gn_stats && !geglu && !p.gelu && gn_stats_shape_ok(g.M, g.N, rpb, g.gn_groups, bn, cg)) {
    p.gn_stats = g.gn_stats;
    p.gn_groups = g.gn_groups;
  }
  if (g.row_stats) {
    B2V_CHECK(!geglu && !p.gelu && g.taps == 1 && !p.gn_stats && bn != 64, "row_stats: plain linear epilogues only");
    // The epilogue writes a row's partials for every lane of a 128-row tile below M. With several slabs whose length is
    // not whole M units, the clipped tail lanes of one slab would alias the first rows of the next and race with them.
    B2V_CHECK(slabs == 1 || rpb % ((int64_t)BM * cg) == 0,
              "row_stats: rows_per_batch=%lld must be a multiple of %d when M holds several slabs", (long long)rpb, BM * cg);
    p.row_stats = g.row_stats;
  }
  if (g.ln_stats) {
    B2V_CHECK(g.ln_csum && g.bias && g.ln_slots > 0 && g.taps == 1 && !g.residual && !g.rowbias && !p.gelu && bn != 64 &&
                  !p.row_stats && !p.gn_stats,
              "ln_stats: needs csum + folded bias, a plain or GEGLU linear epilogue");
    p.ln_stats = g.ln_stats; p.ln_csum = g.ln_csum; p.ln_slots = g.ln_slots; p.ln_eps = g.ln_eps;
  }
  if (cg == 2) return dispatch_gemm<0, 2>(g.adt, bn, geglu, tmA, tmB, p, rpb, slabs, s);
  return dispatch_gemm<0, 1>(g.adt, bn, geglu, tmA, tmB, p, rpb, slabs, s);
}

// 3x3 / pad 1 / stride {1,2} convolution over NHWC through the TMA im2col path.
int conv3x3_im2col(const ConvArgs& c, cudaStream_t s) {
  if (trace_on())
    fprintf(stderr, "B2VTRACE conv N=%d H=%d W=%d Cin=%d Cout=%d stride=%d\n", c.N, c.H, c.W, c.Cin, c.Cout, c.stride);
  B2V_CHECK(c.Cin % BK == 0, "conv Cin=%d must be a multiple of 64", c.Cin);
  B2V_CHECK(act_ok(c.adt), "conv: activation dtype must be bf16 or fp16");
  B2V_CHECK(c.stride == 1 || c.stride == 2, "conv stride must be 1 or 2");
  const int P = (c.H - 1) / c.stride + 1, Q = (c.W - 1) / c.stride + 1;
  const int64_t m_tiles = ((int64_t)c.N * P * Q + BM - 1) / BM;
  const int cg = c.ncfhw_out ? 1 : pick_cg(m_tiles);
  const int bn = pick_bn(c.Cout, false, (m_tiles + cg - 1) / cg, cg);
  B2V_CHECK(bn != 0, "conv Cout=%d must be a multiple of 64", c.Cout);
  CUtensorMap tmA, tmB;
  {
    const uint64_t dims[4] = {(uint64_t)c.Cin, (uint64_t)c.W, (uint64_t)c.H, (uint64_t)c.N};
    const uint64_t strides[3] = {(uint64_t)c.Cin * 2, (uint64_t)c.W * c.Cin * 2, (uint64_t)c.H * c.W * c.Cin * 2};
    const int lower[2] = {-1, -1};
    const int upper[2] = {-1, -1};  // pad - (filter-1)*dilation = 1 - 2
    const uint32_t estr[4] = {1, (uint32_t)c.stride, (uint32_t)c.stride, 1};
    B2V_TRY(tmap_im2col_bf16(&tmA, c.x, 4, dims, strides, lower, upper, BK, BM, estr));
  }
  {
    const uint64_t K = 9ull * c.Cin;
    const uint64_t dims[2] = {K, (uint64_t)c.Cout};
    const uint64_t strides[1] = {K * 2};
    const uint32_t box[2] = {BK, (uint32_t)(bn / cg)};
    B2V_TRY(tmap_tiled_bf16(&tmB, c.w, 2, dims, strides, box));
  }
  GemmParams p{};
  p.M = (int64_t)c.N * P * Q; p.N = c.Cout; p.Kc = c.Cin; p.taps = 9;
  p.rows_per_batch = p.M; p.tap_row_stride = 0;
  p.num_m_units = (int)((p.M + BM * cg - 1) / (BM * cg));
  p.m_units_per_batch = p.num_m_units;
  p.num_n_tiles = c.Cout / bn;
  p.P = P; p.Q = Q; p.conv_stride = c.stride;
  p.bias = c.bias; p.rowbias = c.rowbias; p.rows_per_group = c.rows_per_group > 0 ? c.rows_per_group : 1;
  p.residual = c.residual;
  p.out = c.out;
  p.ldo = c.Cout;
  p.gelu = 0;
  if (c.gn_stats && !c.ncfhw_out && gn_stats_shape_ok(p.M, c.Cout, p.M, c.gn_groups, bn, cg)) {
    p.gn_stats = c.gn_stats;
    p.gn_groups = c.gn_groups;
  }
  if (c.ncfhw_out) {
    B2V_CHECK(c.Cout == 64 && c.ncfhw_cout >= 1 && c.ncfhw_cout <= 16 && c.stride == 1 && c.N % c.ncfhw_F == 0 &&
                  dtype_size(c.ncfhw_dtype) != 0,
              "conv NCFHW epilogue: needs a 64-row padded weight, <= 16 real channels, stride 1");
    p.out_raw = c.ncfhw_out; p.nc_dtype = c.ncfhw_dtype; p.nc_cout = c.ncfhw_cout; p.nc_F = c.ncfhw_F;
    p.nc_HW = (int64_t)P * Q;
    if (c.adt == B2V_F16) return launch_gemm<64, 1, EPI_NCFHW, false, 1, true>(tmA, tmB, p, p.M, 1, s);
    return launch_gemm<64, 1, EPI_NCFHW, false, 1, false>(tmA, tmB, p, p.M, 1, s);
  }
  if (cg == 2) return dispatch_gemm<1, 2>(c.adt, bn, false, tmA, tmB, p, p.M, 1, s);
  return dispatch_gemm<1, 1>(c.adt, bn, false, tmA, tmB, p, p.M, 1, s);
}

}  // namespace b2v
