// tcgen05 / TMEM / TMA GEMM for every dense contraction of the UNet3D
// ([UPSTREAM diffusers] UNet3DConditionModel as called at Distribution/strategies/chunk_only.py:71):
//   * linear layers (q/k/v/out projections, GEGLU feed-forward, proj_in/out)        taps = 1
//   * TemporalConvLayer Conv3d (3,1,1), zero padded in time                           taps = 3
//   * ResnetBlock2D / Down / Upsample Conv2d 3x3 as an implicit GEMM (TMA im2col)   taps = 9
// out[M,N] = sum_tap A_tap[M,Kc] · W[N, tap*Kc:(tap+1)*Kc]^T  (+bias, +row-group bias, +residual | GEGLU)
//
// Kernel shape (persistent, warp specialised, one CTA per SM, 192 threads):
//   warp 0   : TMA producer — A tile {128 rows x 64 ch} + W tile {BN rows x 64 ch} per k-block into a
//              128B-swizzled shared-memory ring (STAGES deep), completion on `full` mbarriers
//   warp 1   : MMA issuer  — one elected lane issues 4 x tcgen05.mma (128 x BN x 16) per k-block into a
//              double-buffered TMEM accumulator; tcgen05.commit releases ring slots / publishes the tile
//   warps 2-5: epilogue    — tcgen05.ld the accumulator (lane = output row), fused bias / time-embedding
//              row bias / residual / GEGLU, bf16 pack, 64-byte row-contiguous global stores
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
constexpr int NUM_THREADS = 192;
constexpr int TMEM_COLS = 512;

template <int BN>
struct TileCfg {
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (BN >= 256) ? 4 : (BN >= 128 ? 6 : 8);
  static constexpr int ACC_STRIDE = (BN > 128) ? 256 : 128;  // TMEM columns between the two accumulators
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 256 /*barriers*/ + 1024 /*alignment slack*/;
};

struct GemmParams {
  int64_t M;               // output rows in total
  int N;                   // GEMM N (before GEGLU halving)
  int Kc;                  // channels per tap
  int taps;                // 1 | 3 | 9
  int64_t rows_per_batch;  // tiled mode: rows of one z-slab (tiles never straddle slabs)
  int64_t tap_row_stride;  // tiled mode, 3 taps: rows per frame
  int m_tiles_per_batch;
  int num_m_tiles;
  int num_n_tiles;
  // im2col geometry (output P x Q per image, conv stride)
  int P, Q, conv_stride;
  // epilogue
  const float* bias;        // [N] (GEGLU: packed like W)
  const float* rowbias;     // [M / rows_per_group, N] or null
  int64_t rows_per_group;
  const __nv_bfloat16* residual;  // [M, ldo] or null
  __nv_bfloat16* out;
  int64_t ldo;
  int gelu;
};

template <int BN, bool GEGLU>
__device__ __forceinline__ void epilogue_tile(const GemmParams& p, uint32_t tacc, int64_t grow, bool valid,
                                              int n_tile) {
  constexpr int OUT_COLS = GEGLU ? BN / 2 : BN;
  const int64_t ocol0 = (int64_t)n_tile * OUT_COLS;
  const int wcol0 = n_tile * BN;  // column in the (packed) weight / bias space
#pragma unroll 1
  for (int c = 0; c < OUT_COLS; c += 32) {
    uint32_t v[32];
    tmem_ld32(tacc + c, v);
    uint32_t g[GEGLU ? 32 : 1];
    if constexpr (GEGLU) tmem_ld32(tacc + BN / 2 + c, g);
    tmem_wait_ld();
    if (valid) {
      float f[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
      if (p.bias) {
        const float4* b4 = reinterpret_cast<const float4*>(p.bias + wcol0 + c);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 b = __ldg(b4 + j);
          f[4 * j] += b.x; f[4 * j + 1] += b.y; f[4 * j + 2] += b.z; f[4 * j + 3] += b.w;
        }
      }
      if constexpr (GEGLU) {
        float gt[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) gt[j] = __uint_as_float(g[j]);
        if (p.bias) {
          const float4* b4 = reinterpret_cast<const float4*>(p.bias + wcol0 + BN / 2 + c);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 b = __ldg(b4 + j);
            gt[4 * j] += b.x; gt[4 * j + 1] += b.y; gt[4 * j + 2] += b.z; gt[4 * j + 3] += b.w;
          }
        }
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] *= gelu_erf_f(gt[j]);
      } else {
        if (p.rowbias) {
          const float4* r4 =
              reinterpret_cast<const float4*>(p.rowbias + (grow / p.rows_per_group) * (int64_t)p.N + wcol0 + c);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 b = __ldg(r4 + j);
            f[4 * j] += b.x; f[4 * j + 1] += b.y; f[4 * j + 2] += b.z; f[4 * j + 3] += b.w;
          }
        }
        if (p.gelu) {
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = gelu_erf_f(f[j]);
        }
      }
      if (p.residual) {
        const uint4* r4 = reinterpret_cast<const uint4*>(p.residual + grow * p.ldo + ocol0 + c);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint4 r = __ldg(r4 + j);
          const float2 a = unpack_bf16x2(r.x), b = unpack_bf16x2(r.y), cc = unpack_bf16x2(r.z), d = unpack_bf16x2(r.w);
          f[8 * j] += a.x; f[8 * j + 1] += a.y; f[8 * j + 2] += b.x; f[8 * j + 3] += b.y;
          f[8 * j + 4] += cc.x; f[8 * j + 5] += cc.y; f[8 * j + 6] += d.x; f[8 * j + 7] += d.y;
        }
      }
      uint4* o4 = reinterpret_cast<uint4*>(p.out + grow * p.ldo + ocol0 + c);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint4 o;
        o.x = pack_bf16x2(f[8 * j], f[8 * j + 1]);
        o.y = pack_bf16x2(f[8 * j + 2], f[8 * j + 3]);
        o.z = pack_bf16x2(f[8 * j + 4], f[8 * j + 5]);
        o.w = pack_bf16x2(f[8 * j + 6], f[8 * j + 7]);
        o4[j] = o;
      }
    }
  }
}

// MODE 0: A through a 3-D tiled tensor map (channels, rows, slabs); MODE 1: A through a 4-D im2col map.
template <int BN, int MODE, bool GEGLU>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmParams p) {
  using Cfg = TileCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * A_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES);
  // bars: [0,STAGES) full, [STAGES,2*STAGES) empty, then tmem_full[2], tmem_empty[2], then the TMEM base slot
  const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + STAGES);
  const uint32_t tfull0 = smem_u32(bars + 2 * STAGES), tempty0 = smem_u32(bars + 2 * STAGES + 2);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kb_per_tap = p.Kc / BK;
  const int num_kb = p.taps * kb_per_tap;
  const int num_tiles = p.num_m_tiles * p.num_n_tiles;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(full0 + 8 * i, 1);
      mbar_init(empty0 + 8 * i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(tfull0 + 8 * i, 1);
      mbar_init(tempty0 + 8 * i, 4);
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_slot), TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int n_tile = tile % p.num_n_tiles, m_tile = tile / p.num_n_tiles;
        int c1 = 0, c2 = 0, c3 = 0;
        if (MODE == 0) {
          c2 = m_tile / p.m_tiles_per_batch;                   // slab
          c1 = (m_tile % p.m_tiles_per_batch) * BM;            // first row inside the slab
        } else {
          const int64_t m0 = (int64_t)m_tile * BM;
          const int pq = p.P * p.Q;
          c3 = (int)(m0 / pq);
          const int rem = (int)(m0 % pq);
          c2 = (rem / p.Q) * p.conv_stride - 1;                // h of the window's top-left (padding 1)
          c1 = (rem % p.Q) * p.conv_stride - 1;                // w
        }
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(empty0 + 8 * stage, phase ^ 1);
          const uint32_t fb = full0 + 8 * stage;
          mbar_expect_tx(fb, Cfg::STAGE_BYTES);
          const int tap = kb / kb_per_tap, kc = (kb % kb_per_tap) * BK;
          const uint32_t dstA = smem_u32(sA + stage * A_BYTES);
          if (MODE == 0) {
            const int64_t r = (int64_t)c1 + (int64_t)(tap - (p.taps >> 1)) * p.tap_row_stride;
            tma_load_3d(dstA, &tmA, fb, kc, (int)r, c2);
          } else {
            tma_load_im2col_4d(dstA, &tmA, fb, kc, c1, c2, c3, (uint16_t)(tap % 3), (uint16_t)(tap / 3));
          }
          tma_load_2d(smem_u32(sB + stage * Cfg::B_BYTES), &tmB, fb, kb * BK, n_tile * BN);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = umma_idesc_bf16(BM, BN);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
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
            umma_bf16_ss(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb | k) != 0);
          }
          umma_commit(empty0 + 8 * stage);                    // slot reusable when these MMAs retire
          if (kb == num_kb - 1) umma_commit(tfull0 + 8 * acc);  // accumulator complete
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else {
    // ===================== epilogue (warps 2..5 -> TMEM lane quarters 2,3,0,1) =====================
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int n_tile = tile % p.num_n_tiles, m_tile = tile / p.num_n_tiles;
      int64_t grow;
      bool valid;
      if (MODE == 0) {
        const int slab = m_tile / p.m_tiles_per_batch;
        const int64_t r = (int64_t)(m_tile % p.m_tiles_per_batch) * BM + row;
        valid = r < p.rows_per_batch;
        grow = (int64_t)slab * p.rows_per_batch + r;
      } else {
        grow = (int64_t)m_tile * BM + row;
        valid = grow < p.M;
      }
      mbar_wait(tfull0 + 8 * acc, acc_phase);
      tc_fence_after();
      const uint32_t tacc = tmem_base + acc * Cfg::ACC_STRIDE + ((uint32_t)(quarter * 32) << 16);
      epilogue_tile<BN, GEGLU>(p, tacc, grow, valid, n_tile);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty0 + 8 * acc);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

template <int BN, int MODE, bool GEGLU>
int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, cudaStream_t s) {
  using Cfg = TileCfg<BN>;
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [] {
    attr_err = cudaFuncSetAttribute(gemm_kernel<BN, MODE, GEGLU>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    Cfg::SMEM_BYTES);
  });
  if (attr_err != cudaSuccess) return fail(B2V_ERR_CUDA, "gemm smem attr: %s", cudaGetErrorString(attr_err));
  const int tiles = p.num_m_tiles * p.num_n_tiles;
  const int grid = tiles < sm_count() ? tiles : sm_count();
  gemm_kernel<BN, MODE, GEGLU><<<grid, NUM_THREADS, Cfg::SMEM_BYTES, s>>>(tmA, tmB, p);
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

template <int MODE>
int dispatch_gemm(int bn, bool geglu, const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p,
                  cudaStream_t s) {
  if (geglu) {
    if (bn == 256) return launch_gemm<256, MODE, true>(tmA, tmB, p, s);
    return fail(B2V_ERR_ARG, "GEGLU epilogue needs N %% 256 == 0");
  }
  switch (bn) {
    case 256: return launch_gemm<256, MODE, false>(tmA, tmB, p, s);
    case 160: return launch_gemm<160, MODE, false>(tmA, tmB, p, s);
    case 128: return launch_gemm<128, MODE, false>(tmA, tmB, p, s);
    case 64: return launch_gemm<64, MODE, false>(tmA, tmB, p, s);
  }
  return fail(B2V_ERR_ARG, "no tile for N");
}

}  // namespace

int pick_bn(int N, bool geglu) {
  if (geglu) return (N % 256 == 0) ? 256 : 0;
  if (N % 256 == 0) return 256;
  if (N % 160 == 0 && N % 128 != 0) return 160;
  if (N % 128 == 0) return 128;
  if (N % 64 == 0) return 64;
  return 0;
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
  const int bn = pick_bn(g.N, geglu);
  B2V_CHECK(bn != 0, "GEMM N=%d has no supported tile (needs a multiple of 64; GEGLU of 256)", g.N);
  B2V_CHECK(g.Kc % BK == 0 && g.Kc > 0, "GEMM Kc=%d must be a multiple of 64", g.Kc);
  B2V_CHECK(g.taps == 1 || g.taps == 3, "tiled GEMM supports 1 or 3 taps");
  const int64_t rpb = g.rows_per_batch > 0 ? g.rows_per_batch : g.M;
  B2V_CHECK(g.M > 0 && g.M % rpb == 0, "M must be slabs x rows_per_batch");
  const int64_t lda = g.lda > 0 ? g.lda : g.Kc;
  const int64_t slabs = g.M / rpb;
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
    const uint32_t box[2] = {BK, (uint32_t)bn};
    B2V_TRY(tmap_tiled_bf16(&tmB, g.w, 2, dims, strides, box));
  }
  GemmParams p{};
  p.M = g.M; p.N = g.N; p.Kc = g.Kc; p.taps = g.taps;
  p.rows_per_batch = rpb;
  p.tap_row_stride = g.tap_row_stride;
  p.m_tiles_per_batch = (int)((rpb + BM - 1) / BM);
  p.num_m_tiles = (int)(slabs * p.m_tiles_per_batch);
  p.num_n_tiles = g.N / bn;
  p.bias = g.bias; p.rowbias = g.rowbias; p.rows_per_group = g.rows_per_group > 0 ? g.rows_per_group : 1;
  p.residual = (const __nv_bfloat16*)g.residual;
  p.out = (__nv_bfloat16*)g.out;
  p.ldo = g.ldo > 0 ? g.ldo : (geglu ? g.N / 2 : g.N);
  p.gelu = (g.flags & B2V_EPI_GELU) ? 1 : 0;
  return dispatch_gemm<0>(bn, geglu, tmA, tmB, p, s);
}

// 3x3 / pad 1 / stride {1,2} convolution over NHWC through the TMA im2col path.
int conv3x3_im2col(const ConvArgs& c, cudaStream_t s) {
  if (trace_on())
    fprintf(stderr, "B2VTRACE conv N=%d H=%d W=%d Cin=%d Cout=%d stride=%d\n", c.N, c.H, c.W, c.Cin, c.Cout, c.stride);
  const int bn = pick_bn(c.Cout, false);
  B2V_CHECK(bn != 0, "conv Cout=%d must be a multiple of 64", c.Cout);
  B2V_CHECK(c.Cin % BK == 0, "conv Cin=%d must be a multiple of 64", c.Cin);
  B2V_CHECK(c.stride == 1 || c.stride == 2, "conv stride must be 1 or 2");
  const int P = (c.H - 1) / c.stride + 1, Q = (c.W - 1) / c.stride + 1;
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
    const uint32_t box[2] = {BK, (uint32_t)bn};
    B2V_TRY(tmap_tiled_bf16(&tmB, c.w, 2, dims, strides, box));
  }
  GemmParams p{};
  p.M = (int64_t)c.N * P * Q; p.N = c.Cout; p.Kc = c.Cin; p.taps = 9;
  p.rows_per_batch = p.M; p.tap_row_stride = 0;
  p.num_m_tiles = (int)((p.M + BM - 1) / BM);
  p.m_tiles_per_batch = p.num_m_tiles;
  p.num_n_tiles = c.Cout / bn;
  p.P = P; p.Q = Q; p.conv_stride = c.stride;
  p.bias = c.bias; p.rowbias = c.rowbias; p.rows_per_group = c.rows_per_group > 0 ? c.rows_per_group : 1;
  p.residual = (const __nv_bfloat16*)c.residual;
  p.out = (__nv_bfloat16*)c.out;
  p.ldo = c.Cout;
  p.gelu = 0;
  return dispatch_gemm<1>(bn, false, tmA, tmB, p, s);
}

}  // namespace b2v
