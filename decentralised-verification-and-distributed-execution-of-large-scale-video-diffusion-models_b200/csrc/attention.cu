// Attention kernels of the UNet3D transformer blocks ([UPSTREAM diffusers] Attention inside
// Transformer2DModel / TransformerTemporalModel, reached from chunk_only.py:71):
//   * attention():          spatial self-attention (seq = h*w up to 9216) and text cross-attention (kv = 77),
//                           flash-style on tcgen05: S = Q K^T and O_blk = P V on the tensor cores with TMEM
//                           accumulators, online softmax in registers, TMA-fed K/V ring.
//   * temporal_attention(): attention over the frame axis (seq = F <= 32, batch = B*h*w). FLOPs are
//                           negligible (0.2 %), the kernel is HBM-bound: one warp per (pixel, head), K/V in
//                           shared memory, one lane per query frame, no permute copies of the activations.
#include <stdlib.h>

#include <mutex>

#include "common.cuh"
#include "ops.h"
#include "ptx.cuh"

namespace b2v {
namespace {

using namespace ptx;

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// ------------------------------------------------------------------------------------------------
// flash attention, head_dim 64, q tile 128, kv tile 128, 2 CTAs / SM (the co-resident CTA's MMAs run under this
// CTA's softmax and vice versa)
//   warp 0: TMA producer   warp 1: MMA issuer   warps 2-5: softmax (thread = query row)
// smem: Q 16K | K 3x16K | V 3x16K
// TMEM: S cols [0,128) fp32 | O cols [128,192) fp32 (accumulates over kv tiles) | P cols [192,256) packed 16-bit pairs
// S(j+1) is computed while softmax(j) runs: the softmax warps release the S columns right after loading them.
// Softmax is single pass: the S row (128 fp32) is held in registers; the exponent uses a *stale* running maximum
// that is only refreshed (and O rescaled in TMEM) when the true maximum has moved by more than 2^8 — the final
// O / l is exact either way, and O never round-trips through registers in the steady state.
// P never touches shared memory: the softmax warps write the rounded probabilities back to TMEM (tcgen05.st) and the
// P V MMA takes its A operand from there — no st.shared, no MEMBAR + proxy fence, no operand fetch from shared
// memory for P (measured with the FMA-pipe exponentials below: 4.49 -> 4.01 ms on the L0 shape,
// profiles/rd2a_attn_ab_final.log).
// ------------------------------------------------------------------------------------------------
constexpr int AT_THREADS = 192;
constexpr int AT_TILE = 16384;  // 128 rows x 64 bf16
constexpr int AT_STAGES = 3;    // K / V ring depth
constexpr int AT_SMEM = AT_TILE * (1 + 2 * AT_STAGES) + 128;  // 2 CTAs/SM: 2 x (this + 1 KiB system) <= 228 KiB
constexpr int AT_TMEM_COLS = 256;
constexpr float AT_RESCALE_LOG2 = 8.0f;
// exp2 on the FMA pipe for a share of the scores. The MUFU does 4 ex2 / clk / sub-partition (8.0 cycles per warp
// instruction, tools/ubench_pipes.cu) and one per score makes it the busiest pipe at head_dim 64 (70 %); ex2 on packed
// 16-bit pairs is no way out (ex2.approx.ftz.bf16x2 / f16x2 compile to two MUFU.EX2 + PRMT on sm_100a). Round-to-nearest
// split x = n + f by the 1.5 * 2^23 trick, degree-3 minimax 2^f on [-0.5, 0.5] (max relative error 7.5e-5: below half
// an ulp of the bf16 / fp16 P), n added into the exponent field; all of it on packed fp32 pairs (FFMA2).
constexpr float AT_MAGIC = 12582912.0f;
constexpr float AT_C0 = 0.9999280572f, AT_C1 = 0.6932609677f, AT_C2 = 0.2426111251f, AT_C3 = 0.0551716685f;
constexpr int AT_DEFAULT_POLY = 2;

struct AttnParams {
  __nv_bfloat16* out;
  int64_t ldo;
  int Sq, Skv, kv_group, nkv;
  float scale_log2;      // scale * log2(e)
  float inv_scale_log2;  // 1 / scale_log2
};

// POLY = column pairs out of every 8 whose exponentials are computed on the FMA pipe (0 = all on the MUFU)
template <bool F16, int POLY>
__global__ void __launch_bounds__(AT_THREADS, 2)
attn_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
            const __grid_constant__ CUtensorMap tmV, const AttnParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sQ = smem;
  uint8_t* sK = smem + AT_TILE;
  uint8_t* sV = smem + (1 + AT_STAGES) * AT_TILE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (1 + 2 * AT_STAGES) * AT_TILE);
  const uint32_t q_full = smem_u32(bars + 0), s_full = smem_u32(bars + 1), p_full = smem_u32(bars + 2),
                 o_full = smem_u32(bars + 3), s_free = smem_u32(bars + 4), k_full0 = smem_u32(bars + 5),
                 v_full0 = smem_u32(bars + 5 + AT_STAGES), kv_empty0 = smem_u32(bars + 5 + 2 * AT_STAGES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 5 + 3 * AT_STAGES);
  static_assert((5 + 3 * AT_STAGES) * 8 + 4 <= 128, "barrier block");

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * 128, h = blockIdx.y, b = blockIdx.z;
  const int nkv = p.nkv;

  if (warp == 0 && lane == 0) {
    if (smem_u32(smem) & 1023u) {  // the 128B-swizzled tiles need a 1 KiB aligned base (no slack bytes are reserved)
      printf("b2v: attention shared memory base is not 1 KiB aligned\n");
      __trap();
    }
    prefetch_tmap(&tmQ); prefetch_tmap(&tmK); prefetch_tmap(&tmV);
    mbar_init(q_full, 1);
    for (int i = 0; i < AT_STAGES; ++i) { mbar_init(k_full0 + 8 * i, 1); mbar_init(v_full0 + 8 * i, 1); mbar_init(kv_empty0 + 8 * i, 1); }
    mbar_init(s_full, 1);
    mbar_init(p_full, 4);
    mbar_init(o_full, 1);
    mbar_init(s_free, 4);
    fence_barrier_init();
  }
  if (warp == 1) { tmem_alloc(smem_u32(tmem_slot), AT_TMEM_COLS); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base, tO = tmem_base + 128, tP = tmem_base + 192;
  pdl_wait();

  if (warp == 0) {
    if (elect_one()) {
      mbar_expect_tx(q_full, AT_TILE);
      tma_load_3d(smem_u32(sQ), &tmQ, q_full, h * 64, q0, b);
      const int bk = b / p.kv_group;
      int st = 0, ph = 0;
      for (int j = 0; j < nkv; ++j) {
        mbar_wait(kv_empty0 + 8 * st, ph ^ 1);
        mbar_expect_tx(k_full0 + 8 * st, AT_TILE);
        tma_load_3d(smem_u32(sK + st * AT_TILE), &tmK, k_full0 + 8 * st, h * 64, j * 128, bk);
        mbar_expect_tx(v_full0 + 8 * st, AT_TILE);
        tma_load_3d(smem_u32(sV + st * AT_TILE), &tmV, v_full0 + 8 * st, h * 64, j * 128, bk);
        if (++st == AT_STAGES) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc_qk = umma_idesc_16(128, 128, F16, F16);
    constexpr uint32_t idesc_pv = umma_idesc_16(128, 64, F16, F16) | (1u << 16);  // B (=V) is MN-major
    int qst = 0, qph = 0;  // ring position of the next S = Q K^T
    auto issue_qk = [&]() {
      mbar_wait(k_full0 + 8 * qst, qph);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t a = umma_desc_sw128(smem_u32(sQ)), bd = umma_desc_sw128(smem_u32(sK + qst * AT_TILE));
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_bf16_ss(tS, a + 2 * k, bd + 2 * k, idesc_qk, k != 0);
        umma_commit(s_full);
      }
      __syncwarp();
      if (++qst == AT_STAGES) { qst = 0; qph ^= 1; }
    };
    mbar_wait(q_full, 0);
    issue_qk();
    int st = 0, ph = 0;
    for (int j = 0; j < nkv; ++j) {
      // S(j+1) = Q K(j+1)^T is issued as soon as the softmax warps have pulled S(j) into registers, so it runs on the
      // tensor core UNDER softmax(j) and the softmax warps never wait for a score tile in the steady state
      if (j + 1 < nkv) {
        mbar_wait(s_free, j & 1);
        tc_fence_after();
        issue_qk();
      }
      mbar_wait(p_full, j & 1);   // P(j) is in TMEM, O has been rescaled if it had to be
      mbar_wait(v_full0 + 8 * st, ph);
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          // A = P from TMEM: 128 rows (lanes) x 16 kv per step = 8 columns of packed pairs; B = V: 16 kv rows (2 KiB)
          // per k-step, MN-major
          const uint64_t bd = umma_desc_sw128(smem_u32(sV + st * AT_TILE + k * 2048));
          umma_bf16_ts(tO, tP + 8 * k, bd, idesc_pv, (j | k) != 0);
        }
        umma_commit(o_full);
        umma_commit(kv_empty0 + 8 * st);
      }
      __syncwarp();
      if (++st == AT_STAGES) { st = 0; ph ^= 1; }
    }
  } else {
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
    const float sl2 = p.scale_log2;
    // reference the exponents are taken against, in log2 units and ALWAYS an integer (ceil of the scaled row maximum when
    // it was last refreshed): MAGIC - mb is then exact in fp32, which the FMA-pipe exp2 needs, and rescales of O are
    // exact powers of two
    float mb = -INFINITY;
    float l = 0.f;
    for (int j = 0; j < nkv; ++j) {
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      uint32_t sv[128];
      tmem_ld32(tS + lane_off, sv);
      tmem_ld32(tS + lane_off + 32, sv + 32);
      tmem_ld32(tS + lane_off + 64, sv + 64);
      tmem_ld32(tS + lane_off + 96, sv + 96);
      tmem_wait_ld();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(s_free);  // the S tile is in registers: the tensor core may overwrite it
      const int kv_left = p.Skv - j * 128;  // valid kv columns in this block
      if (kv_left < 128) {
#pragma unroll
        for (int i = 0; i < 128; ++i)
          if (i >= kv_left) sv[i] = 0xff800000u;  // -inf: exp2 -> 0 (2^-125 on the FMA pipe), never the maximum
      }
      float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
      for (int i = 0; i < 128; i += 8) {  // FMNMX3: two new values per instruction
        mx0 = max3(mx0, __uint_as_float(sv[i]), __uint_as_float(sv[i + 1]));
        mx1 = max3(mx1, __uint_as_float(sv[i + 2]), __uint_as_float(sv[i + 3]));
        mx2 = max3(mx2, __uint_as_float(sv[i + 4]), __uint_as_float(sv[i + 5]));
        mx3 = max3(mx3, __uint_as_float(sv[i + 6]), __uint_as_float(sv[i + 7]));
      }
      const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
      // refresh the reference only when it is stale by more than 2^AT_RESCALE_LOG2 (always on the first tile)
      const float mxs = mx * sl2;
      const bool stale = mxs - mb > AT_RESCALE_LOG2;
      float alpha = 1.0f;
      if (stale) {
        const float nb = ceilf(mxs);
        alpha = ex2(mb - nb);  // mb = -inf -> 0
        mb = nb;
      }
      // exponent = s * scale*log2e - mb: two columns per FFMA2 (half the issue slots of scalar fp32)
      const uint64_t sl2_2 = pack2(sl2, sl2), nmb2 = pack2(-mb, -mb);
      const uint64_t mg2 = pack2(AT_MAGIC - mb, AT_MAGIC - mb), neg1 = pack2(-1.f, -1.f);
      const uint64_t c0 = pack2(AT_C0, AT_C0), c1 = pack2(AT_C1, AT_C1), c2 = pack2(AT_C2, AT_C2), c3 = pack2(AT_C3, AT_C3);
      const float lo = (mb - 125.0f) * p.inv_scale_log2;  // scores below it would leave the exponent field: clamp
      uint64_t sumA = 0ull, sumB = 0ull;  // (0.f, 0.f)
      uint32_t pk[64];
#pragma unroll
      for (int i = 0; i < 64; ++i) {
        float e0, e1;
        if (POLY > 0 && ((i & 7) * POLY) % 8 < POLY) {
          const uint64_t s2 = pack2(fmaxf(__uint_as_float(sv[2 * i]), lo), fmaxf(__uint_as_float(sv[2 * i + 1]), lo));
          const uint64_t t2 = fma2(s2, sl2_2, mg2);   // MAGIC + n, n = round(x)
          const uint64_t nm = fma2(t2, neg1, mg2);    // -mb - n (exact)
          const uint64_t fr = fma2(s2, sl2_2, nm);    // f = x - n in [-0.5, 0.5]
          const uint64_t pl = fma2(fma2(fma2(c3, fr, c2), fr, c1), fr, c0);
          uint32_t tb0, tb1, pb0, pb1;
          unpack2u(t2, tb0, tb1);
          unpack2u(pl, pb0, pb1);
          e0 = __uint_as_float(pb0 + (tb0 << 23));   // 2^f * 2^n: n sits in the low mantissa bits of MAGIC + n
          e1 = __uint_as_float(pb1 + (tb1 << 23));
        } else {
          const uint64_t x2 = fma2(pack2u(sv[2 * i], sv[2 * i + 1]), sl2_2, nmb2);
          float x0, x1;
          unpack2(x2, x0, x1);
          e0 = ex2(x0);
          e1 = ex2(x1);
        }
        if (i & 1) sumB = add2(sumB, pack2(e0, e1));
        else sumA = add2(sumA, pack2(e0, e1));
        pk[i] = pack_act2<F16>(e0, e1);
      }
      float s0, s1;
      unpack2(add2(sumA, sumB), s0, s1);
      l = l * alpha + (s0 + s1);
      if (j > 0) {
        mbar_wait(o_full, (j - 1) & 1);  // PV(j-1) retired: P may be overwritten, O may be touched
        tc_fence_after();
        if (__any_sync(0xffffffffu, stale)) {
          // rescale this row of the O accumulator in TMEM (rare after the first few kv tiles)
#pragma unroll
          for (int c = 0; c < 64; c += 32) {
            uint32_t ov[32];
            tmem_ld32(tO + lane_off + c, ov);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) ov[i] = __float_as_uint(__uint_as_float(ov[i]) * alpha);
            tmem_st16(tO + lane_off + c, ov);
            tmem_st16(tO + lane_off + c + 16, ov + 16);
          }
        }
      }
      // P(j): 64 columns of packed pairs, row = lane
      tmem_st16(tP + lane_off, pk);
      tmem_st16(tP + lane_off + 16, pk + 16);
      tmem_st16(tP + lane_off + 32, pk + 32);
      tmem_st16(tP + lane_off + 48, pk + 48);
      tmem_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
    }
    mbar_wait(o_full, (nkv - 1) & 1);
    tc_fence_after();
    const float inv = 1.0f / l;
    uint32_t ov[64];
    tmem_ld32(tO + lane_off, ov);
    tmem_ld32(tO + lane_off + 32, ov + 32);
    tmem_wait_ld();
    if (q0 + row < p.Sq) {
      uint4* dst = reinterpret_cast<uint4*>(p.out + ((int64_t)b * p.Sq + q0 + row) * p.ldo + h * 64);
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        uint4 u;
        u.x = pack_act2<F16>(__uint_as_float(ov[8 * q]) * inv, __uint_as_float(ov[8 * q + 1]) * inv);
        u.y = pack_act2<F16>(__uint_as_float(ov[8 * q + 2]) * inv, __uint_as_float(ov[8 * q + 3]) * inv);
        u.z = pack_act2<F16>(__uint_as_float(ov[8 * q + 4]) * inv, __uint_as_float(ov[8 * q + 5]) * inv);
        u.w = pack_act2<F16>(__uint_as_float(ov[8 * q + 6]) * inv, __uint_as_float(ov[8 * q + 7]) * inv);
        dst[q] = u;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, AT_TMEM_COLS); }
}

// ------------------------------------------------------------------------------------------------
// temporal attention: sequence = frames (F <= 32), head_dim 64, batch = B*h*w*heads. HBM-bound (0.2 % of the FLOPs):
// one warp per (batch, pixel, head) item, mma.sync m16n8k16 straight on registers loaded from global memory
// (tcgen05 needs M >= 64 per CTA tile and shared-memory operands: wrong shape for 16x16x64 problems).
// No shared memory, no permute copies: frame f of a pixel is row0 + f*HW of the [B,F,H,W,C] activation.
// The contraction index of an MMA may be permuted freely as long as A and B agree, so every lane loads whole
// 16/32-byte pieces of a row:
//   S = Q K^T : lane (g, tg) holds d in [16 tg, 16 tg + 16) of frames g / g+8 (Q) and of kv frames 8 nt + g (K)
//   O = P V   : P is the S accumulator re-packed (FA2 register trick); lane (g, tg) holds d in [8 g, 8 g + 8) of
//               kv frames 2 tg, 2 tg + 1, 2 tg + 8, 2 tg + 9 (V), byte-permuted into (frame, frame+1) pairs
// ------------------------------------------------------------------------------------------------
template <bool F16>
__device__ __forceinline__ void mma_16816(float* d, const uint32_t* a, uint32_t b0, uint32_t b1) {
  if constexpr (F16)
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  else
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

constexpr int TA_WARPS = 8;

template <int FT, bool F16>  // FT = number of 16-frame tiles: F <= 16 * FT
__global__ void __launch_bounds__(TA_WARPS * 32)
temporal_attn_kernel(const __nv_bfloat16* __restrict__ qkv, int64_t ld, int q_off, int k_off, int v_off,
                     __nv_bfloat16* __restrict__ out, int64_t ldo, int B, int F, int64_t HW, int heads,
                     float scale_log2) {
  pdl_wait();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, tg = lane & 3;
  const int64_t total = (int64_t)B * HW * heads;
  const uint4 zero4 = make_uint4(0, 0, 0, 0);
  for (int64_t item = (int64_t)blockIdx.x * TA_WARPS + warp; item < total; item += (int64_t)gridDim.x * TA_WARPS) {
    const int h = (int)(item % heads);
    const int64_t bp = item / heads;
    const int64_t pix = bp % HW;
    const int b = (int)(bp / HW);
    const __nv_bfloat16* base = qkv + ((int64_t)b * F * HW + pix) * ld + h * 64;  // frame f at base + f*HW*ld
    const int64_t fstride = HW * ld;
    // ---- K: kv frames 8 nt + g, d in [16 tg, 16 tg + 16) ; V: kv frames 16 ks + {2tg, 2tg+1, 2tg+8, 2tg+9}, d in [8g, 8g+8)
    uint32_t kr[2 * FT][8];
    uint4 vr[FT][4];
#pragma unroll
    for (int nt = 0; nt < 2 * FT; ++nt) {
      const int f = nt * 8 + g;
      uint4 u0 = zero4, u1 = zero4;
      if (f < F) {
        const uint4* p = reinterpret_cast<const uint4*>(base + f * fstride + k_off + tg * 16);
        u0 = __ldg(p);
        u1 = __ldg(p + 1);
      }
      kr[nt][0] = u0.x; kr[nt][1] = u0.y; kr[nt][2] = u0.z; kr[nt][3] = u0.w;
      kr[nt][4] = u1.x; kr[nt][5] = u1.y; kr[nt][6] = u1.z; kr[nt][7] = u1.w;
    }
#pragma unroll
    for (int ks = 0; ks < FT; ++ks) {
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int f = ks * 16 + tg * 2 + (r & 1) + (r >> 1) * 8;
        vr[ks][r] = f < F ? __ldg(reinterpret_cast<const uint4*>(base + f * fstride + v_off + g * 8)) : zero4;
      }
    }
#pragma unroll
    for (int mt = 0; mt < FT; ++mt) {
      if (mt * 16 >= F) break;
      // ---- Q rows: frames 16 mt + g and + 8
      uint32_t q0[8], q1[8];
      {
        const int f0 = mt * 16 + g, f1 = f0 + 8;
        uint4 a0 = zero4, a1 = zero4, c0 = zero4, c1 = zero4;
        if (f0 < F) {
          const uint4* p = reinterpret_cast<const uint4*>(base + f0 * fstride + q_off + tg * 16);
          a0 = __ldg(p); a1 = __ldg(p + 1);
        }
        if (f1 < F) {
          const uint4* p = reinterpret_cast<const uint4*>(base + f1 * fstride + q_off + tg * 16);
          c0 = __ldg(p); c1 = __ldg(p + 1);
        }
        q0[0] = a0.x; q0[1] = a0.y; q0[2] = a0.z; q0[3] = a0.w; q0[4] = a1.x; q0[5] = a1.y; q0[6] = a1.z; q0[7] = a1.w;
        q1[0] = c0.x; q1[1] = c0.y; q1[2] = c0.z; q1[3] = c0.w; q1[4] = c1.x; q1[5] = c1.y; q1[6] = c1.z; q1[7] = c1.w;
      }
      // ---- S = Q K^T
      float sacc[2 * FT][4];
#pragma unroll
      for (int nt = 0; nt < 2 * FT; ++nt) {
        sacc[nt][0] = sacc[nt][1] = sacc[nt][2] = sacc[nt][3] = 0.f;
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          const uint32_t a[4] = {q0[2 * s], q1[2 * s], q0[2 * s + 1], q1[2 * s + 1]};
          mma_16816<F16>(sacc[nt], a, kr[nt][2 * s], kr[nt][2 * s + 1]);
        }
      }
      // ---- softmax over kv frames (row g: [0],[1]; row g+8: [2],[3]); columns nt*8 + 2 tg + {0,1}
      float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 2 * FT; ++nt) {
        const int c = nt * 8 + tg * 2;
        if (c >= F) { sacc[nt][0] = -INFINITY; sacc[nt][2] = -INFINITY; }
        if (c + 1 >= F) { sacc[nt][1] = -INFINITY; sacc[nt][3] = -INFINITY; }
        m0 = fmaxf(m0, fmaxf(sacc[nt][0], sacc[nt][1]));
        m1 = fmaxf(m1, fmaxf(sacc[nt][2], sacc[nt][3]));
      }
      m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1)); m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
      m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1)); m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
      float l0 = 0.f, l1 = 0.f;
#pragma unroll
      for (int nt = 0; nt < 2 * FT; ++nt) {
        sacc[nt][0] = ex2((sacc[nt][0] - m0) * scale_log2); sacc[nt][1] = ex2((sacc[nt][1] - m0) * scale_log2);
        sacc[nt][2] = ex2((sacc[nt][2] - m1) * scale_log2); sacc[nt][3] = ex2((sacc[nt][3] - m1) * scale_log2);
        l0 += sacc[nt][0] + sacc[nt][1];
        l1 += sacc[nt][2] + sacc[nt][3];
      }
      l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
      l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
      const float i0 = 1.0f / l0, i1 = 1.0f / l1;
      // ---- O = P V
      float oacc[8][4];
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) oacc[nt][0] = oacc[nt][1] = oacc[nt][2] = oacc[nt][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < FT; ++ks) {
        const uint32_t a[4] = {pack_act2<F16>(sacc[2 * ks][0] * i0, sacc[2 * ks][1] * i0),
                               pack_act2<F16>(sacc[2 * ks][2] * i1, sacc[2 * ks][3] * i1),
                               pack_act2<F16>(sacc[2 * ks + 1][0] * i0, sacc[2 * ks + 1][1] * i0),
                               pack_act2<F16>(sacc[2 * ks + 1][2] * i1, sacc[2 * ks + 1][3] * i1)};
        const uint32_t v0[4] = {vr[ks][0].x, vr[ks][0].y, vr[ks][0].z, vr[ks][0].w};
        const uint32_t v1[4] = {vr[ks][1].x, vr[ks][1].y, vr[ks][1].z, vr[ks][1].w};
        const uint32_t v2[4] = {vr[ks][2].x, vr[ks][2].y, vr[ks][2].z, vr[ks][2].w};
        const uint32_t v3[4] = {vr[ks][3].x, vr[ks][3].y, vr[ks][3].z, vr[ks][3].w};
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
          const uint32_t sel = (nt & 1) ? 0x7632u : 0x5410u;  // (frame, frame+1) pair of element nt
          const uint32_t b0 = __byte_perm(v0[nt >> 1], v1[nt >> 1], sel);
          const uint32_t b1 = __byte_perm(v2[nt >> 1], v3[nt >> 1], sel);
          mma_16816<F16>(oacc[nt], a, b0, b1);
        }
      }
      // ---- store: row g holds d = 16 tg + nt ([0]) and 16 tg + 8 + nt ([1]); row g+8: [2], [3]
      const int f0 = mt * 16 + g, f1 = f0 + 8;
      const int64_t orow = ((int64_t)b * F * HW + pix) * ldo + h * 64 + tg * 16;
      if (f0 < F) {
        uint4* dst = reinterpret_cast<uint4*>(out + orow + f0 * HW * ldo);
        dst[0] = make_uint4(pack_act2<F16>(oacc[0][0], oacc[1][0]), pack_act2<F16>(oacc[2][0], oacc[3][0]),
                            pack_act2<F16>(oacc[4][0], oacc[5][0]), pack_act2<F16>(oacc[6][0], oacc[7][0]));
        dst[1] = make_uint4(pack_act2<F16>(oacc[0][1], oacc[1][1]), pack_act2<F16>(oacc[2][1], oacc[3][1]),
                            pack_act2<F16>(oacc[4][1], oacc[5][1]), pack_act2<F16>(oacc[6][1], oacc[7][1]));
      }
      if (f1 < F) {
        uint4* dst = reinterpret_cast<uint4*>(out + orow + f1 * HW * ldo);
        dst[0] = make_uint4(pack_act2<F16>(oacc[0][2], oacc[1][2]), pack_act2<F16>(oacc[2][2], oacc[3][2]),
                            pack_act2<F16>(oacc[4][2], oacc[5][2]), pack_act2<F16>(oacc[6][2], oacc[7][2]));
        dst[1] = make_uint4(pack_act2<F16>(oacc[0][3], oacc[1][3]), pack_act2<F16>(oacc[2][3], oacc[3][3]),
                            pack_act2<F16>(oacc[4][3], oacc[5][3]), pack_act2<F16>(oacc[6][3], oacc[7][3]));
      }
    }
  }
}

}  // namespace

int attention(const AttnArgs& a, cudaStream_t s) {
  B2V_CHECK(a.q && a.k && a.v && a.out, "attention: null pointer");
  B2V_CHECK(a.nb > 0 && a.heads > 0 && a.Sq > 0 && a.Skv > 0 && a.kv_group > 0 && a.nb % a.kv_group == 0,
            "attention: bad shape");
  B2V_CHECK(a.ldq % 8 == 0 && a.ldkv % 8 == 0 && a.ldo % 8 == 0, "attention: leading dims must be multiples of 8");
  B2V_CHECK(a.heads <= 65535 && a.nb <= 65535, "attention: grid too large");
  B2V_CHECK(act_ok(a.adt), "attention: activation dtype must be bf16 or fp16");
  // B2V_ATTN_POLY = 0..3 picks the share of FMA-pipe exponentials for A/B runs (tools/attn_ab.py); the default is the
  // fastest measured (profiles/r02_attn_ab.md)
  using Kern = void (*)(CUtensorMap, CUtensorMap, CUtensorMap, AttnParams);
  struct Variant { Kern bf16, f16; };
  static const Variant variants[] = {
#define B2V_AV(P) {attn_kernel<false, P>, attn_kernel<true, P>}
      B2V_AV(0), B2V_AV(1), B2V_AV(2), B2V_AV(3),
#undef B2V_AV
  };
  constexpr int n_variants = (int)(sizeof(variants) / sizeof(variants[0]));
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  static int variant = AT_DEFAULT_POLY;
  std::call_once(once, [] {
    if (const char* e = getenv("B2V_ATTN_POLY")) {
      const int v = atoi(e);
      if (v >= 0 && v < n_variants) variant = v;
    }
    for (int i = 0; i < n_variants && attr_err == cudaSuccess; ++i) {
      attr_err = cudaFuncSetAttribute(variants[i].bf16, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM);
      if (attr_err == cudaSuccess)
        attr_err = cudaFuncSetAttribute(variants[i].f16, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM);
    }
  });
  if (attr_err != cudaSuccess) return fail(B2V_ERR_CUDA, "attention smem attr: %s", cudaGetErrorString(attr_err));
  CUtensorMap tmQ, tmK, tmV;
  const uint32_t box[3] = {64, 128, 1};
  {
    const uint64_t dims[3] = {(uint64_t)a.heads * 64, (uint64_t)a.Sq, (uint64_t)a.nb};
    const uint64_t str[2] = {(uint64_t)a.ldq * 2, (uint64_t)a.Sq * a.ldq * 2};
    B2V_TRY(tmap_tiled_bf16(&tmQ, a.q, 3, dims, str, box));
  }
  {
    const uint64_t dims[3] = {(uint64_t)a.heads * 64, (uint64_t)a.Skv, (uint64_t)(a.nb / a.kv_group)};
    const uint64_t str[2] = {(uint64_t)a.ldkv * 2, (uint64_t)a.Skv * a.ldkv * 2};
    B2V_TRY(tmap_tiled_bf16(&tmK, a.k, 3, dims, str, box));
    B2V_TRY(tmap_tiled_bf16(&tmV, a.v, 3, dims, str, box));
  }
  AttnParams p;
  p.out = (__nv_bfloat16*)a.out;
  p.ldo = a.ldo;
  p.Sq = a.Sq; p.Skv = a.Skv; p.kv_group = a.kv_group;
  p.nkv = (a.Skv + 127) / 128;
  p.scale_log2 = a.scale * 1.4426950408889634f;
  p.inv_scale_log2 = 1.0f / p.scale_log2;
  dim3 grid((unsigned)((a.Sq + 127) / 128), (unsigned)a.heads, (unsigned)a.nb);
  const Kern kern = a.adt == B2V_F16 ? variants[variant].f16 : variants[variant].bf16;
  B2V_CUDA_OK(launch_pdl(kern, grid, dim3(AT_THREADS), AT_SMEM, s, tmQ, tmK, tmV, p));
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

int temporal_attention(int adt, const void* qkv, int64_t ld, int q_off, int k_off, int v_off, void* out, int64_t ldo,
                       int B, int F, int64_t HW, int heads, float scale, cudaStream_t s) {
  B2V_CHECK(qkv && out, "temporal_attention: null pointer");
  B2V_CHECK(act_ok(adt), "temporal_attention: activation dtype must be bf16 or fp16");
  B2V_CHECK(F >= 1 && F <= 32, "temporal_attention: F=%d must be in [1,32]", F);
  B2V_CHECK(ld % 8 == 0 && ldo % 8 == 0 && q_off % 8 == 0 && k_off % 8 == 0 && v_off % 8 == 0,
            "temporal_attention: offsets must be multiples of 8");
  const int64_t total = (int64_t)B * HW * heads;
  int64_t grid = (total + TA_WARPS - 1) / TA_WARPS;
  const int64_t cap = (int64_t)sm_count() * 8;
  if (grid > cap) grid = cap;
  B2V_CHECK(ld % 8 == 0 && ((uintptr_t)qkv & 15) == 0 && ((uintptr_t)out & 15) == 0, "temporal_attention: 16-byte alignment");
#define B2V_TA(FT, F16)                                                                                             \
  B2V_CUDA_OK(launch_pdl(temporal_attn_kernel<FT, F16>, dim3((unsigned)grid), dim3(TA_WARPS * 32), 0, s,                \
                         (const __nv_bfloat16*)qkv, ld, q_off, k_off, v_off, (__nv_bfloat16*)out, ldo, B, F, HW, heads, \
                         scale * 1.4426950408889634f))
  if (adt == B2V_F16) { if (F <= 16) B2V_TA(1, true); else B2V_TA(2, true); }
  else { if (F <= 16) B2V_TA(1, false); else B2V_TA(2, false); }
#undef B2V_TA
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

}  // namespace b2v
