// Attention kernels of the UNet3D transformer blocks ([UPSTREAM diffusers] Attention inside
// Transformer2DModel / TransformerTemporalModel, reached from chunk_only.py:71):
//   * attention():          spatial self-attention (seq = h*w up to 9216) and text cross-attention (kv = 77),
//                           flash-style on tcgen05: S = Q K^T and O_blk = P V on the tensor cores with TMEM
//                           accumulators, online softmax in registers, TMA-fed K/V ring.
//   * temporal_attention(): attention over the frame axis (seq = F <= 32, batch = B*h*w). FLOPs are
//                           negligible (0.2 %), the kernel is HBM-bound: one warp per (pixel, head), K/V in
//                           shared memory, one lane per query frame, no permute copies of the activations.
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
// flash attention, head_dim 64, q tile 128, kv tile 128, 2 CTAs / SM
//   warp 0: TMA producer   warp 1: MMA issuer   warps 2-5: softmax (thread = query row)
// smem: Q 16K | K 2x16K | V 2x16K | P 32K   TMEM: S cols [0,128)  O_blk cols [128,192)
// ------------------------------------------------------------------------------------------------
constexpr int AT_THREADS = 192;
constexpr int AT_TILE = 16384;  // 128 rows x 64 bf16
constexpr int AT_SMEM = AT_TILE * (1 + 2 + 2 + 2) + 128 + 1024;
constexpr int AT_TMEM_COLS = 256;

struct AttnParams {
  __nv_bfloat16* out;
  int64_t ldo;
  int Sq, Skv, kv_group, nkv;
  float scale_log2;  // scale * log2(e)
};

__global__ void __launch_bounds__(AT_THREADS, 2)
attn_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
            const __grid_constant__ CUtensorMap tmV, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sQ = smem;
  uint8_t* sK = smem + AT_TILE;
  uint8_t* sV = smem + 3 * AT_TILE;
  uint8_t* sP = smem + 5 * AT_TILE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 7 * AT_TILE);
  const uint32_t q_full = smem_u32(bars + 0), k_full0 = smem_u32(bars + 1), v_full0 = smem_u32(bars + 3),
                 kv_empty0 = smem_u32(bars + 5), s_full = smem_u32(bars + 7), p_full = smem_u32(bars + 8),
                 o_full = smem_u32(bars + 9);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * 128, h = blockIdx.y, b = blockIdx.z;
  const int nkv = p.nkv;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmQ); prefetch_tmap(&tmK); prefetch_tmap(&tmV);
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(k_full0 + 8 * i, 1); mbar_init(v_full0 + 8 * i, 1); mbar_init(kv_empty0 + 8 * i, 1); }
    mbar_init(s_full, 1);
    mbar_init(p_full, 4);
    mbar_init(o_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) { tmem_alloc(smem_u32(tmem_slot), AT_TMEM_COLS); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base, tO = tmem_base + 128;

  if (warp == 0) {
    if (elect_one()) {
      mbar_expect_tx(q_full, AT_TILE);
      tma_load_3d(smem_u32(sQ), &tmQ, q_full, h * 64, q0, b);
      const int bk = b / p.kv_group;
      for (int j = 0; j < nkv; ++j) {
        const int st = j & 1;
        mbar_wait(kv_empty0 + 8 * st, ((j >> 1) & 1) ^ 1);
        mbar_expect_tx(k_full0 + 8 * st, AT_TILE);
        tma_load_3d(smem_u32(sK + st * AT_TILE), &tmK, k_full0 + 8 * st, h * 64, j * 128, bk);
        mbar_expect_tx(v_full0 + 8 * st, AT_TILE);
        tma_load_3d(smem_u32(sV + st * AT_TILE), &tmV, v_full0 + 8 * st, h * 64, j * 128, bk);
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc_qk = umma_idesc_bf16(128, 128);
    constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 64) | (1u << 16);  // B (=V) is MN-major
    auto issue_qk = [&](int j) {
      const int st = j & 1;
      mbar_wait(k_full0 + 8 * st, (j >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t a = umma_desc_sw128(smem_u32(sQ)), bd = umma_desc_sw128(smem_u32(sK + st * AT_TILE));
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_bf16_ss(tS, a + 2 * k, bd + 2 * k, idesc_qk, k != 0);
        umma_commit(s_full);
      }
      __syncwarp();
    };
    mbar_wait(q_full, 0);
    issue_qk(0);
    for (int j = 0; j < nkv; ++j) {
      const int st = j & 1;
      mbar_wait(p_full, j & 1);
      mbar_wait(v_full0 + 8 * st, (j >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          // A = P: two 64-wide K blocks of 16 KiB; B = V: 16 kv rows (2 KiB) per k-step, MN-major
          const uint64_t a = umma_desc_sw128(smem_u32(sP + (k >> 2) * AT_TILE)) + 2 * (k & 3);
          const uint64_t bd = umma_desc_sw128(smem_u32(sV + st * AT_TILE + k * 2048));
          umma_bf16_ss(tO, a, bd, idesc_pv, k != 0);
        }
        umma_commit(o_full);
        umma_commit(kv_empty0 + 8 * st);
      }
      __syncwarp();
      if (j + 1 < nkv) issue_qk(j + 1);
    }
  } else {
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
    float m = -INFINITY, l = 0.f;
    float o[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) o[i] = 0.f;
    const float sl2 = p.scale_log2;
    uint8_t* prow = sP + row * 128;
    const int sw = row & 7;
    for (int j = 0; j < nkv; ++j) {
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      const int kv_left = p.Skv - j * 128;  // valid kv columns in this block
      float mx = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < 128; c += 32) {
        uint32_t v[32];
        tmem_ld32(tS + lane_off + c, v);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (c + i < kv_left) mx = fmaxf(mx, __uint_as_float(v[i]));
      }
      const float m_new = fmaxf(m, mx);
      const float alpha = ex2((m - m_new) * sl2);  // m = -inf -> 0
      const float mb = m_new * sl2;
      float sum = 0.f;
#pragma unroll 1
      for (int c = 0; c < 128; c += 32) {
        uint32_t v[32];
        tmem_ld32(tS + lane_off + c, v);
        tmem_wait_ld();
        float pr[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float e = (c + i < kv_left) ? ex2(__uint_as_float(v[i]) * sl2 - mb) : 0.f;
          pr[i] = e;
          sum += e;
        }
        // 4 x 16-byte chunks of the row (8 bf16 each) into the 128B-swizzled K-major P tile
        uint8_t* blk = prow + (c >> 6) * AT_TILE;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint4 u;
          u.x = pack_bf16x2(pr[8 * q], pr[8 * q + 1]);
          u.y = pack_bf16x2(pr[8 * q + 2], pr[8 * q + 3]);
          u.z = pack_bf16x2(pr[8 * q + 4], pr[8 * q + 5]);
          u.w = pack_bf16x2(pr[8 * q + 6], pr[8 * q + 7]);
          const int chunk = ((c & 63) >> 3) + q;
          *reinterpret_cast<uint4*>(blk + ((chunk ^ sw) << 4)) = u;
        }
      }
      fence_proxy_async_smem();  // generic-proxy smem writes -> visible to the tensor core (async proxy)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
      l = l * alpha + sum;
      m = m_new;
      mbar_wait(o_full, j & 1);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < 64; c += 32) {
        uint32_t v[32];
        tmem_ld32(tO + lane_off + c, v);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[c + i] = o[c + i] * alpha + __uint_as_float(v[i]);
      }
      tc_fence_before();
    }
    if (q0 + row < p.Sq) {
      const float inv = 1.0f / l;
      uint4* dst = reinterpret_cast<uint4*>(p.out + ((int64_t)b * p.Sq + q0 + row) * p.ldo + h * 64);
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        uint4 u;
        u.x = pack_bf16x2(o[8 * q] * inv, o[8 * q + 1] * inv);
        u.y = pack_bf16x2(o[8 * q + 2] * inv, o[8 * q + 3] * inv);
        u.z = pack_bf16x2(o[8 * q + 4] * inv, o[8 * q + 5] * inv);
        u.w = pack_bf16x2(o[8 * q + 6] * inv, o[8 * q + 7] * inv);
        dst[q] = u;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, AT_TMEM_COLS); }
}

// ------------------------------------------------------------------------------------------------
// temporal attention: one warp per (batch, pixel, head); lane f = query frame f
// ------------------------------------------------------------------------------------------------
constexpr int TA_WARPS = 4;

__global__ void __launch_bounds__(TA_WARPS * 32)
temporal_attn_kernel(const __nv_bfloat16* __restrict__ qkv, int64_t ld, int q_off, int k_off, int v_off,
                     __nv_bfloat16* __restrict__ out, int64_t ldo, int B, int F, int64_t HW, int heads,
                     float scale_log2) {
  __shared__ __align__(16) __nv_bfloat16 sK[TA_WARPS][32][64];
  __shared__ __align__(16) __nv_bfloat16 sV[TA_WARPS][32][64];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t total = (int64_t)B * HW * heads;
  for (int64_t item = (int64_t)blockIdx.x * TA_WARPS + warp; item < total; item += (int64_t)gridDim.x * TA_WARPS) {
    const int h = (int)(item % heads);
    const int64_t bp = item / heads;
    const int64_t pix = bp % HW;
    const int b = (int)(bp / HW);
    const int64_t row0 = (int64_t)b * F * HW + pix;  // frame f -> row0 + f*HW
    __syncwarp();
    // cooperative K/V load: 8 lanes x 16 B per frame row, 4 frames per pass
    for (int f = lane >> 3; f < F; f += 4) {
      const __nv_bfloat16* src = qkv + (row0 + (int64_t)f * HW) * ld + h * 64 + (lane & 7) * 8;
      *reinterpret_cast<uint4*>(&sK[warp][f][(lane & 7) * 8]) = __ldg(reinterpret_cast<const uint4*>(src + k_off));
      *reinterpret_cast<uint4*>(&sV[warp][f][(lane & 7) * 8]) = __ldg(reinterpret_cast<const uint4*>(src + v_off));
    }
    __syncwarp();
    if (lane < F) {
      float q[64];
      const uint4* qs = reinterpret_cast<const uint4*>(qkv + (row0 + (int64_t)lane * HW) * ld + q_off + h * 64);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const uint4 u = __ldg(qs + i);
        const float2 a = unpack_bf16x2(u.x), bb = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
        q[8 * i] = a.x; q[8 * i + 1] = a.y; q[8 * i + 2] = bb.x; q[8 * i + 3] = bb.y;
        q[8 * i + 4] = c.x; q[8 * i + 5] = c.y; q[8 * i + 6] = d.x; q[8 * i + 7] = d.y;
      }
      float sc[32];
      float mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        sc[j] = -INFINITY;
        if (j < F) {
          const uint4* kr = reinterpret_cast<const uint4*>(&sK[warp][j][0]);
          float acc = 0.f;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint4 u = kr[i];
            const float2 a = unpack_bf16x2(u.x), bb = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
            acc += q[8 * i] * a.x + q[8 * i + 1] * a.y + q[8 * i + 2] * bb.x + q[8 * i + 3] * bb.y +
                   q[8 * i + 4] * c.x + q[8 * i + 5] * c.y + q[8 * i + 6] * d.x + q[8 * i + 7] * d.y;
          }
          sc[j] = acc;
          mx = fmaxf(mx, acc);
        }
      }
      float sum = 0.f;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const float e = (j < F) ? ex2((sc[j] - mx) * scale_log2) : 0.f;
        sc[j] = e;
        sum += e;
      }
      const float inv = 1.0f / sum;
      float o[64];
#pragma unroll
      for (int i = 0; i < 64; ++i) o[i] = 0.f;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        if (j < F) {
          const float pj = sc[j] * inv;
          const uint4* vr = reinterpret_cast<const uint4*>(&sV[warp][j][0]);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint4 u = vr[i];
            const float2 a = unpack_bf16x2(u.x), bb = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
            o[8 * i] += pj * a.x; o[8 * i + 1] += pj * a.y; o[8 * i + 2] += pj * bb.x; o[8 * i + 3] += pj * bb.y;
            o[8 * i + 4] += pj * c.x; o[8 * i + 5] += pj * c.y; o[8 * i + 6] += pj * d.x; o[8 * i + 7] += pj * d.y;
          }
        }
      }
      uint4* dst = reinterpret_cast<uint4*>(out + (row0 + (int64_t)lane * HW) * ldo + h * 64);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        uint4 u;
        u.x = pack_bf16x2(o[8 * i], o[8 * i + 1]); u.y = pack_bf16x2(o[8 * i + 2], o[8 * i + 3]);
        u.z = pack_bf16x2(o[8 * i + 4], o[8 * i + 5]); u.w = pack_bf16x2(o[8 * i + 6], o[8 * i + 7]);
        dst[i] = u;
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
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [] {
    attr_err = cudaFuncSetAttribute(attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM);
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
  dim3 grid((unsigned)((a.Sq + 127) / 128), (unsigned)a.heads, (unsigned)a.nb);
  attn_kernel<<<grid, AT_THREADS, AT_SMEM, s>>>(tmQ, tmK, tmV, p);
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

int temporal_attention(const void* qkv, int64_t ld, int q_off, int k_off, int v_off, void* out, int64_t ldo, int B,
                       int F, int64_t HW, int heads, float scale, cudaStream_t s) {
  B2V_CHECK(qkv && out, "temporal_attention: null pointer");
  B2V_CHECK(F >= 1 && F <= 32, "temporal_attention: F=%d must be in [1,32]", F);
  B2V_CHECK(ld % 8 == 0 && ldo % 8 == 0 && q_off % 8 == 0 && k_off % 8 == 0 && v_off % 8 == 0,
            "temporal_attention: offsets must be multiples of 8");
  const int64_t total = (int64_t)B * HW * heads;
  int64_t grid = (total + TA_WARPS - 1) / TA_WARPS;
  const int64_t cap = (int64_t)sm_count() * 8;
  if (grid > cap) grid = cap;
  temporal_attn_kernel<<<(unsigned)grid, TA_WARPS * 32, 0, s>>>((const __nv_bfloat16*)qkv, ld, q_off, k_off, v_off,
                                                                (__nv_bfloat16*)out, ldo, B, F, HW, heads,
                                                                scale * 1.4426950408889634f);
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

}  // namespace b2v
