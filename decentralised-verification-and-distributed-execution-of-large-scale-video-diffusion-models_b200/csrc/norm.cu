// HBM-bound normalisation / layout kernels of the UNet3D ([UPSTREAM diffusers] ResnetBlock2D,
// TemporalConvLayer, Transformer2DModel, TransformerTemporalModel, BasicTransformerBlock):
//   * GroupNorm(32) (+SiLU) over channels-last rows, statistics per (item, group) where an item is a
//     frame (spatial norms) or a whole batch element (temporal norms over [B,C,F,H,W])
//   * LayerNorm over channels
//   * nearest 2x upsample, channel concat, explicit im2col (fallback / stride-2 cross-check)
// All 128-bit vectorised, 8 bf16 per thread per access, fp32 math.
#include <stdlib.h>

#include "common.cuh"
#include "ops.h"

namespace b2v {
namespace {

// ---------------- GroupNorm ----------------
// Thread mapping of both passes: a CTA is vecs x rpp threads (vecs = C/8 16-byte vectors per row, rpp rows per pass), so
// EVERY lane owns one fixed 8-channel vector and walks rows; GN_UNROLL rows are loaded before any is consumed (the
// kernels are pure HBM streams: the loads in flight per SM set the bandwidth). grid = (row_splits, n_items).
// pass 1: per-(item, group) sum / sum of squares. C/groups is always even, so a bf16 pair never straddles two groups.
// DETERMINISTIC on purpose (the commitment re-executes steps bit-for-bit): per-thread pair partials go to shared
// memory, one thread per group folds them in a fixed order, each CTA writes its partial to its own slot, and
// the apply pass folds the slots in a fixed order in fp64 (gn_fold) — no atomics anywhere.
constexpr int GN_MAX_THREADS = 384, GN_UNROLL = 4;

__device__ __forceinline__ uint4 gn_load(const __nv_bfloat16* __restrict__ x1, int C1, const __nv_bfloat16* __restrict__ x2,
                                         int C2, int64_t grow, int c) {
  if (c < C1) return __ldg(reinterpret_cast<const uint4*>(x1 + grow * C1 + c));
  return __ldg(reinterpret_cast<const uint4*>(x2 + grow * C2 + (c - C1)));
}

template <bool F16>
__global__ void __launch_bounds__(GN_MAX_THREADS)
gn_stats_kernel(const __nv_bfloat16* __restrict__ x1, int C1, const __nv_bfloat16* __restrict__ x2, int C2,
                int64_t rows_per_item, int groups, int rpp, float* __restrict__ partial /*[item][split][group][2]*/) {
  extern __shared__ float sh[];  // [2][rpp][C/2]
  pdl_wait();
  const int C = C1 + C2;
  const int vecs = C >> 3, pairs = C >> 1;
  const int cpg = C / groups;
  const int item = blockIdx.y;
  const int ty = threadIdx.x / vecs, tx = threadIdx.x - ty * vecs;
  const int c = tx << 3;
  const int64_t rows_per_cta = (rows_per_item + gridDim.x - 1) / gridDim.x;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_cta;
  int64_t r1 = r0 + rows_per_cta;
  if (r1 > rows_per_item) r1 = rows_per_item;
  const int64_t base = (int64_t)item * rows_per_item;
  float s[4] = {0.f, 0.f, 0.f, 0.f}, ss[4] = {0.f, 0.f, 0.f, 0.f};
  for (int64_t r = r0 + ty; r < r1; r += (int64_t)rpp * GN_UNROLL) {
    uint4 u[GN_UNROLL];
#pragma unroll
    for (int k = 0; k < GN_UNROLL; ++k) {
      const int64_t rr = r + (int64_t)k * rpp;
      u[k] = rr < r1 ? gn_load(x1, C1, x2, C2, base + rr, c) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int k = 0; k < GN_UNROLL; ++k) {
      const uint32_t uu[4] = {u[k].x, u[k].y, u[k].z, u[k].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = unpack_act2<F16>(uu[j]);
        s[j] += f.x + f.y;
        ss[j] += f.x * f.x + f.y * f.y;
      }
    }
  }
  float* shs = sh;
  float* shq = sh + rpp * pairs;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    shs[ty * pairs + (c >> 1) + j] = s[j];
    shq[ty * pairs + (c >> 1) + j] = ss[j];
  }
  __syncthreads();
  const int ppg = cpg >> 1;
  for (int g = threadIdx.x; g < groups; g += blockDim.x) {
    float a = 0.f, q = 0.f;
    for (int y = 0; y < rpp; ++y)
      for (int k = 0; k < ppg; ++k) {
        a += shs[y * pairs + g * ppg + k];
        q += shq[y * pairs + g * ppg + k];
      }
    float* dst = partial + ((((int64_t)item * gridDim.x + blockIdx.x) * groups + g) << 1);
    dst[0] = a;
    dst[1] = q;
  }
}

// Fold the per-CTA partials [split][group]{sum, sumsq} of one item into {mean, rstd} per group, inside every CTA of the
// apply pass (a few KB out of L2; the separate finalize launch it replaces cost 5 us x 166 norms per forward).
// lane = group, warp = a stride over the splits: each warp reads whole 256-byte rows of the partial table (coalesced),
// adds its splits in increasing order in fp64 and parks the result in shared memory; warp 0 then adds the per-warp
// results in warp order. The order is fixed by (splits, warps) alone, so every CTA and every run gets the same bits.
__device__ __forceinline__ void gn_fold_item(const float* __restrict__ partial, int64_t item, int splits, int groups,
                                             double cnt, float eps, double2* __restrict__ scratch /*[nfull][64]*/,
                                             float2* __restrict__ stats /*[64]*/) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nfull = blockDim.x >> 5;
  if (warp < nfull) {
    for (int g = lane; g < groups; g += 32) {
      double a = 0.0, q = 0.0;
      for (int sp = warp; sp < splits; sp += nfull) {
        const float2 p = __ldg(reinterpret_cast<const float2*>(partial + (((item * splits + sp) * groups + g) << 1)));
        a += (double)p.x;
        q += (double)p.y;
      }
      scratch[warp * 64 + g] = make_double2(a, q);
    }
  }
  __syncthreads();
  if (warp == 0) {
    for (int g = lane; g < groups; g += 32) {
      double a = 0.0, q = 0.0;
      for (int w = 0; w < nfull; ++w) {
        const double2 v = scratch[w * 64 + g];
        a += v.x;
        q += v.y;
      }
      const double m = a / cnt;
      double var = q / cnt - m * m;
      if (var < 0.0) var = 0.0;
      stats[g] = make_float2((float)m, (float)(1.0 / sqrt(var + (double)eps)));
    }
  }
  __syncthreads();
}

// Statistics that arrive from the producing GEMM's epilogue come as one {sum, sumsq} per 32-row block and group. An apply
// CTA that folded all of them itself paid for it (ncu rd2j: 288 blocks per frame made gn_apply 90 us instead of 68), so
// this kernel folds them first: grid = (parts, items), each CTA a contiguous range of blocks, lane = group (a warp reads
// whole 256-byte rows of the table: coalesced), warps stride over the blocks, fp64, fixed order; the apply pass then folds
// `parts` (<= 16) partials per item.
__global__ void __launch_bounds__(256)
gn_prefold_kernel(const float* __restrict__ pre, int splits, int groups, float* __restrict__ out /*[item][part][group][2]*/) {
  __shared__ double2 sh[8][64];
  pdl_wait();
  const int part = blockIdx.x, parts = gridDim.x, item = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int per = (splits + parts - 1) / parts;
  const int s0 = part * per, s1 = (s0 + per < splits) ? s0 + per : splits;
  for (int g = lane; g < groups; g += 32) {
    double a = 0.0, q = 0.0;
    for (int sp = s0 + warp; sp < s1; sp += 8) {
      const float2 v = __ldg(reinterpret_cast<const float2*>(pre + ((((int64_t)item * splits + sp) * groups + g) << 1)));
      a += (double)v.x;
      q += (double)v.y;
    }
    sh[warp][g] = make_double2(a, q);
  }
  __syncthreads();
  if (warp == 0) {
    for (int g = lane; g < groups; g += 32) {
      double a = 0.0, q = 0.0;
#pragma unroll
      for (int w = 0; w < 8; ++w) { a += sh[w][g].x; q += sh[w][g].y; }
      float* dst = out + ((((int64_t)item * parts + part) * groups + g) << 1);
      dst[0] = (float)a;  // re-read by the apply pass as float partials
      dst[1] = (float)q;
    }
  }
}

__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// SiLU with ONE MUFU op per element (x * sigmoid(x) = 0.5 x (1 + tanh(x/2))): with exp + reciprocal the two MUFU ops
// per element cost more than the HBM traffic of the kernel.
__device__ __forceinline__ float silu_fast(float x) {
  const float hx = 0.5f * x;
  return fmaf(hx, tanh_approx(hx), hx);
}

// pass 2: y = x * scale + shift with scale = rstd*gamma, shift = beta - mean*rstd*gamma folded once per thread
// (a thread owns one fixed 8-channel vector and walks rows), optional SiLU. Same mapping / grid as the stats pass.
template <bool F16>
__global__ void __launch_bounds__(GN_MAX_THREADS)
gn_apply_kernel(const __nv_bfloat16* __restrict__ x1, int C1, const __nv_bfloat16* __restrict__ x2, int C2,
                const float* __restrict__ gamma, const float* __restrict__ beta, __nv_bfloat16* __restrict__ y,
                int64_t rows_per_item, int groups, int rpp, int silu, const float* __restrict__ partial, int splits,
                double cnt, float eps) {
  __shared__ float2 stats[64];  // {mean, rstd} of this CTA's item, per group
  __shared__ double2 fold_scratch[(GN_MAX_THREADS / 32) * 64];
  pdl_wait();
  gn_fold_item(partial, blockIdx.y, splits, groups, cnt, eps, fold_scratch, stats);
  const int C = C1 + C2;
  const int vecs = C >> 3;
  const int cpg = C / groups;
  const int item = blockIdx.y;
  const int ty = threadIdx.x / vecs, tx = threadIdx.x - ty * vecs;
  const int c = tx << 3;
  const int64_t rows_per_cta = (rows_per_item + gridDim.x - 1) / gridDim.x;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_cta;
  int64_t r1 = r0 + rows_per_cta;
  if (r1 > rows_per_item) r1 = rows_per_item;
  const int64_t base = (int64_t)item * rows_per_item;
  float sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float2 st = stats[(c + j) / cpg];
    const float gm = __ldg(gamma + c + j), bt = __ldg(beta + c + j);
    sc[j] = st.y * gm;
    sh[j] = fmaf(-st.x, st.y * gm, bt);
  }
  for (int64_t r = r0 + ty; r < r1; r += (int64_t)rpp * GN_UNROLL) {
    uint4 u[GN_UNROLL];
#pragma unroll
    for (int k = 0; k < GN_UNROLL; ++k) {
      const int64_t rr = r + (int64_t)k * rpp;
      if (rr < r1) u[k] = gn_load(x1, C1, x2, C2, base + rr, c);
    }
#pragma unroll
    for (int k = 0; k < GN_UNROLL; ++k) {
      const int64_t rr = r + (int64_t)k * rpp;
      if (rr < r1) {
        const uint32_t uu[4] = {u[k].x, u[k].y, u[k].z, u[k].w};
        float f[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 t = unpack_act2<F16>(uu[j]);
          f[2 * j] = fmaf(t.x, sc[2 * j], sh[2 * j]);
          f[2 * j + 1] = fmaf(t.y, sc[2 * j + 1], sh[2 * j + 1]);
        }
        if (silu) {
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] = silu_fast(f[j]);
        }
        uint4 o;
        o.x = pack_act2<F16>(f[0], f[1]); o.y = pack_act2<F16>(f[2], f[3]);
        o.z = pack_act2<F16>(f[4], f[5]); o.w = pack_act2<F16>(f[6], f[7]);
        *reinterpret_cast<uint4*>(y + (base + rr) * C + c) = o;
      }
    }
  }
}

// ---------------- LayerNorm: LPR lanes per row, V 16-byte vectors per lane, row cached in registers ----------------
// A warp handles 32/LPR rows at once; lane l of a row group loads vectors l, l+LPR, ... (each load instruction reads
// LPR*16 contiguous bytes per row: full sectors, every lane busy, V independent loads in flight per lane).
template <int LPR, int V, bool F16>
__global__ void __launch_bounds__(256)
layer_norm_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
                  __nv_bfloat16* __restrict__ y, int64_t rows, int C, float eps) {
  constexpr int RPW = 32 / LPR;  // rows per warp
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int sub = lane / LPR, l = lane % LPR;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const float inv_c = 1.0f / (float)C;
  for (int64_t r0 = warp * RPW; r0 < rows; r0 += nwarps * RPW) {
    const int64_t r = r0 + sub;
    const bool live = r < rows;
    uint4 u[V];
#pragma unroll
    for (int i = 0; i < V; ++i)
      u[i] = live ? __ldg(reinterpret_cast<const uint4*>(x + r * C + ((l + LPR * i) << 3))) : make_uint4(0, 0, 0, 0);
    float f[V][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < V; ++i) {
      const uint32_t uu[4] = {u[i].x, u[i].y, u[i].z, u[i].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 t = unpack_act2<F16>(uu[j]);
        f[i][2 * j] = t.x; f[i][2 * j + 1] = t.y;
        s += t.x + t.y;
      }
    }
#pragma unroll
    for (int o = LPR / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s * inv_c;
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < V; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) { const float d = f[i][j] - mean; ss += d * d; }
#pragma unroll
    for (int o = LPR / 2; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    const float rstd = rsqrtf(ss * inv_c + eps);
    if (live) {
#pragma unroll
      for (int i = 0; i < V; ++i) {
        const int c = (l + LPR * i) << 3;
        const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + c)), g1 = __ldg(reinterpret_cast<const float4*>(gamma + c + 4));
        const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta + c)), b1 = __ldg(reinterpret_cast<const float4*>(beta + c + 4));
        const float gm[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
        const float bt[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (f[i][j] - mean) * rstd * gm[j] + bt[j];
        uint4 w;
        w.x = pack_act2<F16>(o[0], o[1]); w.y = pack_act2<F16>(o[2], o[3]);
        w.z = pack_act2<F16>(o[4], o[5]); w.w = pack_act2<F16>(o[6], o[7]);
        *reinterpret_cast<uint4*>(y + r * C + c) = w;
      }
    }
  }
}

// ---------------- layout helpers ----------------
__global__ void upsample2x_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, int N, int H, int W, int vecs) {
  const int64_t total = (int64_t)N * (2 * H) * (2 * W) * vecs;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int v = (int)(i % vecs);
    int64_t p = i / vecs;
    const int ox = (int)(p % (2 * W)); p /= (2 * W);
    const int oy = (int)(p % (2 * H));
    const int n = (int)(p / (2 * H));
    y[i] = __ldg(x + (((int64_t)n * H + (oy >> 1)) * W + (ox >> 1)) * vecs + v);
  }
}

__global__ void concat_kernel(const uint4* __restrict__ a, int va, const uint4* __restrict__ b, int vb,
                              uint4* __restrict__ y, int64_t rows) {
  const int vt = va + vb;
  const int64_t total = rows * vt;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / vt;
    const int v = (int)(i % vt);
    y[i] = v < va ? __ldg(a + r * va + v) : __ldg(b + r * vb + (v - va));
  }
}

// explicit im2col: [N,H,W,C] -> [N*P*Q, 9*C] (tap-major, then channel), zero padded
__global__ void im2col3x3_kernel(const uint4* __restrict__ x, uint4* __restrict__ col, int N, int H, int W, int vecs,
                                 int stride, int P, int Q) {
  const int64_t total = (int64_t)N * P * Q * 9 * vecs;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int v = (int)(i % vecs);
    int64_t t = i / vecs;
    const int tap = (int)(t % 9); t /= 9;
    const int q = (int)(t % Q); t /= Q;
    const int p = (int)(t % P);
    const int n = (int)(t / P);
    const int iy = p * stride - 1 + tap / 3, ix = q * stride - 1 + tap % 3;
    uint4 val = make_uint4(0, 0, 0, 0);
    if (iy >= 0 && iy < H && ix >= 0 && ix < W) val = __ldg(x + (((int64_t)n * H + iy) * W + ix) * vecs + v);
    col[i] = val;
  }
}

inline int grid_1d(int64_t work_items, int block, int per_sm) {
  int64_t g = (work_items + block - 1) / block;
  const int64_t cap = (int64_t)sm_count() * per_sm;
  if (g > cap) g = cap;
  return g < 1 ? 1 : (int)g;
}

// CTAs of `kernel` the device holds at once at this block size (registers / threads / shared memory, asked of the
// runtime). The HBM-stream kernels size their grids to whole waves of it: a grid of 2.02 waves costs 3 (ncu r01b: the
// GroupNorm passes ran 896 CTAs on 444 slots).
template <class K>
int resident_slots(K kernel, int threads, size_t smem) {
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, threads, smem) != cudaSuccess || n < 1) n = 1;
  return n * sm_count();
}

// grid of a grid-stride kernel: everything resident at once, never more CTAs than there is work
template <class K>
int grid_resident(K kernel, int64_t work_items, int block, size_t smem = 0) {
  const int64_t need = (work_items + block - 1) / block;
  const int64_t slots = resident_slots(kernel, block, smem);
  const int64_t g = need < slots ? need : slots;
  return g < 1 ? 1 : (int)g;
}

}  // namespace

// rows per pass: as many as fit GN_MAX_THREADS threads of one 16-byte vector each
static int gn_rpp(int C) {
  const int vecs = C >> 3;
  int rpp = GN_MAX_THREADS / vecs;
  return rpp < 1 ? 1 : rpp;
}

// CTAs per item. The grid (splits x n_items) is sized to 1-3 WHOLE waves of the resident slots, whichever wastes least
// (each CTA walks a contiguous row range, so a partial last wave costs a full one), at least 32 rows per CTA.
static int gn_splits(int64_t n_items, int64_t rows_per_item, int64_t slots) {
  const int64_t max_splits = (rows_per_item + 31) / 32;
  int64_t best = 1;
  double best_util = -1.0;
  for (int w = 1; w <= 3; ++w) {
    int64_t sp = (w * slots) / n_items;
    if (sp > max_splits) sp = max_splits;
    if (sp < 1) sp = 1;
    const int64_t ctas = sp * n_items;
    const int64_t waves = (ctas + slots - 1) / slots;
    const double util = (double)ctas / (double)(waves * slots);
    if (util > best_util + 0.02) { best_util = util; best = sp; }
  }
  return (int)best;
}

// upper bound of gn_splits over every block size the kernels use (<= 10 CTAs of >= 193 threads per SM, 3 waves)
static int64_t gn_splits_bound(int64_t n_items, int64_t rows_per_item) {
  int64_t sp = ((int64_t)sm_count() * 30 + n_items - 1) / n_items;
  const int64_t max_splits = (rows_per_item + 31) / 32;
  if (sp > max_splits) sp = max_splits;
  return sp < 1 ? 1 : sp;
}

size_t group_norm_scratch_bytes(int64_t n_items, int64_t rows_per_item, int groups) {
  const size_t partial = sizeof(float) * 2 * (size_t)n_items * gn_splits_bound(n_items, rows_per_item) * groups;
  return ((partial + 255) & ~(size_t)255) + sizeof(float2) * (size_t)n_items * groups;
}

constexpr int GN_DIRECT_FOLD_MAX = 24;   // 32-row blocks per item an apply CTA folds by itself; more go through gn_prefold

template <bool F16>
static int group_norm_t(const void* x1, int C1, const void* x2, int C2, const float* gamma, const float* beta, void* y,
                        int64_t n_items, int64_t rows_per_item, int groups, float eps, int silu, void* scratch,
                        const float* pre, cudaStream_t s) {
  const int C = C1 + C2;
  B2V_CHECK(C % 8 == 0 && C1 % 8 == 0 && C % groups == 0 && (C / groups) % 2 == 0, "group_norm: C=%d (C1=%d) groups=%d unsupported", C, C1, groups);
  B2V_CHECK(C <= GN_MAX_THREADS * 8, "group_norm: C=%d too wide (max %d)", C, GN_MAX_THREADS * 8);
  B2V_CHECK(groups <= 64, "group_norm: at most 64 groups");
  B2V_CHECK(n_items > 0 && n_items < 65536 && rows_per_item > 0, "group_norm: bad item count");
  const int rpp = gn_rpp(C);
  const int threads = (C >> 3) * rpp;
  const size_t shb = sizeof(float) * 2 * rpp * (C >> 1);
  const int s1 = resident_slots(gn_stats_kernel<F16>, threads, shb), s2 = resident_slots(gn_apply_kernel<F16>, threads, 0);
  const int splits = gn_splits(n_items, rows_per_item, s1 < s2 ? s1 : s2);
  float* partial = (float*)scratch;
  const size_t pbytes = sizeof(float) * 2 * (size_t)n_items * splits * groups;
  B2V_CHECK(pbytes + 256 + sizeof(float2) * (size_t)n_items * groups <= group_norm_scratch_bytes(n_items, rows_per_item, groups) + 256,
            "group_norm: scratch too small for %d splits", splits);
  dim3 grid((unsigned)splits, (unsigned)n_items);
  const float* fold_src = partial;
  int fold_splits = splits;
  if (pre) {
    // statistics emitted by the producing GEMM (one partial per 32-row block): no pass over the activation
    B2V_CHECK(C2 == 0 && rows_per_item % 32 == 0, "group_norm: precomputed statistics need whole 32-row blocks, no concat");
    const int64_t blocks = rows_per_item / 32;
    static int direct_max = -1;  // B2V_GN_DIRECT_FOLD=<blocks>: A/B knob
    if (direct_max < 0) { const char* e = getenv("B2V_GN_DIRECT_FOLD"); direct_max = e ? atoi(e) : GN_DIRECT_FOLD_MAX; }
    if (blocks <= direct_max) {
      fold_src = pre;
      fold_splits = (int)blocks;
    } else {
      // enough CTAs to pull the table out of L2 quickly (>= 64 blocks each), at most 16 partials per item for the apply
      int parts = (int)((blocks + 63) / 64);
      const int want = (int)((2 * (int64_t)sm_count() + n_items - 1) / n_items);
      if (parts > want) parts = want;
      if (parts > 16) parts = 16;
      if (parts < 1) parts = 1;
      B2V_CHECK(sizeof(float) * 2 * (size_t)n_items * parts * groups <= group_norm_scratch_bytes(n_items, rows_per_item, groups),
                "group_norm: scratch too small for the pre-fold");
      B2V_CUDA_OK(launch_pdl(gn_prefold_kernel, dim3((unsigned)parts, (unsigned)n_items), dim3(256), 0, s, pre,
                             (int)blocks, groups, partial));
      count_launch();
      fold_splits = parts;
    }
  } else {
    B2V_CUDA_OK(launch_pdl(gn_stats_kernel<F16>, grid, dim3(threads), shb, s, (const __nv_bfloat16*)x1, C1,
                           (const __nv_bfloat16*)x2, C2, rows_per_item, groups, rpp, partial));
    count_launch();
  }
  B2V_CUDA_OK(launch_pdl(gn_apply_kernel<F16>, grid, dim3(threads), 0, s, (const __nv_bfloat16*)x1, C1,
                         (const __nv_bfloat16*)x2, C2, gamma, beta, (__nv_bfloat16*)y, rows_per_item, groups, rpp, silu,
                         fold_src, fold_splits, (double)rows_per_item * (double)(C / groups), eps));
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

int group_norm(int adt, const void* x1, int C1, const void* x2, int C2, const float* gamma, const float* beta, void* y,
               int64_t n_items, int64_t rows_per_item, int groups, float eps, int silu, void* scratch,
               cudaStream_t s, const float* pre) {
  B2V_DISPATCH_ACT(adt, F16, return group_norm_t<F16>(x1, C1, x2, C2, gamma, beta, y, n_items, rows_per_item, groups, eps,
                                                      silu, scratch, pre, s));
  return 0;
}

template <bool F16>
static int layer_norm_t(const void* x, const float* gamma, const float* beta, void* y, int64_t rows, int C, float eps,
                        cudaStream_t s) {
  B2V_CHECK(C % 64 == 0 && C <= 2048 && rows > 0, "layer_norm: C=%d unsupported (multiple of 64, <= 2048)", C);
  const int vecs = C >> 3;
  // lanes per row x vectors per lane = vecs, with <= 8 vectors per lane and every lane of the warp busy
  int lpr = 8;
  while (lpr < 32 && (vecs % lpr != 0 || vecs / lpr > 8 || (vecs % (lpr * 2) == 0 && vecs / lpr > 5))) lpr *= 2;
  B2V_CHECK(vecs % lpr == 0 && vecs / lpr <= 8, "layer_norm: C=%d has no lane mapping", C);
  const int v = vecs / lpr;
  const int64_t warps = (rows + (32 / lpr) - 1) / (32 / lpr);
#define B2V_LN(L, V)                                                                                                 \
  if (lpr == L && v == V) {                                                                                          \
    const int grid = grid_resident(layer_norm_kernel<L, V, F16>, warps * 32, 256);                                   \
    B2V_CUDA_OK(launch_pdl(layer_norm_kernel<L, V, F16>, dim3(grid), dim3(256), 0, s, (const __nv_bfloat16*)x, gamma,   \
                           beta, (__nv_bfloat16*)y, rows, C, eps));                                                   \
    launched = true;                                                                                                 \
  }
  bool launched = false;
  B2V_LN(8, 1) B2V_LN(8, 2) B2V_LN(8, 3) B2V_LN(8, 4) B2V_LN(8, 5)
  B2V_LN(16, 3) B2V_LN(16, 4) B2V_LN(16, 5) B2V_LN(16, 7)
  B2V_LN(32, 3) B2V_LN(32, 4) B2V_LN(32, 5) B2V_LN(32, 6) B2V_LN(32, 7) B2V_LN(32, 8)
#undef B2V_LN
  if (!launched) return fail(B2V_ERR_ARG, "layer_norm: C=%d (lanes %d x vectors %d) not instantiated", C, lpr, v);
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

int layer_norm(int adt, const void* x, const float* gamma, const float* beta, void* y, int64_t rows, int C, float eps,
               cudaStream_t s) {
  B2V_DISPATCH_ACT(adt, F16, return layer_norm_t<F16>(x, gamma, beta, y, rows, C, eps, s));
  return 0;
}

int upsample2x(const void* x, void* y, int N, int H, int W, int C, cudaStream_t s) {
  B2V_CHECK(C % 8 == 0, "upsample2x: C %% 8");
  const int64_t total = (int64_t)N * 4 * H * W * (C >> 3);
  upsample2x_kernel<<<grid_1d(total, 256, 8), 256, 0, s>>>((const uint4*)x, (uint4*)y, N, H, W, C >> 3);
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

int concat_channels(const void* a, int Ca, const void* b, int Cb, void* y, int64_t rows, cudaStream_t s) {
  B2V_CHECK(Ca % 8 == 0 && Cb % 8 == 0, "concat: C %% 8");
  const int64_t total = rows * ((Ca + Cb) >> 3);
  concat_kernel<<<grid_1d(total, 256, 8), 256, 0, s>>>((const uint4*)a, Ca >> 3, (const uint4*)b, Cb >> 3, (uint4*)y,
                                                       rows);
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

int im2col3x3(const void* x, void* col, int N, int H, int W, int C, int stride, cudaStream_t s) {
  B2V_CHECK(C % 8 == 0, "im2col: C %% 8");
  const int P = (H - 1) / stride + 1, Q = (W - 1) / stride + 1;
  const int64_t total = (int64_t)N * P * Q * 9 * (C >> 3);
  im2col3x3_kernel<<<grid_1d(total, 256, 8), 256, 0, s>>>((const uint4*)x, (uint4*)col, N, H, W, C >> 3, stride, P, Q);
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

}  // namespace b2v
