// Kernels of the frame decode that follows the denoise loop:
//   per frame `vae.decode(z / 0.18215).sample` -> `(img*0.5+0.5).clamp(0,1)*255` -> uint8 [H,W,3]
//   (Distribution/strategies/chunk_only.py:152-158, InferNet/neurons/miner.py:594-600).
// The decoder's convolutions, 1x1 projections and GroupNorms run on the UNet's kernels (gemm.cu / norm.cu); this file
// adds what the SD VAE needs on top: the row softmax of the single 512-wide attention head of the mid block (its two
// matrix products are plain tcgen05 GEMMs: S = Q K^T, O = P V), and the frame quantisation to bytes.
// Both are HBM-bound: one read + one write of their operand.
#include "common.cuh"
#include "ops.h"

namespace b2v {
namespace {

// P[r, :] = softmax(scale * S[r, :]) in place, bf16 storage, fp32 arithmetic. One CTA per row; the row is read from
// HBM once into shared memory (fp32), reduced twice there and written once.
__global__ void __launch_bounds__(256)
softmax_rows_kernel(__nv_bfloat16* __restrict__ x, int64_t ld, int cols, float scale_log2e) {
  extern __shared__ __align__(16) float row[];
  __shared__ float red[8];
  __nv_bfloat16* p = x + (int64_t)blockIdx.x * ld;
  const int vecs = cols >> 3, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float m = -INFINITY;
  for (int v = tid; v < vecs; v += 256) {
    const uint4 u = *reinterpret_cast<const uint4*>(p + v * 8);
    const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
    *reinterpret_cast<float4*>(row + v * 8) = make_float4(a.x, a.y, b.x, b.y);
    *reinterpret_cast<float4*>(row + v * 8 + 4) = make_float4(c.x, c.y, d.x, d.y);
    m = fmaxf(m, fmaxf(fmaxf(fmaxf(a.x, a.y), fmaxf(b.x, b.y)), fmaxf(fmaxf(c.x, c.y), fmaxf(d.x, d.y))));
  }
  m = warp_max(m);
  if (lane == 0) red[warp] = m;
  __syncthreads();
  m = red[0];
#pragma unroll
  for (int i = 1; i < 8; ++i) m = fmaxf(m, red[i]);
  __syncthreads();
  // scale > 0: max(scale * s) = scale * max(s)
  const float off = m * scale_log2e;
  float sum = 0.f;
  for (int c = tid; c < cols; c += 256) {
    const float e = exp2f(fmaf(row[c], scale_log2e, -off));
    row[c] = e;
    sum += e;
  }
  sum = warp_sum(sum);
  if (lane == 0) red[warp] = sum;
  __syncthreads();
  sum = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) sum += red[i];  // fixed order: bit-reproducible
  const float inv = 1.0f / sum;
  for (int v = tid; v < vecs; v += 256) {
    const float4 a = *reinterpret_cast<const float4*>(row + v * 8), b = *reinterpret_cast<const float4*>(row + v * 8 + 4);
    uint4 u;
    u.x = pack_bf16x2(a.x * inv, a.y * inv);
    u.y = pack_bf16x2(a.z * inv, a.w * inv);
    u.z = pack_bf16x2(b.x * inv, b.y * inv);
    u.w = pack_bf16x2(b.z * inv, b.w * inv);
    *reinterpret_cast<uint4*>(p + v * 8) = u;
  }
}

// rows [n_px, ld] bf16 (channels 0..2 = R,G,B of the decoder output in [-1,1]) -> [n_px, 3] bytes:
// (v*0.5+0.5).clamp(0,1)*255 truncated, the reference's `.byte()` (chunk_only.py:157-158). Four pixels per thread:
// 12 output bytes leave as three 32-bit stores.
__global__ void __launch_bounds__(256)
frames_u8_kernel(const __nv_bfloat16* __restrict__ x, int64_t ld, int64_t n_px, uint8_t* __restrict__ out) {
  const int64_t quads = (n_px + 3) >> 2;
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < quads; q += (int64_t)gridDim.x * blockDim.x) {
    uint8_t b[12];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t px = q * 4 + i;
      float r = 0.f, g = 0.f, bl = 0.f;
      if (px < n_px) {
        const uint2 u = *reinterpret_cast<const uint2*>(x + px * ld);  // 4 channels, 8-byte aligned (ld % 4 == 0)
        const float2 rg = unpack_bf16x2(u.x), bx = unpack_bf16x2(u.y);
        r = rg.x; g = rg.y; bl = bx.x;
      }
      b[3 * i + 0] = (uint8_t)(fminf(fmaxf(fmaf(r, 0.5f, 0.5f), 0.f), 1.f) * 255.f);
      b[3 * i + 1] = (uint8_t)(fminf(fmaxf(fmaf(g, 0.5f, 0.5f), 0.f), 1.f) * 255.f);
      b[3 * i + 2] = (uint8_t)(fminf(fmaxf(fmaf(bl, 0.5f, 0.5f), 0.f), 1.f) * 255.f);
    }
    if (q * 4 + 3 < n_px) {
      uint32_t* o = reinterpret_cast<uint32_t*>(out + q * 12);
#pragma unroll
      for (int i = 0; i < 3; ++i)
        o[i] = (uint32_t)b[4 * i] | ((uint32_t)b[4 * i + 1] << 8) | ((uint32_t)b[4 * i + 2] << 16) |
               ((uint32_t)b[4 * i + 3] << 24);
    } else {
      for (int64_t px = q * 4, i = 0; px < n_px; ++px, ++i)
        for (int c = 0; c < 3; ++c) out[px * 3 + c] = b[3 * i + c];
    }
  }
}

}  // namespace
}  // namespace b2v

using namespace b2v;

extern "C" int b2v_op_softmax_rows(void* x_dev, int64_t rows, int cols, int64_t ld, float scale, void* stream) {
  B2V_CHECK(x_dev && rows > 0 && rows < (1ll << 31), "softmax_rows: bad rows");
  B2V_CHECK(cols > 0 && cols % 8 == 0 && ld >= cols && ld % 8 == 0, "softmax_rows: cols=%d ld=%lld must be multiples of 8", cols,
            (long long)ld);
  B2V_CHECK(scale > 0.f, "softmax_rows: scale must be positive");
  B2V_TRY(b2v_device_check());
  const size_t smem = (size_t)cols * sizeof(float);
  B2V_CHECK(smem <= 200 * 1024, "softmax_rows: a row of %d columns does not fit in shared memory", cols);
  static bool attr_done = false;
  if (!attr_done) {
    B2V_CUDA_OK(cudaFuncSetAttribute(softmax_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_done = true;
  }
  softmax_rows_kernel<<<(unsigned)rows, 256, smem, (cudaStream_t)stream>>>((__nv_bfloat16*)x_dev, ld, cols,
                                                                            scale * 1.4426950408889634f);
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

extern "C" int b2v_op_frames_u8(const void* rows_dev, int64_t ld, int64_t n_px, uint8_t* out_dev, void* stream) {
  B2V_CHECK(rows_dev && out_dev && n_px > 0, "frames_u8: bad argument");
  B2V_CHECK(ld >= 4 && ld % 4 == 0, "frames_u8: ld=%lld must be a multiple of 4 channels", (long long)ld);
  B2V_CHECK(((uintptr_t)out_dev & 3) == 0, "frames_u8: output must be 4-byte aligned");
  B2V_TRY(b2v_device_check());
  const int64_t quads = (n_px + 3) >> 2;
  int64_t blocks = (quads + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  frames_u8_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)rows_dev, ld, n_px, out_dev);
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

extern "C" int b2v_op_upsample2x(const void* x_dev, void* y_dev, int N, int H, int W, int C, void* stream) {
  B2V_CHECK(x_dev && y_dev && N > 0 && H > 0 && W > 0, "upsample2x: bad argument");
  B2V_TRY(b2v_device_check());
  return upsample2x(x_dev, y_dev, N, H, W, C, (cudaStream_t)stream);
}

extern "C" int b2v_op_latent_rows(const void* x_dev, int io_dtype, void* rows_dev, int B, int C, int F, int64_t hw, int Cpad,
                                  int act_dtype, void* stream) {
  B2V_CHECK(x_dev && rows_dev && B > 0 && C > 0 && F > 0 && hw > 0, "latent_rows: bad argument");
  B2V_TRY(b2v_device_check());
  return ncfhw_to_rows(x_dev, io_dtype, rows_dev, act_dtype, B, C, F, hw, Cpad, (cudaStream_t)stream);
}
