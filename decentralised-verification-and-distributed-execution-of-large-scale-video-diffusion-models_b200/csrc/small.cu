// SIMT kernels for the thin ends of the UNet3D ([UPSTREAM diffusers] UNet3DConditionModel.forward):
// sinusoidal timestep embedding + time MLP, per-resnet time projections (M = batch, far too small for the
// tensor cores), conv_in (4 -> 320 channels, NCFHW -> channels-last) and conv_out (320 -> 4, back to NCFHW).
#include "common.cuh"
#include "ops.h"

namespace b2v {
namespace {

// out[b,n] = act_out( bias[n] + sum_k W[n,k] * in[b,k] ), one warp per output
__global__ void small_linear_kernel(const float* __restrict__ in, const float* __restrict__ w,
                                    const float* __restrict__ bias, float* __restrict__ out, int B, int N, int K,
                                    int silu_out) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (warp >= (int64_t)B * N) return;
  const int b = (int)(warp / N), n = (int)(warp % N);
  const float* wr = w + (int64_t)n * K;
  const float* xr = in + (int64_t)b * K;
  float acc = 0.f;
  for (int k = lane; k < K; k += 32) acc += __ldg(wr + k) * __ldg(xr + k);
  acc = warp_sum(acc);
  if (lane == 0) {
    float v = acc + (bias ? bias[n] : 0.f);
    if (silu_out) v = silu_f(v);
    out[(int64_t)b * N + n] = v;
  }
}

// Timesteps(dim, flip_sin_to_cos=True, downscale_freq_shift=0): [cos(t f_i), sin(t f_i)], f_i = exp(-ln(1e4) i/half)
__global__ void timestep_sinusoid_kernel(float t, int dim, int B, float* __restrict__ out) {
  const int half = dim / 2;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < B * dim; i += gridDim.x * blockDim.x) {
    const int j = i % dim;
    const int k = j < half ? j : j - half;
    const float freq = expf(-logf(10000.0f) * (float)k / (float)half);
    const float a = t * freq;
    out[i] = j < half ? cosf(a) : sinf(a);
  }
}

// conv_in: one thread per (pixel, 8 output channels); weights staged in shared memory as [k][Cout]
template <typename T>
__global__ void __launch_bounds__(256)
conv_in_kernel(const T* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
               __nv_bfloat16* __restrict__ y, int B, int Cin, int F, int H, int W, int Cout) {
  extern __shared__ float sw[];  // [Cin*9][Cout]
  const int K = Cin * 9;
  for (int i = threadIdx.x; i < K * Cout; i += blockDim.x) {
    const int co = i % Cout, k = i / Cout;  // k = ci*9 + tap  (PyTorch [Cout][Cin][3][3])
    sw[i] = w[(int64_t)co * K + k];
  }
  __syncthreads();
  const int vecs = Cout >> 3;
  const int64_t total = (int64_t)B * F * H * W * vecs;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int v = (int)(i % vecs);
    int64_t p = i / vecs;
    const int xx = (int)(p % W); p /= W;
    const int yy = (int)(p % H); p /= H;
    const int f = (int)(p % F);
    const int b = (int)(p / F);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = bias[v * 8 + j];
    for (int ci = 0; ci < Cin; ++ci) {
      const T* plane = x + (((int64_t)b * Cin + ci) * F + f) * H * W;
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) {
        const int iy = yy + tap / 3 - 1, ix = xx + tap % 3 - 1;
        if (iy < 0 || iy >= H || ix < 0 || ix >= W) continue;
        const float xv = DT<T>::ld(plane + (int64_t)iy * W + ix);
        const float* wr = sw + (ci * 9 + tap) * Cout + v * 8;
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += xv * wr[j];
      }
    }
    uint4 o;
    o.x = pack_bf16x2(acc[0], acc[1]); o.y = pack_bf16x2(acc[2], acc[3]);
    o.z = pack_bf16x2(acc[4], acc[5]); o.w = pack_bf16x2(acc[6], acc[7]);
    // row = ((b*F + f)*H + yy)*W + xx  == i / vecs
    *reinterpret_cast<uint4*>(y + (i / vecs) * Cout + v * 8) = o;
  }
}

// conv_out: one warp per output pixel, lanes split the input channels; weights [Cout][9][Cin] in smem
template <typename T>
__global__ void __launch_bounds__(256)
conv_out_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                T* __restrict__ y, int B, int Cin, int F, int H, int W, int Cout) {
  extern __shared__ float sw[];  // [Cout][9][Cin]
  for (int i = threadIdx.x; i < Cout * 9 * Cin; i += blockDim.x) sw[i] = w[i];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int vecs = Cin >> 3;
  const int64_t npix = (int64_t)B * F * H * W;
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t pix = warp0; pix < npix; pix += nwarps) {
    int64_t p = pix;
    const int xx = (int)(p % W); p /= W;
    const int yy = (int)(p % H); p /= H;
    const int f = (int)(p % F);
    const int b = (int)(p / F);
    float acc[4] = {0.f, 0.f, 0.f, 0.f};  // Cout <= 4
    for (int tap = 0; tap < 9; ++tap) {
      const int iy = yy + tap / 3 - 1, ix = xx + tap % 3 - 1;
      if (iy < 0 || iy >= H || ix < 0 || ix >= W) continue;
      const __nv_bfloat16* src = x + ((((int64_t)b * F + f) * H + iy) * W + ix) * Cin;
      for (int v = lane; v < vecs; v += 32) {
        const uint4 u = __ldg(reinterpret_cast<const uint4*>(src + v * 8));
        const float2 a = unpack_bf16x2(u.x), bb = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
        const float xv[8] = {a.x, a.y, bb.x, bb.y, c.x, c.y, d.x, d.y};
        for (int co = 0; co < Cout; ++co) {
          const float* wr = sw + ((co * 9 + tap) * Cin) + v * 8;
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[co] += xv[j] * wr[j];
        }
      }
    }
    for (int co = 0; co < Cout; ++co) acc[co] = warp_sum(acc[co]);
    if (lane == 0) {
      for (int co = 0; co < Cout; ++co)
        DT<T>::st(y + ((((int64_t)b * Cout + co) * F + f) * H + yy) * W + xx, acc[co] + bias[co]);
    }
  }
}

template <typename T>
__global__ void to_bf16_kernel(const T* __restrict__ x, __nv_bfloat16* __restrict__ y, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    y[i] = __float2bfloat16_rn(DT<T>::ld(x + i));
}

}  // namespace

int small_linear(const float* in, const float* w, const float* bias, float* out, int B, int N, int K,
                 cudaStream_t s) {
  const int64_t threads = (int64_t)B * N * 32;
  small_linear_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(in, w, bias, out, B, N, K, 0);
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

// SiLU(time_embedding(Timesteps(t))) — every consumer (ResnetBlock2D.time_emb_proj) applies SiLU first.
// scratch layout inside silu_emb_out's caller-provided buffer: [B,temb] result, then [B,dim0] + [B,temb] temporaries.
int time_embedding(float timestep, int B, int dim0, const float* w1, const float* b1, const float* w2,
                   const float* b2, int temb_dim, float* silu_emb_out, cudaStream_t s) {
  float* sinus = silu_emb_out + (int64_t)B * temb_dim;
  float* hid = sinus + (int64_t)B * dim0;
  timestep_sinusoid_kernel<<<(B * dim0 + 255) / 256, 256, 0, s>>>(timestep, dim0, B, sinus);
  count_launch();
  {
    const int64_t threads = (int64_t)B * temb_dim * 32;
    small_linear_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(sinus, w1, b1, hid, B, temb_dim, dim0, 1);
    count_launch();
    small_linear_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(hid, w2, b2, silu_emb_out, B, temb_dim,
                                                                        temb_dim, 1);
    count_launch();
  }
  B2V_CUDA_OK(launch_check());
  return 0;
}

int conv_in_3x3(const void* x, int io_dtype, const float* w, const float* bias, void* y, int B, int Cin, int F,
                int H, int W, int Cout, cudaStream_t s) {
  B2V_CHECK(Cout % 8 == 0 && Cin * 9 * Cout * 4 <= 200 * 1024, "conv_in: unsupported channels %d -> %d", Cin, Cout);
  const size_t shb = sizeof(float) * (size_t)Cin * 9 * Cout;
  const int64_t total = (int64_t)B * F * H * W * (Cout >> 3);
  int64_t grid = (total + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 4;
  if (grid > cap) grid = cap;
  B2V_DISPATCH_DTYPE(io_dtype, T, {
    B2V_CUDA_OK(cudaFuncSetAttribute(conv_in_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shb));
    conv_in_kernel<T><<<(unsigned)grid, 256, shb, s>>>((const T*)x, w, bias, (__nv_bfloat16*)y, B, Cin, F, H, W, Cout);
  });
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

int conv_out_3x3(const void* x, const float* w, const float* bias, void* y, int io_dtype, int B, int Cin, int F,
                 int H, int W, int Cout, cudaStream_t s) {
  B2V_CHECK(Cin % 8 == 0 && Cout <= 4 && Cout * 9 * Cin * 4 <= 200 * 1024, "conv_out: unsupported channels %d -> %d",
            Cin, Cout);
  const size_t shb = sizeof(float) * (size_t)Cout * 9 * Cin;
  const int64_t npix = (int64_t)B * F * H * W;
  int64_t grid = (npix + 7) / 8;
  const int64_t cap = (int64_t)sm_count() * 4;
  if (grid > cap) grid = cap;
  B2V_DISPATCH_DTYPE(io_dtype, T, {
    B2V_CUDA_OK(cudaFuncSetAttribute(conv_out_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shb));
    conv_out_kernel<T><<<(unsigned)grid, 256, shb, s>>>((const __nv_bfloat16*)x, w, bias, (T*)y, B, Cin, F, H, W, Cout);
  });
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

int to_bf16(const void* x, int dtype, void* y, int64_t n, cudaStream_t s) {
  int64_t grid = (n + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 8;
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  B2V_DISPATCH_DTYPE(dtype, T, (to_bf16_kernel<T><<<(unsigned)grid, 256, 0, s>>>((const T*)x, (__nv_bfloat16*)y, n)));
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

}  // namespace b2v
