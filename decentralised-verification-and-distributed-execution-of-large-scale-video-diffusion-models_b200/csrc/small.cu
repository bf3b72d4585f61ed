// SIMT kernels for the thin ends of the UNet3D ([UPSTREAM diffusers] UNet3DConditionModel.forward):
// sinusoidal timestep embedding + time MLP, per-resnet time projections (M = batch, far too small for the
// tensor cores) and the NCFHW -> channels-last staging of the latent for conv_in (conv_in / conv_out themselves run on
// the tcgen05 implicit-GEMM kernel with zero-padded channels, gemm.cu).
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
__global__ void timestep_sinusoid_kernel(const float* __restrict__ t_dev, int dim, int B, float* __restrict__ out) {
  const float t = *t_dev;  // read from device memory: the kernel is replayed from a CUDA graph with a new timestep
  const int half = dim / 2;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < B * dim; i += gridDim.x * blockDim.x) {
    const int j = i % dim;
    const int k = j < half ? j : j - half;
    const float freq = expf(-logf(10000.0f) * (float)k / (float)half);
    const float a = t * freq;
    out[i] = j < half ? cosf(a) : sinf(a);
  }
}

// NCFHW (io dtype) -> channels-last bf16 rows [B*F*HW, Cpad] (channels >= Cin are zero): the tensor-core conv_in reads
// the latent as a 64-channel image. One thread per (pixel, 8-channel vector): 16-byte coalesced stores.
template <typename T, bool F16>
__global__ void ncfhw_to_rows_kernel(const T* __restrict__ x, uint16_t* __restrict__ y, int B, int Cin, int F,
                                     int64_t HW, int Cpad) {
  const int vecs = Cpad >> 3;
  const int64_t total = (int64_t)B * F * HW * vecs;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int v = (int)(i % vecs);
    const int64_t row = i / vecs;
    const int64_t pix = row % HW;
    const int64_t n = row / HW;
    const int64_t b = n / F, f = n % F;
    float e[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int ci = v * 8 + j;
      e[j] = ci < Cin ? DT<T>::ld(x + ((b * Cin + ci) * F + f) * HW + pix) : 0.f;
    }
    uint4 o;
    o.x = pack_act2<F16>(e[0], e[1]); o.y = pack_act2<F16>(e[2], e[3]);
    o.z = pack_act2<F16>(e[4], e[5]); o.w = pack_act2<F16>(e[6], e[7]);
    reinterpret_cast<uint4*>(y)[i] = o;
  }
}

template <typename T, bool F16>
__global__ void to_act_kernel(const T* __restrict__ x, uint16_t* __restrict__ y, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    y[i] = to_act<F16>(DT<T>::ld(x + i));
}

__global__ void set_scalar_kernel(float* dst, float v) { *dst = v; }

}  // namespace

int set_scalar(float* dst_dev, float v, cudaStream_t s) {
  set_scalar_kernel<<<1, 1, 0, s>>>(dst_dev, v);
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

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
int time_embedding(const float* timestep_dev, int B, int dim0, const float* w1, const float* b1, const float* w2,
                   const float* b2, int temb_dim, float* silu_emb_out, cudaStream_t s) {
  float* sinus = silu_emb_out + (int64_t)B * temb_dim;
  float* hid = sinus + (int64_t)B * dim0;
  timestep_sinusoid_kernel<<<(B * dim0 + 255) / 256, 256, 0, s>>>(timestep_dev, dim0, B, sinus);
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

int ncfhw_to_rows(const void* x, int io_dtype, void* y, int adt, int B, int Cin, int F, int64_t HW, int Cpad,
                  cudaStream_t s) {
  B2V_CHECK(Cpad % 8 == 0 && Cin <= Cpad, "ncfhw_to_rows: bad channels %d -> %d", Cin, Cpad);
  const int64_t total = (int64_t)B * F * HW * (Cpad >> 3);
  int64_t grid = (total + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (grid > cap) grid = cap;
  B2V_DISPATCH_ACT(adt, F16, B2V_DISPATCH_DTYPE(io_dtype, T, (ncfhw_to_rows_kernel<T, F16><<<(unsigned)grid, 256, 0, s>>>(
                                                                  (const T*)x, (uint16_t*)y, B, Cin, F, HW, Cpad))));
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

int convert_to_act(const void* x, int dtype, void* y, int adt, int64_t n, cudaStream_t s) {
  int64_t grid = (n + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 8;
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  B2V_DISPATCH_ACT(adt, F16, B2V_DISPATCH_DTYPE(dtype, T, (to_act_kernel<T, F16><<<(unsigned)grid, 256, 0, s>>>(
                                                               (const T*)x, (uint16_t*)y, n))));
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

}  // namespace b2v
