// Pipe micro-benchmark for the attention softmax design (sm_100a): how many cycles one SM sub-partition needs per
// warp-instruction of each kind, alone and in the mixes the softmax uses, and cycles per score of complete softmax-tile
// bodies (register-resident, no TMEM / tensor core) with a share of the exponentials on an FMA-pipe polynomial.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_build/ubench_pipes tools/ubench_pipes.cu
//   tools/_build/ubench_pipes            (prints one line per test; run on the GPU box)
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define ITER 512

__device__ __forceinline__ float ex2f(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint64_t pack2(float a, float b) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ uint64_t pack2u(uint32_t a, uint32_t b) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(a), "r"(b)); return r; }
__device__ __forceinline__ void unpack2(uint64_t r, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(r)); }
__device__ __forceinline__ void unpack2u(uint64_t r, uint32_t& a, uint32_t& b) { asm("mov.b64 {%0, %1}, %2;" : "=r"(a), "=r"(b) : "l"(r)); }
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) { uint64_t d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) { uint64_t d; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ float max3(float a, float b, float c) { float d; asm volatile("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }
__device__ __forceinline__ uint32_t cvt_bf2(float lo, float hi) { uint32_t r; asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo)); return r; }

// ---- single-instruction throughput: 8 independent chains per thread ----
template <int OP>
__global__ void op_kernel(float* out, long long* cyc, float seed) {
  float a[8];
  uint64_t p[8];
  uint32_t u[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { a[i] = seed + i * 0.01f + threadIdx.x * 1e-4f; p[i] = pack2(a[i], a[i] + 0.5f); u[i] = __float_as_uint(a[i]); }
  const float c1 = seed * 0.999f, c2 = seed * 0.001f;
  const uint64_t pc1 = pack2(c1, c1), pc2 = pack2(c2, c2);
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITER; ++it) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (OP == 0) a[i] = ex2f(a[i]);
        if (OP == 1) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(c1), "f"(c2));
        if (OP == 2) p[i] = fma2(p[i], pc1, pc2);
        if (OP == 3) p[i] = add2(p[i], pc2);
        if (OP == 4) asm volatile("max.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(c1));
        if (OP == 5) a[i] = max3(a[i], c1, c2);
        if (OP == 6) asm volatile("{.reg .b32 t; shl.b32 t, %0, 23; add.u32 %0, t, %1;}" : "+r"(u[i]) : "r"(u[(i + 1) & 7]));
        if (OP == 7) u[i] = cvt_bf2(a[i], __uint_as_float(u[i]));
        if (OP == 8) { uint16_t h = (uint16_t)u[i]; asm volatile("ex2.approx.ftz.bf16 %0, %0;" : "+h"(h)); u[i] = h; }
        if (OP == 9) { uint16_t h = (uint16_t)u[i]; asm volatile("ex2.approx.f16 %0, %0;" : "+h"(h)); u[i] = h; }
        if (OP == 10) asm volatile("mad.lo.u32 %0, %0, 8388608, %1;" : "+r"(u[i]) : "r"(u[(i + 1) & 7]));
        if (OP == 11) asm volatile("max.s32 %0, %0, %1;" : "+r"(u[i]) : "r"(u[(i + 1) & 7]));
        if (OP == 12) { float y = ex2f(a[i]); asm volatile("add.f32 %0, %1, -1.0;" : "=f"(a[i]) : "f"(y)); }  // stays in [0,1]
        if (OP == 13) { float y = ex2f(c1 - (float)i); asm volatile("add.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(y)); }  // independent
        if (OP == 14) { a[i] = ex2f(a[i]); p[i] = fma2(p[i], pc1, pc2); }
        if (OP == 15) { a[i] = ex2f(a[i]); p[i] = fma2(p[i], pc1, pc2); p[i] = fma2(p[i], pc1, pc2); p[i] = fma2(p[i], pc1, pc2); }
        if (OP == 16) { float y = ex2f(c1 - (float)i); asm volatile("add.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(y)); u[i] = cvt_bf2(y, y); }
      }
    }
  }
  const long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) { float x, y; unpack2(p[i], x, y); s += a[i] + x + y + __uint_as_float(u[i]); }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// ---- a softmax tile body as in attn_kernel: 128 scores per thread in registers -> 64 packed bf16x2 ----
// POLY = pairs out of every 8 whose exponentials go through the FMA-pipe polynomial; SUM = row sum in registers
// (FADD2) or not at all (left to the tensor core); DEG = polynomial degree; CLAMP = per-score lower clamp
template <int POLY, bool SUM, int DEG, bool CLAMP>
__global__ void softmax_kernel(uint32_t* out, long long* cyc, float seed, int iters) {
  // the 128 scores of the row-tile are re-read from shared memory every iteration (the real kernel reads them from TMEM):
  // nothing is loop-invariant, nothing can be hoisted or CSE'd by ptxas
  __shared__ __align__(16) float row[128];
  if (threadIdx.x < 128) row[threadIdx.x] = seed * (float)((threadIdx.x * 37) & 127) - 3.0f;
  float m_used = -1e30f, l = 0.f;
  const float sl2 = 0.125f * 1.4426950408889634f;
  uint32_t acc = 0;
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < iters; ++it) {
    uint32_t sv[128];
#pragma unroll
    for (int i = 0; i < 128; i += 4)
      asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(sv[i]), "=r"(sv[i + 1]), "=r"(sv[i + 2]), "=r"(sv[i + 3])
                   : "r"((uint32_t)__cvta_generic_to_shared(row + i)) : "memory");
    float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
    for (int i = 0; i < 128; i += 8) {
      mx0 = max3(mx0, __uint_as_float(sv[i]), __uint_as_float(sv[i + 1]));
      mx1 = max3(mx1, __uint_as_float(sv[i + 2]), __uint_as_float(sv[i + 3]));
      mx2 = max3(mx2, __uint_as_float(sv[i + 4]), __uint_as_float(sv[i + 5]));
      mx3 = max3(mx3, __uint_as_float(sv[i + 6]), __uint_as_float(sv[i + 7]));
    }
    const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
    if (mx * sl2 - m_used > 8.0f) { const float nb = ceilf(mx * sl2); l *= ex2f(m_used - nb); m_used = nb; }
    const float mb = m_used;
    const float MAGIC = 12582912.0f;  // 1.5 * 2^23
    const float lo = (mb - 125.0f) / sl2;
    const uint64_t sl2_2 = pack2(sl2, sl2), nmb2 = pack2(-mb, -mb), mg2 = pack2(MAGIC - mb, MAGIC - mb);
    const uint64_t neg1 = pack2(-1.f, -1.f);
    const uint64_t c0 = pack2(0.9999280572f, 0.9999280572f), c1 = pack2(0.6932609677f, 0.6932609677f),
                   c2 = pack2(0.2426111251f, 0.2426111251f), c3 = pack2(0.0551716685f, 0.0551716685f);
    const uint64_t d0 = pack2(1.0004431f, 1.0004431f), d1 = pack2(0.70344806f, 0.70344806f), d2 = pack2(0.23842894f, 0.23842894f);
    uint64_t sumA = 0ull, sumB = 0ull;
#pragma unroll
    for (int i = 0; i < 64; ++i) {
      float e0, e1;
      if ((i & 7) < POLY) {
        uint32_t a = sv[2 * i], b = sv[2 * i + 1];
        if (CLAMP) { a = __float_as_uint(fmaxf(__uint_as_float(a), lo)); b = __float_as_uint(fmaxf(__uint_as_float(b), lo)); }
        const uint64_t s2 = pack2u(a, b);
        const uint64_t t2 = fma2(s2, sl2_2, mg2);          // MAGIC + round(x)
        const uint64_t nm = fma2(t2, neg1, mg2);           // -mb - round(x)
        const uint64_t fr = fma2(s2, sl2_2, nm);           // x - round(x) in [-0.5, 0.5]
        uint64_t p;
        if (DEG == 3) p = fma2(fma2(fma2(c3, fr, c2), fr, c1), fr, c0);
        else p = fma2(fma2(d2, fr, d1), fr, d0);
        uint32_t tb0, tb1, pb0, pb1;
        unpack2u(t2, tb0, tb1);
        unpack2u(p, pb0, pb1);
        e0 = __uint_as_float(pb0 + (tb0 << 23));
        e1 = __uint_as_float(pb1 + (tb1 << 23));
      } else {
        const uint64_t x2 = fma2(pack2u(sv[2 * i], sv[2 * i + 1]), sl2_2, nmb2);
        float x0, x1;
        unpack2(x2, x0, x1);
        e0 = ex2f(x0);
        e1 = ex2f(x1);
      }
      if (SUM) {
        if (i & 1) sumB = add2(sumB, pack2(e0, e1));
        else sumA = add2(sumA, pack2(e0, e1));
      }
      const uint32_t pk = cvt_bf2(e0, e1);
      acc ^= pk;
    }
    if (SUM) { float s0, s1; unpack2(add2(sumA, sumB), s0, s1); l += s0 + s1; }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc ^ __float_as_uint(l);
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

static float* d_out;
static long long* d_cyc;

template <int OP>
void run_op(const char* name, int warps_per_smsp) {
  const int threads = 128 * warps_per_smsp;
  op_kernel<OP><<<148, threads>>>(d_out, d_cyc, 0.5f);
  op_kernel<OP><<<148, threads>>>(d_out, d_cyc, 0.5f);
  cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, d_cyc, sizeof(h), cudaMemcpyDeviceToHost);
  double avg = 0;
  for (int i = 0; i < 148; ++i) avg += h[i];
  avg /= 148;
  const double instr = (double)ITER * 32 * warps_per_smsp;  // warp-instructions per SMSP
  printf("op %-28s warps/SMSP %d: %.3f cycles per warp-instruction per SMSP\n", name, warps_per_smsp, avg / instr);
}

template <int POLY, bool SUM, int DEG, bool CLAMP>
void run_sm(int warps_per_smsp) {
  const int threads = 128 * warps_per_smsp, iters = 64;
  softmax_kernel<POLY, SUM, DEG, CLAMP><<<148, threads>>>((uint32_t*)d_out, d_cyc, 0.37f, iters);
  softmax_kernel<POLY, SUM, DEG, CLAMP><<<148, threads>>>((uint32_t*)d_out, d_cyc, 0.37f, iters);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, d_cyc, sizeof(h), cudaMemcpyDeviceToHost);
  double avg = 0;
  for (int i = 0; i < 148; ++i) avg += h[i];
  avg /= 148;
  printf("softmax poly %d/8 sum %d deg %d clamp %d warps/SMSP %d: %.1f cycles per 128-score row-tile per SMSP-warp-slot, "
         "%.2f cycles per score per SMSP  (%s)\n", POLY, (int)SUM, DEG, (int)CLAMP, warps_per_smsp, avg / iters,
         avg / iters / 128.0 , cudaGetErrorString(e));
}

int main() {
  cudaMalloc(&d_out, 148 * 1024 * sizeof(float));
  cudaMalloc(&d_cyc, 148 * sizeof(long long));
  for (int w = 1; w <= 4; w *= 2) {
    run_op<0>("MUFU.EX2 f32", w);
    run_op<8>("ex2 bf16 (scalar)", w);
    run_op<9>("ex2 f16 (scalar)", w);
    run_op<1>("FFMA 3-reg", w);
    run_op<2>("FFMA2 (f32x2)", w);
    run_op<3>("FADD2 (f32x2)", w);
    run_op<4>("FMNMX", w);
    run_op<5>("FMNMX3", w);
    run_op<6>("shl+add (LEA?)", w);
    run_op<10>("IMAD imm", w);
    run_op<11>("IMNMX", w);
    run_op<7>("F2FP bf16x2 pack", w);
    run_op<12>("MUFU.EX2 + FADD (in range)", w);
    run_op<13>("MUFU.EX2 indep + FADD", w);
    run_op<14>("MUFU.EX2 + 1 FFMA2", w);
    run_op<15>("MUFU.EX2 + 3 FFMA2", w);
    run_op<16>("MUFU.EX2 indep + FADD + F2FP", w);
  }
  for (int w = 1; w <= 2; ++w) {
    run_sm<0, true, 3, true>(w);
    run_sm<0, false, 3, true>(w);
    run_sm<2, true, 3, true>(w);
    run_sm<2, false, 3, true>(w);
    run_sm<3, true, 3, true>(w);
    run_sm<3, false, 3, true>(w);
    run_sm<4, true, 3, true>(w);
    run_sm<4, false, 3, true>(w);
    run_sm<4, false, 3, false>(w);
    run_sm<4, false, 2, true>(w);
    run_sm<5, false, 3, true>(w);
    run_sm<5, false, 2, true>(w);
    run_sm<6, false, 3, true>(w);
    run_sm<8, false, 3, true>(w);
  }
  return 0;
}
