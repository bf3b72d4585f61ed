// SHA-256 compression-round micro-benchmark (sm_100a): cycles per 64-byte block of ONE dependent chain on one warp, for
// different ways of spreading the round's integer work over the alu pipe (SHF / LOP3 / IADD3) and the fma pipe (IMAD).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_build/ubench_sha_round tools/ubench_sha_round.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ uint32_t rotr(uint32_t x, int n) { return __funnelshift_r(x, x, n); }
__device__ __forceinline__ uint32_t S0(uint32_t x) { return rotr(x, 2) ^ rotr(x, 13) ^ rotr(x, 22); }
__device__ __forceinline__ uint32_t S1(uint32_t x) { return rotr(x, 6) ^ rotr(x, 11) ^ rotr(x, 25); }
__device__ __forceinline__ uint32_t Ch(uint32_t e, uint32_t f, uint32_t g) { return (e & f) ^ (~e & g); }
__device__ __forceinline__ uint32_t Maj(uint32_t a, uint32_t b, uint32_t c) { return (a & b) ^ (a & c) ^ (b & c); }
// a + b on the FMA pipe: the multiplier is a kernel argument (always 1), so ptxas cannot turn the IMAD back into IADD3
__device__ __forceinline__ uint32_t addf(uint32_t a, uint32_t b, uint32_t one) {
  uint32_t r;
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(one), "r"(b));
  return r;
}

template <int V>
__device__ __forceinline__ void round_fn(uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d, uint32_t& e, uint32_t& f,
                                         uint32_t& g, uint32_t& h, uint32_t kw, uint32_t one) {
  if (V == 0) {  // as written in csrc/sha256.cu
    const uint32_t t1 = h + S1(e) + Ch(e, f, g) + kw;
    const uint32_t t2 = S0(a) + Maj(a, b, c);
    d += t1;
    h = t1 + t2;
  } else if (V == 1) {  // every add on the FMA pipe
    const uint32_t t1 = addf(addf(addf(h, kw, one), Ch(e, f, g), one), S1(e), one);
    const uint32_t t2 = addf(S0(a), Maj(a, b, c), one);
    d = addf(d, t1, one);
    h = addf(t1, t2, one);
  } else {  // FMA-pipe adds, and everything that does not depend on Sigma1(e) folded in ahead of it
    const uint32_t pre = addf(addf(h, kw, one), Ch(e, f, g), one);   // h + kw + Ch
    const uint32_t u = addf(d, pre, one);                            // d + h + kw + Ch
    const uint32_t w = addf(addf(pre, Maj(a, b, c), one), S0(a), one);
    const uint32_t s1 = S1(e);
    d = addf(u, s1, one);   // new e: e -> SHF -> LOP3 -> IMAD
    h = addf(w, s1, one);   // new a
  }
}

template <int V>
__global__ void __launch_bounds__(32) chain_kernel(uint32_t* out, long long* cyc, int nblk, uint32_t one) {
  __shared__ __align__(16) uint32_t kw[64];
  for (int i = threadIdx.x; i < 64; i += 32) kw[i] = 0x9e3779b9u * (i + 1);
  __syncthreads();
  uint32_t a = 0x6a09e667u, b = 0xbb67ae85u, c = 0x3c6ef372u, d = 0xa54ff53au, e = 0x510e527fu, f = 0x9b05688cu,
           g = 0x1f83d9abu, h = 0x5be0cd19u;
  const long long t0 = clock64();
  for (int blk = 0; blk < nblk; ++blk) {
    const uint4* kw4 = reinterpret_cast<const uint4*>(kw);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const uint4 x = kw4[2 * i], y = kw4[2 * i + 1];
      round_fn<V>(a, b, c, d, e, f, g, h, x.x, one); round_fn<V>(h, a, b, c, d, e, f, g, x.y, one);
      round_fn<V>(g, h, a, b, c, d, e, f, x.z, one); round_fn<V>(f, g, h, a, b, c, d, e, x.w, one);
      round_fn<V>(e, f, g, h, a, b, c, d, y.x, one); round_fn<V>(d, e, f, g, h, a, b, c, y.y, one);
      round_fn<V>(c, d, e, f, g, h, a, b, y.z, one); round_fn<V>(b, c, d, e, f, g, h, a, y.w, one);
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) {
    out[blockIdx.x * 8 + 0] = a ^ b ^ c ^ d; out[blockIdx.x * 8 + 1] = e ^ f ^ g ^ h;
    cyc[blockIdx.x] = t1 - t0;
  }
}

template <int V>
void run(uint32_t* o, long long* c, uint32_t* ref) {
  const int nblk = 20000;
  chain_kernel<V><<<50, 32>>>(o, c, nblk, 1u);
  cudaDeviceSynchronize();
  long long h[50];
  uint32_t ho[400];
  cudaMemcpy(h, c, sizeof(h), cudaMemcpyDeviceToHost);
  cudaMemcpy(ho, o, sizeof(ho), cudaMemcpyDeviceToHost);
  if (V == 0) { ref[0] = ho[0]; ref[1] = ho[1]; }
  double avg = 0;
  for (int i = 0; i < 50; ++i) avg += h[i];
  avg /= 50;
  printf("variant %d: %.1f cycles per 64-byte block, %.2f per round (%s)\n", V, avg / nblk, avg / nblk / 64,
         (ho[0] == ref[0] && ho[1] == ref[1]) ? "same digest state as variant 0" : "MISMATCH");
}

int main() {
  uint32_t* o; long long* c; uint32_t ref[2];
  cudaMalloc(&o, 400 * 4); cudaMalloc(&c, 50 * 8);
  run<0>(o, c, ref); run<1>(o, c, ref); run<2>(o, c, ref);
  return 0;
}
