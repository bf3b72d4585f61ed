// TMEM read-bandwidth micro-benchmark (sm_100a): back-to-back tcgen05.ld from an allocated (uninitialised) TMEM region,
// 1 / 2 / 4 warps of one CTA (each warp its own 32-lane quarter) and 1 or 2 CTAs per SM. Prints bytes per clock per SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_build/ubench_tmem tools/ubench_tmem.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16x256(uint32_t taddr, uint32_t* v) {  // 16x256b.x8: 32 regs, lanes 0-15 x 64 columns
  asm volatile(
      "tcgen05.ld.sync.aligned.16x256b.x8.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}

template <int SHAPE, int DEPTH>  // DEPTH = loads in flight before each wait::ld
__global__ void __launch_bounds__(128) tmem_kernel(uint32_t* out, long long* cyc, int iters, int active_warps) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)), "r"(256) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = slot + ((uint32_t)(warp * 32) << 16);
  uint32_t acc = 0;
  long long t0 = 0, t1 = 0;
  if (warp < active_warps) {
    t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      uint32_t v[DEPTH][32];
#pragma unroll
      for (int d = 0; d < DEPTH; ++d) {
        if (SHAPE == 0) tmem_ld32(base + 32 * (d & 3), v[d]);
        else tmem_ld16x256(base + 64 * (d & 1), v[d]);
      }
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int d = 0; d < DEPTH; ++d) acc ^= v[d][0] ^ v[d][13] ^ v[d][31];
    }
    t1 = clock64();
  }
  __syncthreads();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(slot), "r"(256) : "memory");
}

template <int SHAPE, int DEPTH>
void run(uint32_t* d_out, long long* d_cyc, int ctas_per_sm, int warps) {
  const int iters = 2000, grid = 148 * ctas_per_sm;
  tmem_kernel<SHAPE, DEPTH><<<grid, 128>>>(d_out, d_cyc, iters, warps);
  tmem_kernel<SHAPE, DEPTH><<<grid, 128>>>(d_out, d_cyc, iters, warps);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[296];
  cudaMemcpy(h, d_cyc, grid * sizeof(long long), cudaMemcpyDeviceToHost);
  double avg = 0;
  for (int i = 0; i < grid; ++i) avg += h[i];
  avg /= grid;
  const double bytes_per_warp = (double)iters * DEPTH * 4096.0;
  printf("tcgen05.ld %s depth %d, %d CTA/SM x %d warps: %.1f cycles per 4 KiB load per warp, %.1f B/clk per warp, %.1f B/clk per SM (%s)\n",
         SHAPE == 0 ? "32x32b.x32" : "16x256b.x8", DEPTH, ctas_per_sm, warps, avg / (iters * DEPTH), bytes_per_warp / avg,
         bytes_per_warp * warps * ctas_per_sm / avg, cudaGetErrorString(e));
}

int main() {
  uint32_t* d_out;
  long long* d_cyc;
  cudaMalloc(&d_out, 296 * 128 * 4);
  cudaMalloc(&d_cyc, 296 * sizeof(long long));
  for (int c = 1; c <= 2; ++c)
    for (int w = 1; w <= 4; w *= 2) {
      run<0, 1>(d_out, d_cyc, c, w);
      run<0, 4>(d_out, d_cyc, c, w);
      run<1, 2>(d_out, d_cyc, c, w);
    }
  return 0;
}
