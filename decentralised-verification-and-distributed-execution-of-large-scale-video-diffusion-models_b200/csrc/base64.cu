// Wire format of the commitment: base64 of every step's leaf images
// (InferNet/neurons/miner.py:643-657 `base64.b64encode(z_bytes).decode('ascii')` for every z / eps of the trajectory,
// twice — proof.leaf_data and synapse.latents / noise_preds — and again in the spot-check reveal, miner.py:417-423).
// The reference pulls 2 x 88 MB to the host and encodes them with the stdlib; here the rows of the device-resident
// trajectory buffers are encoded in place (RFC 4648 standard alphabet, '=' padding: bit-identical to base64.b64encode)
// and leave the device once, already as ASCII. Pure byte work, HBM-bound: 3 bytes read, 4 written.
//
// One thread = 12 input bytes -> 16 output characters (three 32-bit loads, one 16-byte store; consecutive threads touch
// consecutive addresses). The last (< 12 byte) group of a row takes the scalar path with padding.
#include "common.cuh"

namespace b2v {
namespace {

// 6-bit value -> ASCII, branch-free: 'A'..'Z' 'a'..'z' '0'..'9' '+' '/'
__device__ __forceinline__ uint32_t b64_char(uint32_t v) {
  uint32_t c = v + 65u;                 // 'A' + v
  c += (v >= 26u) ? 6u : 0u;            // 'a' - 'Z' - 1
  c -= (v >= 52u) ? 75u : 0u;           // '0' - 'z' - 1
  c -= (v >= 62u) ? 15u : 0u;           // '+' - '9' - 1
  c += (v >= 63u) ? 3u : 0u;            // '/' - '+' - 1
  return c;
}

// 3 bytes b0 b1 b2 (b0 first in memory) -> 4 characters packed little-endian in one word
__device__ __forceinline__ uint32_t b64_quad(uint32_t b0, uint32_t b1, uint32_t b2) {
  const uint32_t t = (b0 << 16) | (b1 << 8) | b2;
  return b64_char(t >> 18) | (b64_char((t >> 12) & 63u) << 8) | (b64_char((t >> 6) & 63u) << 16) |
         (b64_char(t & 63u) << 24);
}

__global__ void __launch_bounds__(256)
base64_rows_kernel(const uint8_t* __restrict__ src, int64_t n_rows, uint64_t row_bytes, uint64_t row_stride,
                   uint8_t* __restrict__ dst, uint64_t out_row_bytes, int aligned) {
  const uint64_t groups = (row_bytes + 11) / 12;  // 12-byte groups per row (the last may be partial)
  const uint64_t total = (uint64_t)n_rows * groups;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t row = i / groups, gidx = i - row * groups;
    const uint8_t* s = src + row * row_stride + gidx * 12;
    uint8_t* d = dst + row * out_row_bytes + gidx * 16;
    const uint64_t left = row_bytes - gidx * 12;
    if (left >= 12 && aligned) {
      const uint32_t* s4 = reinterpret_cast<const uint32_t*>(s);
      const uint32_t w0 = __ldg(s4), w1 = __ldg(s4 + 1), w2 = __ldg(s4 + 2);
      uint4 o;
      o.x = b64_quad(w0 & 255u, (w0 >> 8) & 255u, (w0 >> 16) & 255u);
      o.y = b64_quad(w0 >> 24, w1 & 255u, (w1 >> 8) & 255u);
      o.z = b64_quad((w1 >> 16) & 255u, w1 >> 24, w2 & 255u);
      o.w = b64_quad((w2 >> 8) & 255u, (w2 >> 16) & 255u, w2 >> 24);
      *reinterpret_cast<uint4*>(d) = o;
    } else {
      const int n = left < 12 ? (int)left : 12;
      for (int k = 0; k < n; k += 3) {
        const int m = n - k;  // bytes in this triple: 1, 2 or >= 3
        const uint32_t b0 = s[k], b1 = m > 1 ? s[k + 1] : 0u, b2 = m > 2 ? s[k + 2] : 0u;
        uint32_t q = b64_quad(b0, b1, b2);
        if (m == 1) q = (q & 0x0000ffffu) | 0x3d3d0000u;       // "xx=="
        else if (m == 2) q = (q & 0x00ffffffu) | 0x3d000000u;  // "xxx="
        uint8_t* o = d + (k / 3) * 4;
        o[0] = (uint8_t)q; o[1] = (uint8_t)(q >> 8); o[2] = (uint8_t)(q >> 16); o[3] = (uint8_t)(q >> 24);
      }
    }
  }
}

// ---- decode (the validator's side of the wire: neurons/validator.py:783-786 `base64.b64decode(z_b)`) ----------------
// ASCII -> 6-bit value, branch-free; 64 = '=' (padding), 255 = not in the alphabet
__device__ __forceinline__ uint32_t b64_val(uint32_t c) {
  uint32_t v = 255u;
  v = (c - 65u < 26u) ? c - 65u : v;   // 'A'..'Z'
  v = (c - 97u < 26u) ? c - 71u : v;   // 'a'..'z'
  v = (c - 48u < 10u) ? c + 4u : v;    // '0'..'9'
  v = (c == 43u) ? 62u : v;            // '+'
  v = (c == 47u) ? 63u : v;            // '/'
  v = (c == 61u) ? 64u : v;            // '='
  return v;
}

// One thread = 16 characters -> 12 bytes. Every row has row_chars characters (a multiple of 4). status[row] gets
// (pad count 0..2) | 0x100 when the row is not canonical base64 (a character outside the alphabet, or '=' anywhere but
// in the last one or two positions): the caller then decodes that row with the reference's own permissive routine.
// Trailing non-zero bits in a padded quad are ignored, as binascii does. Padded byte positions are written as zero.
__global__ void __launch_bounds__(256)
base64_decode_rows_kernel(const uint8_t* __restrict__ src, int64_t n_rows, uint64_t row_chars, uint64_t row_stride,
                          uint8_t* __restrict__ dst, uint64_t dst_stride, int* __restrict__ status, int aligned) {
  const uint64_t groups = (row_chars + 15) / 16;
  const uint64_t total = (uint64_t)n_rows * groups;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t row = i / groups, gidx = i - row * groups;
    const uint8_t* s = src + row * row_stride + gidx * 16;
    uint8_t* d = dst + row * dst_stride + gidx * 12;
    const uint64_t left = row_chars - gidx * 16;
    const int n = left < 16 ? (int)left : 16;  // characters in this group (multiple of 4)
    uint32_t w[4] = {0x41414141u, 0x41414141u, 0x41414141u, 0x41414141u};  // "AAAA": decodes to zero
    if (n == 16 && aligned) {
      const uint4 u = __ldg(reinterpret_cast<const uint4*>(s));
      w[0] = u.x; w[1] = u.y; w[2] = u.z; w[3] = u.w;
    } else {
      for (int k = 0; k < n; ++k) w[k >> 2] = (w[k >> 2] & ~(255u << (8 * (k & 3)))) | ((uint32_t)s[k] << (8 * (k & 3)));
    }
    uint32_t bad = 0, pad = 0;
    uint8_t o[12];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint32_t v[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) v[k] = b64_val((w[q] >> (8 * k)) & 255u);
      const bool last = (uint64_t)(q * 4 + 4) == left;  // the final quad of the row: '=' allowed in positions 2, 3
      uint32_t p = 0;
      if (last && v[3] == 64u) { p = 1; v[3] = 0; if (v[2] == 64u) { p = 2; v[2] = 0; } }
      pad |= p;
      bad |= (v[0] | v[1] | v[2] | v[3]) >> 6;  // 64 ('=' elsewhere) or 255 (outside the alphabet)
      const uint32_t t = ((v[0] & 63u) << 18) | ((v[1] & 63u) << 12) | ((v[2] & 63u) << 6) | (v[3] & 63u);
      o[3 * q] = (uint8_t)(t >> 16); o[3 * q + 1] = (uint8_t)(t >> 8); o[3 * q + 2] = (uint8_t)t;
    }
    const int nb = (n >> 2) * 3;
    if (nb == 12 && aligned) {
      uint32_t* d4 = reinterpret_cast<uint32_t*>(d);
#pragma unroll
      for (int k = 0; k < 3; ++k)
        d4[k] = (uint32_t)o[4 * k] | ((uint32_t)o[4 * k + 1] << 8) | ((uint32_t)o[4 * k + 2] << 16) | ((uint32_t)o[4 * k + 3] << 24);
    } else {
#pragma unroll
      for (int k = 0; k < 12; ++k)
        if (k < nb) d[k] = o[k];
    }
    if (bad) atomicOr(status + row, 0x100);
    if (pad) atomicOr(status + row, (int)pad);
  }
}

}  // namespace
}  // namespace b2v

using namespace b2v;

extern "C" int b2v_base64_decode_rows(const void* src_dev, int64_t n_rows, uint64_t row_chars, uint64_t row_stride,
                                      void* dst_dev, uint64_t dst_stride, int* status_dev, void* stream) {
  B2V_CHECK(n_rows >= 0, "n_rows < 0");
  if (n_rows == 0) return 0;
  B2V_CHECK(status_dev, "null status");
  B2V_CHECK(row_chars % 4 == 0, "base64 rows must hold a multiple of 4 characters (got %llu)", (unsigned long long)row_chars);
  B2V_TRY(b2v_device_check());
  cudaStream_t s = (cudaStream_t)stream;
  B2V_CUDA_OK(cudaMemsetAsync(status_dev, 0, sizeof(int) * (size_t)n_rows, s));
  if (row_chars == 0) return 0;
  B2V_CHECK(src_dev && dst_dev, "null argument");
  B2V_CHECK(row_stride >= row_chars && dst_stride >= row_chars / 4 * 3, "row strides shorter than the rows");
  const int aligned = (((uintptr_t)src_dev | row_stride) & 15) == 0 && (((uintptr_t)dst_dev | dst_stride) & 3) == 0;
  const uint64_t total = (uint64_t)n_rows * ((row_chars + 15) / 16);
  uint64_t blocks = (total + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  base64_decode_rows_kernel<<<(unsigned)blocks, 256, 0, s>>>((const uint8_t*)src_dev, n_rows, row_chars, row_stride,
                                                             (uint8_t*)dst_dev, dst_stride, status_dev, aligned);
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

extern "C" uint64_t b2v_base64_len(uint64_t nbytes) { return 4 * ((nbytes + 2) / 3); }

extern "C" int b2v_base64_rows(const void* src_dev, int64_t n_rows, uint64_t row_bytes, uint64_t row_stride,
                               void* dst_dev, void* stream) {
  B2V_CHECK(n_rows >= 0, "n_rows < 0");
  if (n_rows == 0 || row_bytes == 0) return 0;
  B2V_CHECK(src_dev && dst_dev, "null argument");
  B2V_CHECK(row_stride >= row_bytes, "row_stride < row_bytes");
  B2V_TRY(b2v_device_check());
  const uint64_t out_row = b2v_base64_len(row_bytes);
  // the vector path needs 4-byte aligned sources and 16-byte aligned destinations for every full group of every row
  const int aligned = (((uintptr_t)src_dev | row_stride) & 3) == 0 && (((uintptr_t)dst_dev | out_row) & 15) == 0;
  const uint64_t total = (uint64_t)n_rows * ((row_bytes + 11) / 12);
  uint64_t blocks = (total + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  base64_rows_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const uint8_t*)src_dev, n_rows, row_bytes,
                                                                         row_stride, (uint8_t*)dst_dev, out_row, aligned);
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}
