// Proof-of-Inference commitment kernels: SHA-256 leaf hashing, Merkle levels, sibling paths, batched
// leaf verification.  Replaces the hashlib/Python loops of
//   InferNet/neurons/miner.py:180-264      (Miner._build_merkle_tree)
//   InferNet/template/validator/proof.py:250-304 (build_merkle_root, verify_merkle_leaf)
// Results are bit-exact with hashlib: the message of a leaf is prefix(2 B, optional) ‖ A ‖ B and is
// hashed as ONE Merkle–Damgård chain (the construction cannot be changed without changing the root).
//
// Parallelism: a chain is serial (≈55 k dependent compressions for a 3.5 MB XL leaf), so the device
// parallelism is the number of leaves.
//   * chain kernel  : one 64-thread CTA per leaf. Warp 0 (producer) loads 32 blocks at a time with all
//                     lanes, byte-swaps and expands the 64-entry message schedule (independent per
//                     block) and stores W[t]+K[t] to shared memory; warp 1 (consumer) runs the 64
//                     dependent rounds per block reading the precomputed schedule. Double buffered.
//   * lane kernel   : one lane per leaf for thousands of leaves (all 32 lanes busy in the rounds).
#include <mutex>
#include <vector>

#include "common.cuh"

namespace b2v {
namespace {

__constant__ uint32_t K256[64] = {
    0x428a2f98u, 0x71374491u, 0xb5c0fbcfu, 0xe9b5dba5u, 0x3956c25bu, 0x59f111f1u, 0x923f82a4u, 0xab1c5ed5u,
    0xd807aa98u, 0x12835b01u, 0x243185beu, 0x550c7dc3u, 0x72be5d74u, 0x80deb1feu, 0x9bdc06a7u, 0xc19bf174u,
    0xe49b69c1u, 0xefbe4786u, 0x0fc19dc6u, 0x240ca1ccu, 0x2de92c6fu, 0x4a7484aau, 0x5cb0a9dcu, 0x76f988dau,
    0x983e5152u, 0xa831c66du, 0xb00327c8u, 0xbf597fc7u, 0xc6e00bf3u, 0xd5a79147u, 0x06ca6351u, 0x14292967u,
    0x27b70a85u, 0x2e1b2138u, 0x4d2c6dfcu, 0x53380d13u, 0x650a7354u, 0x766a0abbu, 0x81c2c92eu, 0x92722c85u,
    0xa2bfe8a1u, 0xa81a664bu, 0xc24b8b70u, 0xc76c51a3u, 0xd192e819u, 0xd6990624u, 0xf40e3585u, 0x106aa070u,
    0x19a4c116u, 0x1e376c08u, 0x2748774cu, 0x34b0bcb5u, 0x391c0cb3u, 0x4ed8aa4au, 0x5b9cca4fu, 0x682e6ff3u,
    0x748f82eeu, 0x78a5636fu, 0x84c87814u, 0x8cc70208u, 0x90befffau, 0xa4506cebu, 0xbef9a3f7u, 0xc67178f2u};

struct LeafDesc {
  const uint8_t* a;
  const uint8_t* b;
  uint64_t alen;
  uint64_t blen;
  uint32_t prefix_word;  // (p0 | p1<<8) << 16 : the two prefix bytes in the upper half (see msg_word)
  uint32_t prefix_len;   // 0 or 2
};

__device__ __forceinline__ uint32_t rotr(uint32_t x, int n) { return __funnelshift_r(x, x, n); }
__device__ __forceinline__ uint32_t bsw(uint32_t x) { return __byte_perm(x, 0, 0x0123); }
__device__ __forceinline__ uint32_t s0(uint32_t x) { return rotr(x, 7) ^ rotr(x, 18) ^ (x >> 3); }
__device__ __forceinline__ uint32_t s1(uint32_t x) { return rotr(x, 17) ^ rotr(x, 19) ^ (x >> 10); }
__device__ __forceinline__ uint32_t S0(uint32_t x) { return rotr(x, 2) ^ rotr(x, 13) ^ rotr(x, 22); }
__device__ __forceinline__ uint32_t S1(uint32_t x) { return rotr(x, 6) ^ rotr(x, 11) ^ rotr(x, 25); }
__device__ __forceinline__ uint32_t Ch(uint32_t e, uint32_t f, uint32_t g) { return (e & f) ^ (~e & g); }
__device__ __forceinline__ uint32_t Maj(uint32_t a, uint32_t b, uint32_t c) { return (a & b) ^ (a & c) ^ (b & c); }

__device__ __forceinline__ void sha_init(uint32_t* h) {
  h[0] = 0x6a09e667u; h[1] = 0xbb67ae85u; h[2] = 0x3c6ef372u; h[3] = 0xa54ff53au;
  h[4] = 0x510e527fu; h[5] = 0x9b05688cu; h[6] = 0x1f83d9abu; h[7] = 0x5be0cd19u;
}

// message geometry of one leaf
struct Msg {
  const uint8_t* a;
  const uint8_t* b;
  uint64_t alen, blen, L, nblocks;
  uint32_t prefix_word, prefix_len;
  bool fast;  // word-aligned buffers: message words can be assembled from 32-bit loads
};

__device__ __forceinline__ Msg make_msg(const LeafDesc& d) {
  Msg m;
  m.a = d.a; m.b = d.b; m.alen = d.alen; m.blen = d.blen;
  m.prefix_word = d.prefix_word; m.prefix_len = d.prefix_len;
  m.L = (uint64_t)d.prefix_len + d.alen + d.blen;
  m.nblocks = (m.L + 9 + 63) >> 6;
  m.fast = ((((uintptr_t)d.a) | ((uintptr_t)d.b) | d.alen | d.blen) & 3) == 0;
  return m;
}

__device__ __forceinline__ uint32_t msg_byte(const Msg& m, uint64_t o) {
  if (o < m.prefix_len) return (m.prefix_word >> (16 + 8 * (uint32_t)o)) & 0xffu;
  o -= m.prefix_len;
  if (o < m.alen) return m.a[o];
  o -= m.alen;
  return m.b[o];
}

// little-endian data word i of the stream A‖B (i = -1 is the prefix pseudo-word)
__device__ __forceinline__ uint32_t data_word(const Msg& m, int64_t i) {
  if (i < 0) return m.prefix_word;
  const uint64_t na = m.alen >> 2;
  if ((uint64_t)i < na) return __ldg(reinterpret_cast<const uint32_t*>(m.a) + i);
  return __ldg(reinterpret_cast<const uint32_t*>(m.b) + ((uint64_t)i - na));
}

// big-endian 32-bit word j of the padded message
__device__ __forceinline__ uint32_t msg_word(const Msg& m, uint64_t j) {
  const uint64_t o = j << 2;
  if (m.fast && o + 4 <= m.L) {
    if (m.prefix_len == 0) return bsw(data_word(m, (int64_t)j));
    // 2-byte prefix: message bytes 4j..4j+3 are data bytes 4j-2..4j+1
    const uint32_t lo = data_word(m, (int64_t)j - 1), hi = data_word(m, (int64_t)j);
    return bsw(__funnelshift_r(lo, hi, 16));
  }
  const uint64_t P = m.nblocks << 6;
  const uint64_t bitlen = m.L << 3;
  uint32_t w = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint64_t p = o + k;
    uint32_t byte;
    if (p < m.L) byte = msg_byte(m, p);
    else if (p == m.L) byte = 0x80u;
    else if (p >= P - 8) byte = (uint32_t)((bitlen >> (8 * (P - 1 - p))) & 0xffu);
    else byte = 0u;
    w = (w << 8) | byte;
  }
  return w;
}

#define SHA_ROUND(a, b, c, d, e, f, g, h, kw)                 \
  {                                                           \
    const uint32_t t1 = h + S1(e) + Ch(e, f, g) + (kw);       \
    const uint32_t t2 = S0(a) + Maj(a, b, c);                 \
    d += t1;                                                  \
    h = t1 + t2;                                              \
  }

// 64 rounds given the 64 precomputed (W[t]+K[t]) values in shared memory (16-byte aligned row).
__device__ __forceinline__ void compress_kw(uint32_t* st, const uint32_t* kw) {
  uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
  const uint4* kw4 = reinterpret_cast<const uint4*>(kw);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const uint4 x = kw4[2 * i], y = kw4[2 * i + 1];
    SHA_ROUND(a, b, c, d, e, f, g, h, x.x);
    SHA_ROUND(h, a, b, c, d, e, f, g, x.y);
    SHA_ROUND(g, h, a, b, c, d, e, f, x.z);
    SHA_ROUND(f, g, h, a, b, c, d, e, x.w);
    SHA_ROUND(e, f, g, h, a, b, c, d, y.x);
    SHA_ROUND(d, e, f, g, h, a, b, c, y.y);
    SHA_ROUND(c, d, e, f, g, h, a, b, y.z);
    SHA_ROUND(b, c, d, e, f, g, h, a, y.w);
  }
  st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}

// full compression of one block held in registers (rolling schedule)
__device__ __forceinline__ void compress_block(uint32_t* st, uint32_t* w) {
  uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
#pragma unroll
  for (int i = 0; i < 64; i += 8) {
    if (i >= 16) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int t = (i + k) & 15;
        w[t] = w[t] + s0(w[(t + 1) & 15]) + w[(t + 9) & 15] + s1(w[(t + 14) & 15]);
      }
    }
    SHA_ROUND(a, b, c, d, e, f, g, h, w[(i + 0) & 15] + K256[i + 0]);
    SHA_ROUND(h, a, b, c, d, e, f, g, w[(i + 1) & 15] + K256[i + 1]);
    SHA_ROUND(g, h, a, b, c, d, e, f, w[(i + 2) & 15] + K256[i + 2]);
    SHA_ROUND(f, g, h, a, b, c, d, e, w[(i + 3) & 15] + K256[i + 3]);
    SHA_ROUND(e, f, g, h, a, b, c, d, w[(i + 4) & 15] + K256[i + 4]);
    SHA_ROUND(d, e, f, g, h, a, b, c, w[(i + 5) & 15] + K256[i + 5]);
    SHA_ROUND(c, d, e, f, g, h, a, b, w[(i + 6) & 15] + K256[i + 6]);
    SHA_ROUND(b, c, d, e, f, g, h, a, w[(i + 7) & 15] + K256[i + 7]);
  }
  st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}

__device__ __forceinline__ void store_digest(uint8_t* out, const uint32_t* st) {
  uint32_t* o = reinterpret_cast<uint32_t*>(out);  // digests are 32-byte aligned rows
#pragma unroll
  for (int i = 0; i < 8; ++i) o[i] = bsw(st[i]);
}

// ---------------------------------------------------------------------------------------------
// chain kernel: one CTA (producer warp + consumer warp) per leaf
// ---------------------------------------------------------------------------------------------
constexpr int KW_STRIDE = 68;  // words per block row in smem: 16-byte aligned, 4-way conflicts at most
constexpr int BATCH = 32;      // blocks per producer pass (one per lane)

__global__ void __launch_bounds__(64) sha256_chain_kernel(const LeafDesc* __restrict__ descs,
                                                          uint8_t* __restrict__ digests) {
  __shared__ __align__(16) uint32_t kw[2][BATCH][KW_STRIDE];
  const Msg m = make_msg(descs[blockIdx.x]);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint64_t nbatches = (m.nblocks + BATCH - 1) / BATCH;

  auto produce = [&](uint64_t batch, int buf) {
    const uint64_t blk = batch * BATCH + lane;
    if (blk >= m.nblocks) return;
    uint32_t w[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) w[j] = msg_word(m, blk * 16 + j);
    uint32_t* row = &kw[buf][lane][0];
#pragma unroll
    for (int t = 0; t < 64; ++t) {
      if (t >= 16) {
        const int i = t & 15;
        w[i] = w[i] + s0(w[(i + 1) & 15]) + w[(i + 9) & 15] + s1(w[(i + 14) & 15]);
      }
      row[t] = w[t & 15] + K256[t];
    }
  };

  uint32_t st[8];
  sha_init(st);
  if (warp == 0) produce(0, 0);
  __syncthreads();
  for (uint64_t bt = 0; bt < nbatches; ++bt) {
    const int buf = (int)(bt & 1);
    if (warp == 0) {
      if (bt + 1 < nbatches) produce(bt + 1, buf ^ 1);
    } else {
      const uint64_t left = m.nblocks - bt * BATCH;
      const int cnt = left < (uint64_t)BATCH ? (int)left : BATCH;
      for (int i = 0; i < cnt; ++i) compress_kw(st, &kw[buf][i][0]);  // all lanes redundantly (broadcast LDS)
    }
    __syncthreads();
  }
  if (warp == 1 && lane == 0) store_digest(digests + (size_t)blockIdx.x * 32, st);
}

// ---------------------------------------------------------------------------------------------
// warp kernel: ONE warp per leaf, NO shared memory — the in-loop commitment (b2v.miner.record_trajectory) hashes one
// leaf on a side stream while the UNet runs. The UNet's GEMM kernels are persistent, one CTA per SM, and take the whole
// shared memory of the SM: a hashing CTA that owns even 17 KB of it (the chain kernel) evicts one GEMM CTA from its SM
// for the whole ~50 ms of the chain and stretches every GEMM of that window (measured +27 % per denoise step). This
// kernel needs 32 threads x ~110 registers and nothing else, so it sits BESIDE the resident GEMM / attention CTAs.
// Lane l expands the message schedule (W+K, 64 registers) of block 32*batch + l; the 64 dependent rounds of each block
// then run on all lanes redundantly with the schedule words broadcast by SHFL from the owning lane. The raw words of the
// next batch are loaded before the round loop so the global-memory latency hides under ~50 k cycles of rounds.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32) sha256_warp_kernel(const LeafDesc* __restrict__ descs,
                                                         uint8_t* __restrict__ digests) {
  const Msg m = make_msg(descs[blockIdx.x]);
  const int lane = threadIdx.x;
  const uint64_t nbatches = (m.nblocks + 31) >> 5;
  uint32_t st[8];
  sha_init(st);
  uint32_t wn[16];
  {
    const uint64_t blk = (uint64_t)lane;
#pragma unroll
    for (int j = 0; j < 16; ++j) wn[j] = blk < m.nblocks ? msg_word(m, blk * 16 + j) : 0u;
  }
  for (uint64_t bt = 0; bt < nbatches; ++bt) {
    uint32_t kw[64];
#pragma unroll
    for (int t = 0; t < 64; ++t) {
      if (t >= 16) {
        const int i = t & 15;
        wn[i] = wn[i] + s0(wn[(i + 1) & 15]) + wn[(i + 9) & 15] + s1(wn[(i + 14) & 15]);
      }
      kw[t] = wn[t & 15] + K256[t];
    }
    if (bt + 1 < nbatches) {
      const uint64_t blk = (bt + 1) * 32 + lane;
#pragma unroll
      for (int j = 0; j < 16; ++j) wn[j] = blk < m.nblocks ? msg_word(m, blk * 16 + j) : 0u;
    }
    const uint64_t left = m.nblocks - bt * 32;
    const int cnt = left < 32 ? (int)left : 32;
    for (int src = 0; src < cnt; ++src) {
      uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
#pragma unroll
      for (int t = 0; t < 64; t += 8) {
        SHA_ROUND(a, b, c, d, e, f, g, h, __shfl_sync(0xffffffffu, kw[t + 0], src));
        SHA_ROUND(h, a, b, c, d, e, f, g, __shfl_sync(0xffffffffu, kw[t + 1], src));
        SHA_ROUND(g, h, a, b, c, d, e, f, __shfl_sync(0xffffffffu, kw[t + 2], src));
        SHA_ROUND(f, g, h, a, b, c, d, e, __shfl_sync(0xffffffffu, kw[t + 3], src));
        SHA_ROUND(e, f, g, h, a, b, c, d, __shfl_sync(0xffffffffu, kw[t + 4], src));
        SHA_ROUND(d, e, f, g, h, a, b, c, __shfl_sync(0xffffffffu, kw[t + 5], src));
        SHA_ROUND(c, d, e, f, g, h, a, b, __shfl_sync(0xffffffffu, kw[t + 6], src));
        SHA_ROUND(b, c, d, e, f, g, h, a, __shfl_sync(0xffffffffu, kw[t + 7], src));
      }
      st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
    }
  }
  if (lane == 0) store_digest(digests + (size_t)blockIdx.x * 32, st);
}

// ---------------------------------------------------------------------------------------------
// lane kernel: one lane per leaf
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) sha256_lane_kernel(const LeafDesc* __restrict__ descs, int n,
                                                          uint8_t* __restrict__ digests) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const Msg m = make_msg(descs[i]);
  uint32_t st[8];
  sha_init(st);
  for (uint64_t blk = 0; blk < m.nblocks; ++blk) {
    uint32_t w[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) w[j] = msg_word(m, blk * 16 + j);
    compress_block(st, w);
  }
  store_digest(digests + (size_t)i * 32, st);
}

// ---------------------------------------------------------------------------------------------
// Merkle levels / paths / verification. A node is 32 bytes = 8 big-endian words once loaded with bsw.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void load_node(const uint8_t* p, uint32_t* w) {
  const uint32_t* q = reinterpret_cast<const uint32_t*>(p);
#pragma unroll
  for (int i = 0; i < 8; ++i) w[i] = bsw(q[i]);
}

// byte-wise lexicographic l < r  (proof.py:266 `if left < right`)
__device__ __forceinline__ bool node_less(const uint32_t* l, const uint32_t* r) {
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    if (l[i] != r[i]) return l[i] < r[i];
  }
  return false;
}

// SHA256(min(l,r) ‖ max(l,r)): 64 data bytes + one padding block
__device__ __forceinline__ void hash_pair(const uint32_t* l, const uint32_t* r, uint32_t* out) {
  uint32_t w[16];
  const bool lt = node_less(l, r);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    w[i] = lt ? l[i] : r[i];
    w[8 + i] = lt ? r[i] : l[i];
  }
  sha_init(out);
  compress_block(out, w);
  w[0] = 0x80000000u;
#pragma unroll
  for (int i = 1; i < 15; ++i) w[i] = 0;
  w[15] = 512u;
  compress_block(out, w);
}

__global__ void merkle_level_kernel(const uint8_t* __restrict__ cur, int ncur, uint8_t* __restrict__ nxt) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int nnext = (ncur + 1) >> 1;
  if (i >= nnext) return;
  uint32_t l[8], r[8], o[8];
  load_node(cur + (size_t)(2 * i) * 32, l);
  const int ri = (2 * i + 1 < ncur) ? 2 * i + 1 : 2 * i;  // odd tail pairs with itself (proof.py:264)
  load_node(cur + (size_t)ri * 32, r);
  hash_pair(l, r, o);
  store_digest(nxt + (size_t)i * 32, o);
}

// paths[leaf][level] = sibling (or the node itself when the sibling is missing) (proof.py:278-285)
__global__ void merkle_paths_kernel(const uint8_t* __restrict__ levels, int n, int depth,
                                    uint8_t* __restrict__ paths) {
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // one thread per (leaf, level, word)
  const int64_t total = (int64_t)n * depth * 8;
  if (gid >= total) return;
  const int word = (int)(gid & 7);
  const int64_t ll = gid >> 3;
  const int level = (int)(ll % depth);
  const int leaf = (int)(ll / depth);
  size_t off = 0;
  int len = n;
  for (int k = 0; k < level; ++k) { off += (size_t)len; len = (len + 1) >> 1; }
  const int idx = leaf >> level;
  const int sib = idx ^ 1;
  const int src = sib < len ? sib : idx;
  const uint32_t* s = reinterpret_cast<const uint32_t*>(levels + (off + (size_t)src) * 32);
  reinterpret_cast<uint32_t*>(paths)[gid] = s[word];
}

__global__ void merkle_verify_kernel(const uint8_t* __restrict__ leaf_hashes, const uint8_t* __restrict__ paths,
                                     int m, int depth, const uint8_t* __restrict__ root, int root_stride,
                                     uint8_t* __restrict__ ok) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  uint32_t cur[8], sib[8], o[8];
  load_node(leaf_hashes + (size_t)i * 32, cur);
  for (int k = 0; k < depth; ++k) {
    load_node(paths + ((size_t)i * depth + k) * 32, sib);
    hash_pair(cur, sib, o);
#pragma unroll
    for (int j = 0; j < 8; ++j) cur[j] = o[j];
  }
  load_node(root + (size_t)i * root_stride, sib);  // stride 0: one tree; 32: one root per leaf (many miners)
  bool eq = true;
#pragma unroll
  for (int j = 0; j < 8; ++j) eq = eq && (cur[j] == sib[j]);
  ok[i] = eq ? 1 : 0;
}

}  // namespace
}  // namespace b2v

using namespace b2v;

extern "C" int b2v_sha256_leaves(const uint16_t* prefix_be16_host, const void* const* a_ptrs_host,
                                 const uint64_t* a_lens_host, const void* const* b_ptrs_host,
                                 const uint64_t* b_lens_host, int n, uint8_t* digests_dev, int mode,
                                 void* stream) {
  B2V_CHECK(n >= 0, "n < 0");
  if (n == 0) return 0;
  B2V_CHECK(a_ptrs_host && a_lens_host && digests_dev, "null argument");
  B2V_TRY(b2v_device_check());
  cudaStream_t s = (cudaStream_t)stream;
  std::vector<LeafDesc> h((size_t)n);
  for (int i = 0; i < n; ++i) {
    LeafDesc& d = h[i];
    d.a = (const uint8_t*)a_ptrs_host[i];
    d.alen = a_lens_host[i];
    d.b = b_ptrs_host ? (const uint8_t*)b_ptrs_host[i] : nullptr;
    d.blen = (b_lens_host && d.b) ? b_lens_host[i] : 0;
    if (d.alen == 0) d.a = nullptr;
    if (prefix_be16_host) {
      const uint32_t t = prefix_be16_host[i];
      const uint32_t p0 = (t >> 8) & 0xffu, p1 = t & 0xffu;  // big-endian: high byte first
      d.prefix_word = (p0 | (p1 << 8)) << 16;
      d.prefix_len = 2;
    } else {
      d.prefix_word = 0;
      d.prefix_len = 0;
    }
  }
  LeafDesc* d_dev = nullptr;
  B2V_CUDA_OK(cudaMallocAsync((void**)&d_dev, sizeof(LeafDesc) * (size_t)n, s));
  B2V_CUDA_OK(cudaMemcpyAsync(d_dev, h.data(), sizeof(LeafDesc) * (size_t)n, cudaMemcpyHostToDevice, s));
  // pageable source: make sure the staging copy has consumed h before it goes out of scope
  if (mode == 0) mode = (n > 4 * sm_count() * 4) ? 2 : 1;
  if (mode == 1) {
    sha256_chain_kernel<<<n, 64, 0, s>>>(d_dev, digests_dev);
  } else if (mode == 3) {
    sha256_warp_kernel<<<n, 32, 0, s>>>(d_dev, digests_dev);
  } else {
    sha256_lane_kernel<<<(n + 127) / 128, 128, 0, s>>>(d_dev, n, digests_dev);
  }
  count_launch();
  cudaError_t e = launch_check();
  cudaFreeAsync(d_dev, s);
  if (e != cudaSuccess) return fail(B2V_ERR_CUDA, "sha256 launch: %s", cudaGetErrorString(e));
  // cudaMemcpyAsync from pageable memory returns only after the source has been staged, so `h` may die here.
  return 0;
}

extern "C" int b2v_merkle_depth(int n) {
  int d = 0;
  while (n > 1) { n = (n + 1) >> 1; ++d; }
  return d;
}

extern "C" size_t b2v_merkle_scratch_bytes(int n) {
  size_t tot = 0;
  int len = n;
  while (true) { tot += (size_t)len; if (len <= 1) break; len = (len + 1) >> 1; }
  return tot * 32 + 64;
}

extern "C" int b2v_merkle_build(const uint8_t* leaf_hashes_dev, int n, uint8_t* root_dev, uint8_t* paths_dev,
                                void* scratch_dev, void* stream) {
  B2V_CHECK(n >= 1, "merkle_build needs n >= 1 (the empty tree is handled by the host shim, proof.py:252)");
  B2V_CHECK(leaf_hashes_dev && root_dev && scratch_dev, "null argument");
  B2V_TRY(b2v_device_check());
  cudaStream_t s = (cudaStream_t)stream;
  uint8_t* levels = (uint8_t*)scratch_dev;
  B2V_CUDA_OK(cudaMemcpyAsync(levels, leaf_hashes_dev, (size_t)n * 32, cudaMemcpyDeviceToDevice, s));
  size_t off = 0;
  int len = n;
  while (len > 1) {
    const int nn = (len + 1) >> 1;
    merkle_level_kernel<<<(nn + 63) / 64, 64, 0, s>>>(levels + off * 32, len, levels + (off + (size_t)len) * 32);
    count_launch();
    off += (size_t)len;
    len = nn;
  }
  B2V_CUDA_OK(cudaMemcpyAsync(root_dev, levels + off * 32, 32, cudaMemcpyDeviceToDevice, s));
  const int depth = b2v_merkle_depth(n);
  if (paths_dev && depth > 0) {
    const int64_t total = (int64_t)n * depth * 8;
    merkle_paths_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(levels, n, depth, paths_dev);
    count_launch();
  }
  B2V_CUDA_OK(launch_check());
  return 0;
}

extern "C" int b2v_merkle_verify(const uint8_t* leaf_hashes_dev, const uint8_t* paths_dev, int m, int depth,
                                 const uint8_t* root_dev, uint8_t* ok_dev, void* stream) {
  B2V_CHECK(m >= 0 && depth >= 0, "bad sizes");
  if (m == 0) return 0;
  B2V_CHECK(leaf_hashes_dev && root_dev && ok_dev && (depth == 0 || paths_dev), "null argument");
  B2V_TRY(b2v_device_check());
  merkle_verify_kernel<<<(m + 63) / 64, 64, 0, (cudaStream_t)stream>>>(leaf_hashes_dev, paths_dev, m, depth,
                                                                     root_dev, 0, ok_dev);
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

extern "C" int b2v_merkle_verify_roots(const uint8_t* leaf_hashes_dev, const uint8_t* paths_dev, int m, int depth,
                                       const uint8_t* roots_dev, uint8_t* ok_dev, void* stream) {
  B2V_CHECK(m >= 0 && depth >= 0, "bad sizes");
  if (m == 0) return 0;
  B2V_CHECK(leaf_hashes_dev && roots_dev && ok_dev && (depth == 0 || paths_dev), "null argument");
  B2V_TRY(b2v_device_check());
  merkle_verify_kernel<<<(m + 63) / 64, 64, 0, (cudaStream_t)stream>>>(leaf_hashes_dev, paths_dev, m, depth,
                                                                     roots_dev, 32, ok_dev);
  count_launch();
  B2V_CUDA_OK(launch_check());
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// Merkle build for callers that must not allocate (b2v.h: b2v_merkle_build_static).
//
// The tree levels live in one file-scope device array instead of a caller-supplied scratch buffer. Every device that
// loads libb2v.so carries its own copy of the array (B2V_MERKLE_STATIC_BYTES, about 257 KiB) until the process exits.
// The levels are written in full, bottom-up, before anything reads them, so nothing has to be cleared between calls;
// what the previous call left behind is simply overwritten (and stays readable by nobody: the array is not exported).
// ---------------------------------------------------------------------------------------------------------------
#define B2V_MERKLE_STATIC_BYTES ((2 * (size_t)B2V_MERKLE_STATIC_MAX_LEAVES + 32) * 32 + 64)

static __device__ __align__(256) uint8_t g_merkle_static_levels[B2V_MERKLE_STATIC_BYTES];

namespace {
// `g_merkle_static_levels` in HOST code is only a handle for the runtime's symbol table, not a device address: it must go
// through cudaGetSymbolAddress, once per device and with that device current, before a kernel can be given it.
constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
uint8_t* g_static_levels_ptr[kMaxDevices] = {nullptr};

int merkle_static_levels(uint8_t** out) {
  int dev = -1;
  B2V_CUDA_OK(cudaGetDevice(&dev));
  B2V_CHECK(dev >= 0 && dev < kMaxDevices, "merkle_build_static: device ordinal %d outside [0,%d)", dev, kMaxDevices);
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_levels_ptr[dev]) {
    void* p = nullptr;
    B2V_CUDA_OK(cudaGetSymbolAddress(&p, g_merkle_static_levels));  // resolves on the CURRENT device (= dev)
    B2V_CHECK(p != nullptr, "merkle_build_static: cudaGetSymbolAddress returned a null address on device %d", dev);
    g_static_levels_ptr[dev] = (uint8_t*)p;
  }
  *out = g_static_levels_ptr[dev];
  return 0;
}
}  // namespace

extern "C" int b2v_merkle_static_max_leaves(void) { return B2V_MERKLE_STATIC_MAX_LEAVES; }

extern "C" int b2v_merkle_build_static(const uint8_t* leaf_hashes_dev, int n, uint8_t* root_dev, uint8_t* paths_dev,
                                       void* stream) {
  B2V_CHECK(n >= 1, "merkle_build_static needs n >= 1 (the empty tree is handled by the host shim, proof.py:252)");
  B2V_CHECK(n <= B2V_MERKLE_STATIC_MAX_LEAVES,
            "merkle_build_static: %d leaves exceed the built-in scratch (B2V_MERKLE_STATIC_MAX_LEAVES = %d); "
            "use b2v_merkle_build with a caller-provided scratch buffer for larger trees",
            n, B2V_MERKLE_STATIC_MAX_LEAVES);
  B2V_CHECK(b2v_merkle_scratch_bytes(n) <= B2V_MERKLE_STATIC_BYTES, "merkle_build_static: scratch bound violated");
  B2V_CHECK(leaf_hashes_dev && root_dev, "null argument");
  B2V_TRY(b2v_device_check());
  uint8_t* levels = nullptr;
  B2V_TRY(merkle_static_levels(&levels));  // first call on a device: symbol lookup; afterwards a table read
  cudaStream_t s = (cudaStream_t)stream;
  // Same launches as b2v_merkle_build from here on, against the built-in array. No cudaMalloc / cudaFree / pool use.
  B2V_CUDA_OK(cudaMemcpyAsync(levels, leaf_hashes_dev, (size_t)n * 32, cudaMemcpyDeviceToDevice, s));
  size_t off = 0;
  int len = n;
  while (len > 1) {
    const int nn = (len + 1) >> 1;
    merkle_level_kernel<<<(nn + 63) / 64, 64, 0, s>>>(levels + off * 32, len, levels + (off + (size_t)len) * 32);
    count_launch();
    off += (size_t)len;
    len = nn;
  }
  B2V_CUDA_OK(cudaMemcpyAsync(root_dev, levels + off * 32, 32, cudaMemcpyDeviceToDevice, s));
  const int depth = b2v_merkle_depth(n);
  if (paths_dev && depth > 0) {
    const int64_t total = (int64_t)n * depth * 8;
    merkle_paths_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(levels, n, depth, paths_dev);
    count_launch();
  }
  B2V_CUDA_OK(launch_check());
  return 0;
}
