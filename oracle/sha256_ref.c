/* ORACLE — test infrastructure only (tests/, __graft_entry__.smoke(), bench.py's CPU baseline).
 *
 * Plain-C restatement of the byte arithmetic behind the reference's commitment: FIPS 180-4 SHA-256 (what
 * hashlib.sha256 computes at InferNet/neurons/miner.py:208 and template/validator/proof.py:256,270,300) and the
 * sorted-pair / duplicate-last Merkle levels of template/validator/proof.py:260-272. Scalar, one thread: it is the
 * "port" CPU baseline and a second, hashlib-independent check of the CUDA kernels.
 * Pinned by tests/test_oracle_merkle.py against the FIPS known-answer vectors ("abc", the 448-bit message, one
 * million 'a') and against hashlib / tests/golden/proof_golden.json.
 *
 * build: gcc -O3 -shared -fPIC -o oracle/libsha256_ref.so oracle/sha256_ref.c   (done by __graft_entry__.build())
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

static const uint32_t K[64] = {
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5, 0xd807aa98,
    0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174, 0xe49b69c1, 0xefbe4786,
    0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da, 0x983e5152, 0xa831c66d, 0xb00327c8,
    0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967, 0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13,
    0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85, 0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819,
    0xd6990624, 0xf40e3585, 0x106aa070, 0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a,
    0x5b9cca4f, 0x682e6ff3, 0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7,
    0xc67178f2};

typedef struct {
  uint32_t h[8];
  uint8_t buf[64];
  size_t fill;
  uint64_t total;
} sha_ctx;

static uint32_t ror(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }

static void block(uint32_t* h, const uint8_t* p) {
  uint32_t w[64], a, b, c, d, e, f, g, hh;
  for (int i = 0; i < 16; ++i)
    w[i] = ((uint32_t)p[4 * i] << 24) | ((uint32_t)p[4 * i + 1] << 16) | ((uint32_t)p[4 * i + 2] << 8) | p[4 * i + 3];
  for (int i = 16; i < 64; ++i) {
    uint32_t s0 = ror(w[i - 15], 7) ^ ror(w[i - 15], 18) ^ (w[i - 15] >> 3);
    uint32_t s1 = ror(w[i - 2], 17) ^ ror(w[i - 2], 19) ^ (w[i - 2] >> 10);
    w[i] = w[i - 16] + s0 + w[i - 7] + s1;
  }
  a = h[0]; b = h[1]; c = h[2]; d = h[3]; e = h[4]; f = h[5]; g = h[6]; hh = h[7];
  for (int i = 0; i < 64; ++i) {
    uint32_t t1 = hh + (ror(e, 6) ^ ror(e, 11) ^ ror(e, 25)) + ((e & f) ^ (~e & g)) + K[i] + w[i];
    uint32_t t2 = (ror(a, 2) ^ ror(a, 13) ^ ror(a, 22)) + ((a & b) ^ (a & c) ^ (b & c));
    hh = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
  }
  h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
}

static void sha_init(sha_ctx* c) {
  static const uint32_t iv[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a,
                                 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};
  memcpy(c->h, iv, sizeof iv);
  c->fill = 0;
  c->total = 0;
}

static void sha_update(sha_ctx* c, const uint8_t* p, size_t n) {
  c->total += n;
  if (c->fill) {
    size_t take = 64 - c->fill < n ? 64 - c->fill : n;
    memcpy(c->buf + c->fill, p, take);
    c->fill += take; p += take; n -= take;
    if (c->fill == 64) { block(c->h, c->buf); c->fill = 0; }
  }
  while (n >= 64) { block(c->h, p); p += 64; n -= 64; }
  if (n) { memcpy(c->buf, p, n); c->fill = n; }
}

static void sha_final(sha_ctx* c, uint8_t out[32]) {
  uint64_t bits = c->total * 8;
  uint8_t pad[72] = {0x80};
  size_t padlen = (c->fill < 56) ? 56 - c->fill : 120 - c->fill;
  uint8_t len[8];
  for (int i = 0; i < 8; ++i) len[i] = (uint8_t)(bits >> (56 - 8 * i));
  sha_update(c, pad, padlen);
  sha_update(c, len, 8);
  for (int i = 0; i < 8; ++i) {
    out[4 * i] = (uint8_t)(c->h[i] >> 24); out[4 * i + 1] = (uint8_t)(c->h[i] >> 16);
    out[4 * i + 2] = (uint8_t)(c->h[i] >> 8); out[4 * i + 3] = (uint8_t)c->h[i];
  }
}

/* digest = SHA256(prefix[0:prefix_len] || a || b) */
void ref_sha256_concat(const uint8_t* prefix, size_t prefix_len, const uint8_t* a, size_t alen, const uint8_t* b,
                       size_t blen, uint8_t out[32]) {
  sha_ctx c;
  sha_init(&c);
  if (prefix_len) sha_update(&c, prefix, prefix_len);
  if (alen) sha_update(&c, a, alen);
  if (blen) sha_update(&c, b, blen);
  sha_final(&c, out);
}

/* leaf i = SHA256(be16(t_i) || z_i || eps_i), z/eps rows of `stride` bytes inside two big buffers */
void ref_leaf_hashes(const uint16_t* t, const uint8_t* z, const uint8_t* eps, size_t bytes_each, int n,
                     uint8_t* out) {
  for (int i = 0; i < n; ++i) {
    uint8_t pre[2] = {(uint8_t)(t[i] >> 8), (uint8_t)t[i]};
    ref_sha256_concat(pre, 2, z + (size_t)i * bytes_each, bytes_each, eps + (size_t)i * bytes_each, bytes_each,
                      out + 32 * (size_t)i);
  }
}

static void parent(const uint8_t* l, const uint8_t* r, uint8_t* out) {
  uint8_t cat[64];
  if (memcmp(l, r, 32) < 0) { memcpy(cat, l, 32); memcpy(cat + 32, r, 32); }
  else { memcpy(cat, r, 32); memcpy(cat + 32, l, 32); }
  ref_sha256_concat(0, 0, cat, 64, 0, 0, out);
}

/* levels: caller buffer holding every level back to back (n + ceil(n/2) + ... + 1 nodes); returns node count. */
int ref_merkle_levels(const uint8_t* leaf_hashes, int n, uint8_t* levels) {
  memcpy(levels, leaf_hashes, (size_t)n * 32);
  int off = 0, len = n, total = n;
  while (len > 1) {
    int nn = (len + 1) / 2;
    for (int i = 0; i < nn; ++i) {
      const uint8_t* l = levels + (size_t)(off + 2 * i) * 32;
      const uint8_t* r = (2 * i + 1 < len) ? l + 32 : l;
      parent(l, r, levels + (size_t)(off + len + i) * 32);
    }
    off += len; len = nn; total += nn;
  }
  return total;
}
