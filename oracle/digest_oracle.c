/*
 * digest_oracle.c — CPU ORACLE (test infrastructure, not product; see digest_oracle.h).
 *
 * Plain-C restatement of the three digests the reference's dependencies compute through
 * Node crypto (OpenSSL):
 *   SHA-1   : FIPS 180-4 §6.1  — webtorrent piece verify, reference call site lib/download.js:64-66
 *   MD5     : RFC 1321 §3      — minio-js per-part Content-MD5 / ETag, lib/upload.js:45,55
 *   SHA-256 : FIPS 180-4 §6.2  — minio-js x-amz-content-sha256,       lib/upload.js:45,55
 * plus the S3 framing minio-js applies to them and the synthetic-data generator of SURVEY §8d.
 * PARITY UNPINNED by the reference's own tests; pinned by tests/golden/kat.json and libcrypto.
 */
#include "digest_oracle.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>

static inline uint32_t rol(uint32_t x, int n) { return (x << n) | (x >> (32 - n)); }
static inline uint32_t ror(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }

static inline uint32_t be32(const uint8_t* p) {
  return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3];
}
static inline uint32_t le32(const uint8_t* p) {
  return ((uint32_t)p[3] << 24) | ((uint32_t)p[2] << 16) | ((uint32_t)p[1] << 8) | p[0];
}
static inline void put_be32(uint8_t* p, uint32_t v) {
  p[0] = (uint8_t)(v >> 24); p[1] = (uint8_t)(v >> 16); p[2] = (uint8_t)(v >> 8); p[3] = (uint8_t)v;
}
static inline void put_le32(uint8_t* p, uint32_t v) {
  p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24);
}

/* ---------------------------------------------------------------- SHA-1 (FIPS 180-4 §6.1.2) */
static void sha1_block(uint32_t h[5], const uint8_t* p) {
  uint32_t w[80];
  for (int t = 0; t < 16; t++) w[t] = be32(p + 4 * t);
  for (int t = 16; t < 80; t++) w[t] = rol(w[t - 3] ^ w[t - 8] ^ w[t - 14] ^ w[t - 16], 1);
  uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4];
  for (int t = 0; t < 80; t++) {
    uint32_t f, k;
    if (t < 20)      { f = (b & c) | (~b & d);          k = 0x5a827999u; }
    else if (t < 40) { f = b ^ c ^ d;                   k = 0x6ed9eba1u; }
    else if (t < 60) { f = (b & c) | (b & d) | (c & d); k = 0x8f1bbcdcu; }
    else             { f = b ^ c ^ d;                   k = 0xca62c1d6u; }
    uint32_t tmp = rol(a, 5) + f + e + k + w[t];
    e = d; d = c; c = rol(b, 30); b = a; a = tmp;
  }
  h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e;
}

/* ---------------------------------------------------------------- MD5 (RFC 1321 §3.4) */
static const uint8_t md5_s[64] = {
  7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22,
  5, 9, 14, 20, 5, 9, 14, 20, 5, 9, 14, 20, 5, 9, 14, 20,
  4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23,
  6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21};
/* T[i] = floor(2^32 * |sin(i+1)|), RFC 1321 §3.4 */
static const uint32_t md5_t[64] = {
  0xd76aa478, 0xe8c7b756, 0x242070db, 0xc1bdceee, 0xf57c0faf, 0x4787c62a, 0xa8304613, 0xfd469501,
  0x698098d8, 0x8b44f7af, 0xffff5bb1, 0x895cd7be, 0x6b901122, 0xfd987193, 0xa679438e, 0x49b40821,
  0xf61e2562, 0xc040b340, 0x265e5a51, 0xe9b6c7aa, 0xd62f105d, 0x02441453, 0xd8a1e681, 0xe7d3fbc8,
  0x21e1cde6, 0xc33707d6, 0xf4d50d87, 0x455a14ed, 0xa9e3e905, 0xfcefa3f8, 0x676f02d9, 0x8d2a4c8a,
  0xfffa3942, 0x8771f681, 0x6d9d6122, 0xfde5380c, 0xa4beea44, 0x4bdecfa9, 0xf6bb4b60, 0xbebfbc70,
  0x289b7ec6, 0xeaa127fa, 0xd4ef3085, 0x04881d05, 0xd9d4d039, 0xe6db99e5, 0x1fa27cf8, 0xc4ac5665,
  0xf4292244, 0x432aff97, 0xab9423a7, 0xfc93a039, 0x655b59c3, 0x8f0ccc92, 0xffeff47d, 0x85845dd1,
  0x6fa87e4f, 0xfe2ce6e0, 0xa3014314, 0x4e0811a1, 0xf7537e82, 0xbd3af235, 0x2ad7d2bb, 0xeb86d391};

static void md5_block(uint32_t h[4], const uint8_t* p) {
  uint32_t m[16];
  for (int i = 0; i < 16; i++) m[i] = le32(p + 4 * i);
  uint32_t a = h[0], b = h[1], c = h[2], d = h[3];
  for (int i = 0; i < 64; i++) {
    uint32_t f; int g;
    if (i < 16)      { f = (b & c) | (~b & d); g = i; }
    else if (i < 32) { f = (d & b) | (~d & c); g = (5 * i + 1) & 15; }
    else if (i < 48) { f = b ^ c ^ d;          g = (3 * i + 5) & 15; }
    else             { f = c ^ (b | ~d);       g = (7 * i) & 15; }
    uint32_t tmp = d; d = c; c = b;
    b = b + rol(a + f + md5_t[i] + m[g], md5_s[i]);
    a = tmp;
  }
  h[0] += a; h[1] += b; h[2] += c; h[3] += d;
}

/* ---------------------------------------------------------------- SHA-256 (FIPS 180-4 §6.2.2) */
static const uint32_t sha256_k[64] = {
  0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5,
  0xd807aa98, 0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174,
  0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da,
  0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967,
  0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85,
  0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070,
  0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3,
  0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};

static void sha256_block(uint32_t h[8], const uint8_t* p) {
  uint32_t w[64];
  for (int t = 0; t < 16; t++) w[t] = be32(p + 4 * t);
  for (int t = 16; t < 64; t++) {
    uint32_t s0 = ror(w[t - 15], 7) ^ ror(w[t - 15], 18) ^ (w[t - 15] >> 3);
    uint32_t s1 = ror(w[t - 2], 17) ^ ror(w[t - 2], 19) ^ (w[t - 2] >> 10);
    w[t] = s1 + w[t - 7] + s0 + w[t - 16];
  }
  uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
  for (int t = 0; t < 64; t++) {
    uint32_t S1 = ror(e, 6) ^ ror(e, 11) ^ ror(e, 25);
    uint32_t ch = (e & f) ^ (~e & g);
    uint32_t t1 = hh + S1 + ch + sha256_k[t] + w[t];
    uint32_t S0 = ror(a, 2) ^ ror(a, 13) ^ ror(a, 22);
    uint32_t mj = (a & b) ^ (a & c) ^ (b & c);
    uint32_t t2 = S0 + mj;
    hh = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
  }
  h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
}

/* ---------------------------------------------------------------- streaming front end */
int orc_digest_len(uint32_t alg) {
  return alg == ORC_SHA1 ? 20 : alg == ORC_MD5 ? 16 : alg == ORC_SHA256 ? 32 : -1;
}

int orc_init(orc_ctx* c, uint32_t alg) {
  memset(c, 0, sizeof(*c));
  c->alg = alg;
  switch (alg) {
    case ORC_SHA1:
      c->h[0] = 0x67452301u; c->h[1] = 0xefcdab89u; c->h[2] = 0x98badcfeu; c->h[3] = 0x10325476u;
      c->h[4] = 0xc3d2e1f0u;
      return 0;
    case ORC_MD5:
      c->h[0] = 0x67452301u; c->h[1] = 0xefcdab89u; c->h[2] = 0x98badcfeu; c->h[3] = 0x10325476u;
      return 0;
    case ORC_SHA256:
      c->h[0] = 0x6a09e667u; c->h[1] = 0xbb67ae85u; c->h[2] = 0x3c6ef372u; c->h[3] = 0xa54ff53au;
      c->h[4] = 0x510e527fu; c->h[5] = 0x9b05688cu; c->h[6] = 0x1f83d9abu; c->h[7] = 0x5be0cd19u;
      return 0;
    default:
      return -1;
  }
}

static void one_block(orc_ctx* c, const uint8_t* p) {
  if (c->alg == ORC_SHA1) sha1_block(c->h, p);
  else if (c->alg == ORC_MD5) md5_block(c->h, p);
  else sha256_block(c->h, p);
}

void orc_update(orc_ctx* c, const void* data, size_t len) {
  const uint8_t* p = (const uint8_t*)data;
  c->nbytes += len;
  if (c->fill) {
    size_t take = 64 - c->fill;
    if (take > len) take = len;
    memcpy(c->buf + c->fill, p, take);
    c->fill += (uint32_t)take; p += take; len -= take;
    if (c->fill < 64) return;
    one_block(c, c->buf);
    c->fill = 0;
  }
  while (len >= 64) { one_block(c, p); p += 64; len -= 64; }
  if (len) { memcpy(c->buf, p, len); c->fill = (uint32_t)len; }
}

int orc_final(orc_ctx* c, uint8_t* out) {
  /* Padding: 0x80, zeros to 56 mod 64, 64-bit bit length (LE for MD5, BE for SHA). */
  uint64_t bits = c->nbytes * 8u;
  uint8_t pad[72];
  size_t padlen = (c->fill < 56) ? (56 - c->fill) : (120 - c->fill);
  memset(pad, 0, sizeof(pad));
  pad[0] = 0x80;
  if (c->alg == ORC_MD5) {
    for (int i = 0; i < 8; i++) pad[padlen + i] = (uint8_t)(bits >> (8 * i));
  } else {
    for (int i = 0; i < 8; i++) pad[padlen + i] = (uint8_t)(bits >> (56 - 8 * i));
  }
  uint64_t keep = c->nbytes;
  orc_update(c, pad, padlen + 8);
  c->nbytes = keep;
  int n = orc_digest_len(c->alg);
  for (int i = 0; i < n / 4; i++) {
    if (c->alg == ORC_MD5) put_le32(out + 4 * i, c->h[i]);
    else put_be32(out + 4 * i, c->h[i]);
  }
  return n;
}

static void one_shot(uint32_t alg, const void* data, size_t len, uint8_t* out) {
  orc_ctx c;
  orc_init(&c, alg);
  orc_update(&c, data, len);
  orc_final(&c, out);
}
void orc_sha1(const void* d, size_t n, uint8_t out[20]) { one_shot(ORC_SHA1, d, n, out); }
void orc_md5(const void* d, size_t n, uint8_t out[16]) { one_shot(ORC_MD5, d, n, out); }
void orc_sha256(const void* d, size_t n, uint8_t out[32]) { one_shot(ORC_SHA256, d, n, out); }

/* ---------------------------------------------------------------- batch (optionally threaded) */
typedef struct {
  uint32_t alg; const uint8_t* base; const uint64_t* off; const uint64_t* len;
  size_t n; uint8_t* out; int tid, nthreads;
} batch_arg;

static void* batch_worker(void* vp) {
  batch_arg* a = (batch_arg*)vp;
  int dl = orc_digest_len(a->alg);
  for (size_t i = (size_t)a->tid; i < a->n; i += (size_t)a->nthreads)
    one_shot(a->alg, a->base + a->off[i], (size_t)a->len[i], a->out + i * (size_t)dl);
  return NULL;
}

void orc_batch(uint32_t alg, const uint8_t* base, const uint64_t* offsets, const uint64_t* lengths,
               size_t n, uint8_t* out, int threads) {
  if (threads < 1) threads = 1;
  if ((size_t)threads > n) threads = n ? (int)n : 1;
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)threads);
  batch_arg* args = (batch_arg*)malloc(sizeof(batch_arg) * (size_t)threads);
  for (int t = 0; t < threads; t++) {
    args[t] = (batch_arg){alg, base, offsets, lengths, n, out, t, threads};
    if (t > 0) pthread_create(&th[t], NULL, batch_worker, &args[t]);
  }
  batch_worker(&args[0]);
  for (int t = 1; t < threads; t++) pthread_join(th[t], NULL);
  free(th); free(args);
}

/* ---------------------------------------------------------------- synthetic data */
uint64_t orc_splitmix64(uint64_t x) {
  x += 0x9e3779b97f4a7c15ull;
  x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
  x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
  return x ^ (x >> 31);
}

void orc_synth_fill(uint8_t* dst, uint64_t msg_index, uint64_t byte_off, uint64_t n, uint64_t seed) {
  uint64_t key = seed + (msg_index << 40);
  for (uint64_t k = 0; k < n; k++) {
    uint64_t pos = byte_off + k;
    uint64_t w = orc_splitmix64(key + (pos >> 3));
    dst[k] = (uint8_t)(w >> (8 * (pos & 7)));
  }
}

/* ---------------------------------------------------------------- S3 framing (minio-js 7.0.12) */
void orc_hex(const uint8_t* in, size_t n, char* out) {
  static const char d[] = "0123456789abcdef";
  for (size_t i = 0; i < n; i++) { out[2 * i] = d[in[i] >> 4]; out[2 * i + 1] = d[in[i] & 15]; }
  out[2 * n] = 0;
}

size_t orc_base64(const uint8_t* in, size_t n, char* out) {
  static const char a[] = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/";
  size_t o = 0;
  for (size_t i = 0; i < n; i += 3) {
    uint32_t v = (uint32_t)in[i] << 16;
    if (i + 1 < n) v |= (uint32_t)in[i + 1] << 8;
    if (i + 2 < n) v |= in[i + 2];
    out[o++] = a[(v >> 18) & 63];
    out[o++] = a[(v >> 12) & 63];
    out[o++] = (i + 1 < n) ? a[(v >> 6) & 63] : '=';
    out[o++] = (i + 2 < n) ? a[v & 63] : '=';
  }
  out[o] = 0;
  return o;
}

void orc_etag_multipart(const uint8_t* md5s, size_t n_parts, uint8_t out[16]) {
  orc_md5(md5s, n_parts * 16, out);
}

uint64_t orc_calc_part_size(uint64_t object_size) {
  uint64_t part = 64ull << 20;
  while (part * 10000ull <= object_size) part += 16ull << 20;
  return part;
}
