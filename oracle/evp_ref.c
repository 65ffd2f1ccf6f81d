/*
 * evp_ref.c — libcrypto (OpenSSL 3.0) harness.  TEST INFRASTRUCTURE / CPU BASELINE ONLY.
 *
 * Node's crypto.createHash(alg).update(buf).digest() is a thin binding over libcrypto's
 * EVP_DigestInit_ex / EVP_DigestUpdate / EVP_DigestFinal_ex; the reference reaches it through
 *   lib/download.js:64-66 -> webtorrent@0.107.6 -> simple-sha1@2.1.2 (sha1, one update per piece)
 *   lib/upload.js:45,55   -> minio@7.0.12 getHashSummer (md5|sha256, one update per 64 KiB
 *                            stream chunk, fs.createReadStream default highWaterMark)
 * Node itself is not installable here, so this file drives the same library with the same call
 * pattern.  It is used (a) to pin digest_oracle.c, and (b) as the timed CPU baseline of bench.py.
 * Nothing in the product links this.
 */
#include <openssl/evp.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "digest_oracle.h"

static const EVP_MD* md_of(uint32_t alg) {
  switch (alg) {
    case ORC_SHA1: return EVP_sha1();
    case ORC_MD5: return EVP_md5();
    case ORC_SHA256: return EVP_sha256();
    default: return NULL;
  }
}

/* createHash(alg); for each `chunk` bytes: update(); digest().  chunk==0 -> one update. */
int evp_digest(uint32_t alg, const uint8_t* data, size_t len, size_t chunk, uint8_t* out) {
  const EVP_MD* md = md_of(alg);
  if (!md) return -1;
  EVP_MD_CTX* c = EVP_MD_CTX_new();
  if (!c) return -2;
  unsigned int n = 0;
  int ok = EVP_DigestInit_ex(c, md, NULL);
  if (chunk == 0) chunk = len ? len : 1;
  for (size_t o = 0; ok && o < len; o += chunk)
    ok = EVP_DigestUpdate(c, data + o, (len - o < chunk) ? (len - o) : chunk);
  if (ok) ok = EVP_DigestFinal_ex(c, out, &n);
  EVP_MD_CTX_free(c);
  return ok ? (int)n : -3;
}

typedef struct {
  uint32_t mask; const uint8_t* base; const uint64_t* off; const uint64_t* len; size_t n;
  size_t chunk; uint8_t *o1, *o5, *o256; int tid, nthreads; int rc;
} evp_arg;

static void* evp_worker(void* vp) {
  evp_arg* a = (evp_arg*)vp;
  a->rc = 0;
  for (size_t i = (size_t)a->tid; i < a->n; i += (size_t)a->nthreads) {
    const uint8_t* p = a->base + a->off[i];
    size_t l = (size_t)a->len[i];
    if ((a->mask & ORC_SHA1) && evp_digest(ORC_SHA1, p, l, a->chunk, a->o1 + 20 * i) != 20) a->rc = -1;
    if ((a->mask & ORC_MD5) && evp_digest(ORC_MD5, p, l, a->chunk, a->o5 + 16 * i) != 16) a->rc = -1;
    if ((a->mask & ORC_SHA256) && evp_digest(ORC_SHA256, p, l, a->chunk, a->o256 + 32 * i) != 32) a->rc = -1;
  }
  return NULL;
}

static double now_s(void) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

/* Hash n messages out of one flat buffer with `threads` threads (messages dealt by index).
 * Returns wall seconds spent hashing (>= 0) or a negative error. */
double evp_batch(uint32_t alg_mask, const uint8_t* base, const uint64_t* offsets,
                 const uint64_t* lengths, size_t n, size_t chunk, int threads, uint8_t* out_sha1,
                 uint8_t* out_md5, uint8_t* out_sha256) {
  if (threads < 1) threads = 1;
  if ((size_t)threads > n) threads = n ? (int)n : 1;
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)threads);
  evp_arg* args = (evp_arg*)malloc(sizeof(evp_arg) * (size_t)threads);
  double t0 = now_s();
  for (int t = 0; t < threads; t++) {
    args[t] = (evp_arg){alg_mask, base, offsets, lengths, n, chunk, out_sha1, out_md5, out_sha256,
                        t, threads, 0};
    if (t > 0) pthread_create(&th[t], NULL, evp_worker, &args[t]);
  }
  evp_worker(&args[0]);
  for (int t = 1; t < threads; t++) pthread_join(th[t], NULL);
  double dt = now_s() - t0;
  int rc = 0;
  for (int t = 0; t < threads; t++) rc |= args[t].rc;
  free(th); free(args);
  return rc ? -1.0 : dt;
}

/* Threaded synthetic fill of n equal-length messages laid out back to back (test/bench set-up). */
typedef struct { uint8_t* dst; uint64_t first, n, len, seed; int tid, nthreads; } fill_arg;
static void* fill_worker(void* vp) {
  fill_arg* a = (fill_arg*)vp;
  for (uint64_t i = (uint64_t)a->tid; i < a->n; i += (uint64_t)a->nthreads) {
    /* word-at-a-time fast path; identical bytes to orc_synth_fill */
    uint64_t key = a->seed + ((a->first + i) << 40);
    uint8_t* d = a->dst + i * a->len;
    uint64_t words = a->len / 8;
    for (uint64_t j = 0; j < words; j++) {
      uint64_t w = orc_splitmix64(key + j);
      memcpy(d + 8 * j, &w, 8); /* little-endian host */
    }
    if (a->len & 7) orc_synth_fill(d + 8 * words, a->first + i, 8 * words, a->len & 7, a->seed);
  }
  return NULL;
}
void evp_synth_fill_mt(uint8_t* dst, uint64_t first_msg, uint64_t n, uint64_t len, uint64_t seed,
                       int threads) {
  if (threads < 1) threads = 1;
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)threads);
  fill_arg* args = (fill_arg*)malloc(sizeof(fill_arg) * (size_t)threads);
  for (int t = 0; t < threads; t++) {
    args[t] = (fill_arg){dst, first_msg, n, len, seed, t, threads};
    if (t > 0) pthread_create(&th[t], NULL, fill_worker, &args[t]);
  }
  fill_worker(&args[0]);
  for (int t = 1; t < threads; t++) pthread_join(th[t], NULL);
  free(th); free(args);
}

const char* evp_version(void) { return OpenSSL_version(OPENSSL_VERSION); }
