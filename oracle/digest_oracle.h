/*
 * digest_oracle.h — CPU ORACLE for the byte-integrity path.  TEST INFRASTRUCTURE ONLY.
 *
 * This is the checker, never the product: only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may load it.  The shipped library
 * (downloader_b200/libtritonhash.so) never links or calls anything in oracle/.
 *
 * PARITY UNPINNED BY THE REFERENCE: tritonmedia/downloader has no test, golden vector or
 * fixture that pins a digest (its only tests are test/process/filter_dirs.js:17-82 and
 * test/eslint.js:1-30).  The arithmetic lives in un-vendored dependencies —
 *   simple-sha1@2.1.2 (yarn.lock:3182)  -> Node crypto.createHash('sha1')    -> OpenSSL
 *   minio@7.0.12      (yarn.lock:2152)  -> Node crypto.createHash('md5'|'sha256') -> OpenSSL
 * reached from lib/download.js:64-66 (client.add -> webtorrent piece verify) and
 * lib/upload.js:45,55 (fPutObject / putObject).  The oracle therefore restates the PUBLISHED
 * algorithms (RFC 1321, FIPS 180-4 §6.1/§6.2) and is pinned against (a) the standards'
 * known-answer vectors in tests/golden/kat.json and (b) this image's OpenSSL 3.0 libcrypto
 * (evp_ref.c) — the same library code Node's crypto executes.
 */
#ifndef DIGEST_ORACLE_H
#define DIGEST_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_SHA1 = 1, ORC_MD5 = 2, ORC_SHA256 = 4 };

/* Streaming context: mirrors crypto.createHash(alg).update(chunk)...digest(). */
typedef struct {
  uint32_t alg;        /* ORC_SHA1 | ORC_MD5 | ORC_SHA256 (exactly one) */
  uint32_t h[8];       /* chaining value */
  uint64_t nbytes;     /* total message bytes absorbed */
  uint8_t buf[64];     /* partial block */
  uint32_t fill;       /* bytes in buf */
} orc_ctx;

int orc_init(orc_ctx* c, uint32_t alg);
void orc_update(orc_ctx* c, const void* data, size_t len);
/* Writes 20 / 16 / 32 bytes; returns the digest length. */
int orc_final(orc_ctx* c, uint8_t* out);

/* One-shot helpers. */
void orc_sha1(const void* data, size_t len, uint8_t out[20]);
void orc_md5(const void* data, size_t len, uint8_t out[16]);
void orc_sha256(const void* data, size_t len, uint8_t out[32]);
int orc_digest_len(uint32_t alg);

/* Batch over a flat buffer: message i = base[offsets[i] .. offsets[i]+lengths[i]).
 * out is n * digest_len bytes.  `threads` > 1 distributes messages by index. */
void orc_batch(uint32_t alg, const uint8_t* base, const uint64_t* offsets,
               const uint64_t* lengths, size_t n, uint8_t* out, int threads);

/* Synthetic data: 64-bit word j of message i = splitmix64(seed + (i<<40) + j),
 * bytes taken little-endian.  Fills bytes [byte_off, byte_off+n) of message i. */
uint64_t orc_splitmix64(uint64_t x);
void orc_synth_fill(uint8_t* dst, uint64_t msg_index, uint64_t byte_off, uint64_t n,
                    uint64_t seed);

/* Framing used by the S3 path (minio-js 7.0.12 behaviour). */
void orc_hex(const uint8_t* in, size_t n, char* out /* 2n+1 */);
size_t orc_base64(const uint8_t* in, size_t n, char* out /* 4*ceil(n/3)+1 */);
/* md5(md5_1 || ... || md5_N) as raw 16 bytes; the S3 ETag string is hex(out)+"-"+N. */
void orc_etag_multipart(const uint8_t* md5s, size_t n_parts, uint8_t out[16]);
/* minio-js calculatePartSize: 64 MiB, then +16 MiB steps until partSize*10000 > size. */
uint64_t orc_calc_part_size(uint64_t object_size);

#ifdef __cplusplus
}
#endif
#endif
