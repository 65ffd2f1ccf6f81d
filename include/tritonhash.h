/*
 * tritonhash.h — C ABI of the B200 byte-integrity engine (libtritonhash.so).
 *
 * Drop-in boundary for the one data-parallel path of tritonmedia/downloader: the digests its
 * dependencies compute through Node crypto, batched as thousands of independent messages.
 * Each entry point cites the reference interface it replaces (paths are in the reference tree;
 * [dep] = behaviour of the pinned, un-vendored npm dependency).
 *
 * Conventions
 *   - plain C, no C++ types, no exceptions; every call returns int: 0 = TH_OK, < 0 = TH_E*.
 *   - a digest MISMATCH is data (ok_bits), never an error ([dep] webtorrent re-requests the piece).
 *   - there is NO CPU fallback: without a usable GPU th_init fails with TH_ENODEVICE.
 *   - digests are raw bytes in the standard byte order (what crypto's .digest() Buffer holds):
 *     SHA-1 20 B, MD5 16 B, SHA-256 32 B.  hex / base64 are th_hex / th_base64.
 *   - message start addresses must be 16-byte aligned on the device: equal-length batches are staged
 *     by 2-D TMA tiles (a tensor map over {n rows x len bytes}, 16-byte row pitch) and ragged batches by
 *     16-byte cp.async copies.  Everything the HOST can see is checked and refused with TH_EALIGN — the
 *     base pointer, the stride, and every th_msg.offset of th_submit / th_submit_device.  Offsets that
 *     exist only in DEVICE memory (d_offsets of the launch-level calls and of th_digest_ragged_static)
 *     cannot be inspected without a synchronising copy and are NOT validated: a misaligned one faults
 *     in the kernel (a sticky CUDA error), it does not come back as TH_EALIGN.
 *   - thread safety: every th_* call on a context may be made from any thread.  Each GPU of a context
 *     has its own lock, so threads feeding different GPUs do not wait for each other.
 */
#ifndef TRITONHASH_H
#define TRITONHASH_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TH_ALG_SHA1 1u   /* crypto.createHash('sha1')   — [dep] simple-sha1@2.1.2, webtorrent@0.107.6 */
#define TH_ALG_MD5 2u    /* crypto.createHash('md5')    — [dep] minio@7.0.12 getHashSummer(false)    */
#define TH_ALG_SHA256 4u /* crypto.createHash('sha256') — [dep] minio@7.0.12 getHashSummer(true)     */

enum {
  TH_OK = 0,
  TH_EINVAL = -1,    /* bad argument */
  TH_ENODEVICE = -2, /* no CUDA device / not sm_100 capable; product has no CPU path */
  TH_ECUDA = -3,     /* CUDA runtime error (sticky per context; text via th_last_error) */
  TH_ENOMEM = -4,    /* host or device allocation failed */
  TH_EALIGN = -5,    /* message offset / base pointer not 16-byte aligned */
  TH_EBUSY = -6,     /* ring full and the oldest slot has not been submitted yet */
  TH_ETOOBIG = -7,   /* request larger than the ring / a fixed limit */
  TH_ESTATE = -8     /* unknown ticket / slot, or call out of order */
};

typedef struct th_ctx th_ctx;
typedef uint64_t th_ticket;
typedef uint64_t th_slot;

/* One message = bytes [offset, offset+length) of the batch buffer (a ring slot or a device base).
 * Replaces: one piece Buffer handed to sha1(buf, cb) ([dep] webtorrent lib/torrent.js; call site
 * lib/download.js:64-66) or one part range [start,end] streamed into getHashSummer ([dep] minio
 * fPutObject; call site lib/upload.js:45). */
typedef struct th_msg {
  uint64_t offset; /* multiple of 16 */
  uint64_t length; /* any value, 0 .. 2^61 */
} th_msg;

typedef struct th_stats {
  uint64_t messages;    /* messages hashed (per requested digest) */
  uint64_t bytes;       /* message bytes hashed, counted once per message per submit */
  uint64_t launches;    /* kernels launched by this library */
  uint64_t h2d_bytes;   /* bytes copied host -> device (payload + descriptors) */
  uint64_t d2h_bytes;   /* bytes copied device -> host (digests, ok words) */
  uint64_t submits;
  double kernel_ms;     /* summed device time of digest kernels (CUDA events; only while ticket timing is on) */
  uint64_t overflow_allocs; /* job buffers that did not fit the staging arena and were allocated on the submit
                               path (cudaHostAlloc + cudaMalloc synchronise the device: every running chain
                               stalls).  Non-zero = size the rings larger, or submit fewer tiny messages at once */
} th_stats;

/* Completion callback.  Delivered once per ticket on the context's completion thread (a library thread, not
 * CUDA's own callback thread) with no library lock held — the N-API shim forwards it to the JS thread with a
 * napi_threadsafe_function.
 *   - status is the ticket's real outcome: TH_OK, or TH_ECUDA when the device faulted.  A faulted device never
 *     signals completion by itself, so the completion thread doubles as a watchdog: within about 100 ms of a
 *     fault every ticket that still owes a callback gets it with TH_ECUDA (the reference would throw and nack).
 *   - the callback MAY call th_result / th_wait / th_poll / th_submit / th_ring_* on the same context, and may
 *     retire its own ticket; it must not call th_shutdown, and while it runs no other callback is delivered.
 *   - the ticket must still be retired with th_result whatever the status (it frees the ticket's buffers);
 *     th_result called from another thread waits until the callback has returned, so `user` may be freed
 *     right after th_result. */
typedef void (*th_done_cb)(void* user, th_ticket ticket, int status);

/* ---- context ------------------------------------------------------------------------------
 * Replaces nothing in the reference (Node's crypto needs no set-up); owns, per GPU, a pinned host
 * ring + its device mirror, the copy / compute streams and the descriptor staging buffers.
 * gpu_ids == NULL -> devices 0..n_gpus-1.  ring_bytes_per_gpu may be 0 (device-resident use only). */
int th_init(const int* gpu_ids, int n_gpus, size_t ring_bytes_per_gpu, th_ctx** out);
/* Same, with the device-side staging ring sized separately (0 = same as the pinned ring, which is th_init).
 * A slot's pinned bytes are reusable as soon as its H2D copy has run; its device bytes only when the last
 * kernel reading them has finished.  The digest chain of one 16 MiB part runs for ~0.25 s — several times
 * longer than the copy of a whole slot — so to keep PCIe busy with large parts the bytes in flight
 * (≈ 55 GB/s x chain latency) must fit the DEVICE ring; the pinned ring only needs two or three slots.
 * No counterpart in the reference (Node's Buffers are plain heap memory). */
int th_init_staged(const int* gpu_ids, int n_gpus, size_t ring_bytes_per_gpu, size_t device_bytes_per_gpu,
                   th_ctx** out);
int th_shutdown(th_ctx* ctx);
int th_gpu_count(const th_ctx* ctx);

/* ---- host path: pinned ring -> H2D -> digest kernels -> D2H digests -------------------------
 * th_ring_acquire hands out `bytes` of pinned host memory the caller fills (fs.read target —
 * replaces the GC-owned piece/part Buffers of [dep] webtorrent / minio).  Blocks while the pinned ring
 * is full of slots whose H2D copy has not run yet, or the device ring is full of slots whose tickets have not
 * completed (TH_EBUSY instead if the slot it would wait for has not been submitted).  The pinned bytes return
 * to the ring when the slot's copy completes, the device bytes when all of its tickets have. */
int th_ring_acquire(th_ctx* ctx, int gpu, size_t bytes, void** host_ptr, th_slot* slot);
/* Same, but never waits: TH_EBUSY when the bytes do not fit right now.  For callers that must not block their
 * thread — Node's event loop: the add-on returns null and the JS side awaits its oldest outstanding Promise
 * before trying again.  No counterpart in the reference. */
int th_ring_try_acquire(th_ctx* ctx, int gpu, size_t bytes, void** host_ptr, th_slot* slot);
/* Drops a slot that was acquired but will not be submitted. */
int th_ring_release(th_ctx* ctx, th_slot slot);

/* Non-blocking.  Hashes msgs[0..n) of the slot with every digest in alg_mask.
 * A slot may be submitted more than once while one of its tickets is still outstanding — e.g. once as
 * torrent pieces (SHA-1) and once as S3 parts (MD5 | SHA-256) of the same bytes: the slot crosses PCIe
 * once, with its first submit; the slot id stays valid for further submits until all of its tickets have completed
 * (host_ptr must not be written after the first submit).
 * expected_sha1 (nullable, n*20 raw bytes = info.pieces of the .torrent, [dep] parse-torrent@7.0.0)
 * enables the fused compare: bit i of ok_bits = (sha1(msg i) == expected i), which is webtorrent's
 * `hash === this._hashes[index]` -> bitfield.set(index) ([dep] lib/torrent.js). */
int th_submit(th_ctx* ctx, th_slot slot, const th_msg* msgs, size_t n, uint32_t alg_mask,
              const uint8_t* expected_sha1, th_done_cb cb, void* user, th_ticket* out);

/* Same, for data already resident on `gpu` (benchmarks, chained GPU stages).  d_base is a device
 * pointer (16-byte aligned); offsets in msgs are relative to it.  The caller keeps d_base alive
 * until the ticket completes. */
int th_submit_device(th_ctx* ctx, int gpu, const void* d_base, const th_msg* msgs, size_t n,
                     uint32_t alg_mask, const uint8_t* expected_sha1, th_done_cb cb, void* user,
                     th_ticket* out);

/* Device-side timeline of a ticket in milliseconds since th_init (waits for the ticket; call before th_result):
 * ms7[0] inputs on the device; ms7[1], ms7[2] start / end of the SHA-1 kernel(s); [3], [4] MD5; [5], [6] SHA-256;
 * -1 where a digest was not requested, and everywhere for tickets submitted with ticket timing off (the default).
 * Used to see copy / kernel overlap of the host path (tools/config4.py --timeline). */
int th_ticket_timing(th_ctx* ctx, th_ticket ticket, double* ms7);
/* Turns the per-ticket timeline (and th_stats.kernel_ms) on or off for later submits; off by default, or set by
 * TRITONHASH_TICKET_TIMING=1 at th_init.  It is a diagnostic: a timing-enabled CUDA event recorded behind a kernel
 * holds its hardware queue until the kernel ends, so with it on the library keeps to 30 digest streams (one
 * queue each) instead of 120, and at most 30 long chains (large parts) run at once.  No counterpart in the reference. */
int th_set_ticket_timing(th_ctx* ctx, int on);

/* Blocks until the ticket's digests are in host memory. */
int th_wait(th_ctx* ctx, th_ticket ticket);
/* Never blocks: TH_OK when the ticket's digests are in host memory, TH_EBUSY while it is still running, the
 * ticket's error otherwise.  For callers that may not block and do not use callbacks (a poll from Node's event
 * loop, or one thread collecting from several GPUs in whatever order they finish).  No counterpart in the
 * reference (Node's crypto is synchronous). */
int th_poll(th_ctx* ctx, th_ticket ticket);
/* Copies results out (any pointer may be NULL) and retires the ticket.  Waits if needed.
 * sha1: n*20, md5: n*16, sha256: n*32, ok_bits: ceil(n/32) words, bit (i&31) of word i>>5. */
int th_result(th_ctx* ctx, th_ticket ticket, uint8_t* sha1, uint8_t* md5, uint8_t* sha256,
              uint32_t* ok_bits);

/* Multipart object ETag: out16[j] = md5(md5s of object j's parts, concatenated), computed on the
 * GPU by one warp per object (message words travel by warp shuffle).  counts[j] = parts of object j,
 * md5s = all part digests back to back (host memory).  The S3 ETag string is hex(out16)+"-"+count.
 * Blocks until the result is in out16 but allocates nothing (its buffers come from the context's staging
 * arena), so calling it per object does not stall digest chains that are running on the GPU.
 * Replaces the ETag the S3 server returns to [dep] minio completeMultipartUpload (lib/upload.js:45). */
int th_etag_combine(th_ctx* ctx, int gpu, const uint8_t* md5s, const uint32_t* counts,
                    size_t n_objects, uint8_t* out16);

int th_stats_get(th_ctx* ctx, int gpu /* -1 = sum over GPUs */, th_stats* out);
const char* th_strerror(int code);
/* Text of the last CUDA error seen by this context ("" if none). */
const char* th_last_error(const th_ctx* ctx);

/* ---- encoders ([dep] minio getHashSummer emits base64(md5) and lower-case hex(sha256);
 *      simple-sha1 emits lower-case hex) -------------------------------------------------------- */
void th_hex(const uint8_t* in, size_t n, char* out /* 2n+1 */);
size_t th_base64(const uint8_t* in, size_t n, char* out /* 4*ceil(n/3)+1 */);

/* ---- launch-level API: device pointers in, device pointers out, caller's stream ---------------
 * No context, no allocation, no synchronisation: exactly one kernel launch on `stream`
 * (a cudaStream_t; NULL = default stream) of the current device.  These are what th_submit*
 * launch; benchmarks time them with CUDA events on that stream.
 *   message i = d_base + (d_offsets ? d_offsets[i] : i*stride), length d_lengths ? d_lengths[i] : len.
 *   d_order (nullable): thread slot k hashes message d_order[k] (host-side length bucketing);
 *   with d_order the caller zeroes d_ok_bits first.
 * Replaces: th_sha1_batch -> sha1(buf) per piece + compare ([dep] webtorrent _verifyPieces / onChunk);
 *           th_md5_batch / th_sha256_batch -> getHashSummer per part ([dep] minio fPutObject). */
int th_sha1_batch(const void* d_base, const uint64_t* d_offsets, const uint64_t* d_lengths,
                  const uint32_t* d_order, uint64_t stride, uint64_t len, uint32_t n,
                  uint8_t* d_out20, const uint8_t* d_expected20, uint32_t* d_ok_bits, void* stream);
int th_md5_batch(const void* d_base, const uint64_t* d_offsets, const uint64_t* d_lengths,
                 const uint32_t* d_order, uint64_t stride, uint64_t len, uint32_t n,
                 uint8_t* d_out16, void* stream);
int th_sha256_batch(const void* d_base, const uint64_t* d_offsets, const uint64_t* d_lengths,
                    const uint32_t* d_order, uint64_t stride, uint64_t len, uint32_t n,
                    uint8_t* d_out32, void* stream);
/* MD5 and SHA-256 of the same equal-length, fixed-stride messages — what an S3 part needs (Content-MD5 / ETag and
 * x-amz-content-sha256; [dep] minio getHashSummer, call site lib/upload.js:45).  Where the batch is a few large
 * messages (the SHA-256 plan th_plan reports as "assist": at most 296 groups of 32) ONE kernel produces both: per
 * group a SHA-256 chain warp, the helper warp that expands its message schedule, and an MD5 chain warp, each on a
 * sub-partition of its own; the MD5 warp and the helper read the same shared-memory tiles, so the parts are read from
 * HBM once.  th_md5_sha256_fused says whether a batch gets that kernel (1) or the two separate launches, one after the
 * other on `stream` (0) — callers that want the separate kernels to overlap launch th_md5_batch / th_sha256_batch on
 * two streams themselves, as th_submit does. */
int th_md5_sha256_batch(const void* d_base, uint64_t stride, uint64_t len, uint32_t n, uint8_t* d_out16,
                        uint8_t* d_out32, void* stream);
int th_md5_sha256_fused(uint64_t stride, uint64_t len, uint32_t n);
/* d_part_prefix: n_objects+1 exclusive prefix sums of part counts (device). */
int th_etag_combine_device(const uint8_t* d_md5s, const uint64_t* d_part_prefix, uint32_t n_objects,
                           uint8_t* d_out16, void* stream);
/* Synthetic data: 64-bit word j of message i = splitmix64(seed + (i<<40) + j), LE.
 * Fills n messages of `len` bytes at d_dst + k*stride for message index first_msg+k. */
int th_synth_fill(void* d_dst, uint64_t first_msg, uint64_t n, uint64_t len, uint64_t stride,
                  uint64_t seed, void* stream);

/* No counterpart in the reference.  Ragged batch (pieces / parts of unlike lengths) for callers that
 * must not allocate: every other way into the ragged path (th_submit, th_submit_device) builds its
 * length-bucketing table in device memory the library obtains with cudaMalloc / cudaHostAlloc, and an
 * allocation synchronises the whole device and fails outright when the application has already given
 * all of HBM to its own pools or rings — which is exactly the situation of a plug-in called from a
 * media pipeline's callback, or of the N-API add-on once its rings are sized to fill the GPU.
 * This entry point sorts the messages into length classes on the GPU, in a table that is part of the
 * library image (about 257 KiB of device memory per GPU, held until the process exits), and then runs
 * the same digest kernel th_sha1_batch / th_md5_batch / th_sha256_batch run; it calls no allocator.
 *   - at most TH_STATIC_MAX_MSGS messages per call; more is refused with TH_ETOOBIG;
 *   - alg is one of TH_ALG_*; d_offsets / d_lengths / d_out (/ d_expected20, d_ok_bits for SHA-1
 *     verification) are device pointers as in the launch-level calls; d_ok_bits is cleared here;
 *   - the first call on each device looks up the table's device address (cudaGetSymbolAddress) with
 *     that device current; make that first call before you start capturing the stream;
 *   - "calls no allocator" is about this function's own code.  The CUDA driver itself places a library's
 *     kernels and __device__ arrays in device memory when the library is first used on a device, and
 *     that can be this call.  So make the first call on each device (n = 1 is enough), or any other launch
 *     from this library, BEFORE your application hands the last of HBM to its own pools; if the driver
 *     cannot place the library image any more the call returns TH_ENOMEM;
 *   - arguments the host can check (alg, NULL pointers, n, the alignment of d_base) are refused before
 *     anything is enqueued on the stream; d_offsets lives in device memory and is not inspected: every
 *     offset must be a multiple of 16 or the kernel faults (see the conventions at the top);
 *   - all calls on a device share the one table, which keeps what the last call left in it: the
 *     counters are cleared at the start of each call, and two calls whose work overlaps on the GPU
 *     (different streams) race on it — serialise them. */
#define TH_STATIC_MAX_MSGS 65536u
int th_digest_ragged_static(uint32_t alg, const void* d_base, const uint64_t* d_offsets,
                            const uint64_t* d_lengths, uint32_t n, uint8_t* d_out,
                            const uint8_t* d_expected20, uint32_t* d_ok_bits, void* stream);

/* Staging geometry of the digest kernels: bytes each lane stages per pipeline stage (multiple of
 * 64, 64..1024) and the number of stages (2..4).  Process-wide; returns TH_EINVAL if no such
 * variant was compiled.  Defaults are the tuned values in DESIGN.md. */
int th_set_staging(uint32_t alg, int chunk_bytes, int stages);
int th_get_staging(uint32_t alg, int* chunk_bytes, int* stages);
/* Equal-length, fixed-stride batches (all pieces of a torrent but the last, all parts of an object
 * but the last) take the tensor-map kernels: a {128 B x 32 messages} TMA box per copy.  This picks how
 * many such tiles a warp stages per pipeline stage (1, 2, 4) and the number of stages (2..4). */
int th_set_tiling(uint32_t alg, int tiles_per_stage, int stages);
/* Tuning knobs (process-wide): "path" 0 auto | 1 the ragged engine for every batch | 2 one-warp CTAs on tensor-map
 * tiles | 3 persistent time-sliced kernel | 4 chain + schedule-helper warp per group (SHA-1/SHA-256, few
 * large messages); "workers" 4|8|12|16 warps per SM for path 3; "rounds" scheduling rounds per worker for
 * path 3; "sliced_min_groups" / "assist_max_groups" / "tile_many_groups" auto thresholds (in 32-message groups);
 * "ragged_rows" 1 (default) cooperative cp.async row tiles | 0 one bulk copy per lane (th_set_staging geometry)
 * for batches of unlike lengths; "rows_variant" 0..4 = (tiles, stages) (1,3) (1,4) (2,2) (2,3) (4,2) of the former;
 * "fuse_md5_sha256" 1 (default) | 0: whether MD5 | SHA-256 of few large messages share one kernel; "parts_variant" 0..2:
 * tile geometry of that kernel (2 tiles x 2 stages | 4 x 2 | 2 x 3); "md5_lone" 1 | 0: run-ahead stream for MD5 chains that have a
 * sub-partition to themselves; "assist_variant" 0 (2 ring slots of 2 blocks), 1..5, 7 other geometries. */
int th_set_option(const char* name, long long value);
/* Describes, as text, the kernel and geometry an equal-length batch of n messages would be run
 * with on the current device (what benchmarks record next to their numbers). */
int th_plan(uint32_t alg, uint64_t len, uint64_t stride, uint32_t n, char* out, size_t cap);

#ifdef __cplusplus
}
#endif
#endif /* TRITONHASH_H */
