'use strict'
/**
 * Batched digest front end over the N-API add-on (js/addon).  WRITTEN, NOT EXECUTED on this image.
 *
 * Keeps the call shapes the reference's dependencies use, but batched and asynchronous:
 *   simple-sha1:   sha1(buf, cb) / sha1.sync(buf)             -> sha1Batch(slot, msgs, expected?)
 *   minio-js:      getHashSummer(enableSHA256) -> {md5sum: base64, sha256sum: hex}
 *                                                             -> partDigests(slot, parts, {enableSHA256})
 */
const addon = require('../addon/build/Release/tritonhash.node')

const SHA1 = addon.SHA1
const MD5 = addon.MD5
const SHA256 = addon.SHA256

let started = false

// deviceBytesPerGpu (0 = same as the pinned ring) sizes the device staging ring: th_init_staged
function init (gpuIds = [0], ringBytesPerGpu = 1 << 30, deviceBytesPerGpu = 0) {
  if (!started) {
    addon.init(gpuIds, ringBytesPerGpu, deviceBytesPerGpu) // throws when no sm_100 GPU: there is no CPU fallback
    started = true
  }
}

function shutdown () {
  if (started) addon.shutdown()
  started = false
}

/**
 * pinned slot the caller fills, e.g. `await fd.read(buffer, 0, n, pos)`.  Never blocks the event loop: while the
 * ring is full it awaits the caller's oldest outstanding digest Promise (`pending`, shifted by the caller) and retries.
 */
async function ringAcquire (gpu, bytes, pending = [], collect = () => {}) {
  for (;;) {
    const r = addon.ringAcquire(gpu, bytes)
    if (r) return r
    if (pending.length) collect(await pending.shift())
    else await new Promise(resolve => setImmediate(resolve)) // slots of other callers: let their completions run
  }
}

/**
 * `fh.read` may return fewer bytes than asked for: one read(2) moves at most 0x7ffff000 bytes, and network file
 * systems return short counts at will.  Reads until `length` bytes are in `buffer`, or the file ends.
 * onEof 'throw': a short file is an error (an upload must not hash bytes the file does not have);
 * onEof 'zero':  the missing tail reads as zeros (a verify pass: such a piece simply fails its digest).
 * The slot is recycled pinned memory — whatever is not written here still holds another slot's bytes.
 * @returns {Promise<number>} bytes actually read from the file
 */
async function readFully (fh, buffer, offset, length, position, onEof = 'throw') {
  let got = 0
  while (got < length) {
    const { bytesRead } = await fh.read(buffer, offset + got, length - got, position + got)
    if (bytesRead <= 0) break
    got += bytesRead
  }
  if (got < length) {
    if (onEof === 'throw') throw new Error(`short read: wanted ${length} bytes at ${position}, file ended after ${got}`)
    buffer.fill(0, offset + got, offset + length)
  }
  return got
}

/** gives back a slot that will not be submitted (e.g. the read that was to fill it failed) */
function ringRelease (slot) {
  addon.ringRelease(slot)
}

/**
 * SHA-1 of every message of a slot; with `expected` (n x 20 raw bytes, info.pieces) also the verified
 * bitfield: bit i <=> webtorrent's `hash === this._hashes[i]`.
 * @returns {Promise<{hex: string[], ok: boolean[]|null}>}
 */
async function sha1Batch (slot, msgs, expected) {
  const r = await addon.hashBatch(SHA1, slot, msgs, expected)
  const hex = msgs.map((_, i) => r.sha1.toString('hex', i * 20, i * 20 + 20)) // lower-case hex, as simple-sha1
  let ok = null
  if (r.okBits) ok = msgs.map((_, i) => ((r.okBits.readUInt32LE((i >>> 5) * 4) >>> (i & 31)) & 1) === 1)
  return { hex, ok, raw: r.sha1 }
}

/** getHashSummer for all parts at once: [{md5sum: base64 | '', sha256sum: hex | ''}] plus the raw md5s */
async function partDigests (slot, parts, { enableSHA256 = false, both = false } = {}) {
  const algs = both ? (MD5 | SHA256) : (enableSHA256 ? SHA256 : MD5)
  const r = await addon.hashBatch(algs, slot, parts)
  const out = parts.map((_, i) => ({
    md5sum: r.md5 ? r.md5.toString('base64', i * 16, i * 16 + 16) : '',
    sha256sum: r.sha256 ? r.sha256.toString('hex', i * 32, i * 32 + 32) : ''
  }))
  return { digests: out, md5: r.md5 || null, sha256: r.sha256 || null }
}

/** S3 multipart ETag of one object from its raw part md5s: hex(md5(concat)) + '-' + N */
function multipartEtag (md5s) {
  const n = md5s.length / 16
  return addon.etagCombine(md5s, [n]).toString('hex') + '-' + n
}

module.exports = { SHA1, MD5, SHA256, init, shutdown, ringAcquire, ringRelease, readFully, sha1Batch, partDigests, multipartEtag, stats: () => addon.stats() }
