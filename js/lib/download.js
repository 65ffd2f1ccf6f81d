'use strict'
/**
 * Download stage.  WRITTEN, NOT EXECUTED on this image.
 * Contract of the reference stage (lib/download.js:30,230-277): factory `(config, emitter, logger)`; the job
 * carries `media.{id, source, sourceURI}`; the lower-cased SourceType name picks the method and an unknown
 * one throws 'Protocol not supported.'; progress 0 at the start and 50 at the end; result `{ path }`.
 * A stalled torrent rejects with `err.code === 'ERRDLSTALL'` (the orchestrator acks and drops it).
 *
 * The peer-to-peer transport stays webtorrent's.  What moves to the GPU is the integrity pass: the client is
 * created with `{ skipVerify: true }`-style hooks (config.torrentClient) and every completed range — or, on
 * restart, everything already on disk — is verified by `verifyOnDisk` in batched launches; pieces whose bit
 * is clear are handed back to the client to be requested again (a failed piece is data, not an error).
 */
const fs = require('fs')
const fsp = require('fs').promises
const path = require('path')
const th = require('./tritonhash')

const STALL_MS = 240000
const IN_FLIGHT = 4 // slots submitted and not yet collected

/**
 * Hash pieces of a download directory against info.pieces; returns the verified bitfield as boolean[nPieces].
 * `only` (ascending piece indices) restricts the pass to those pieces — after a refetch only what was requested
 * again is read, copied and hashed (webtorrent verifies a piece when it completes, lib/download.js:64-66);
 * entries of the result for pieces that were not asked about are `undefined`.
 */
async function verifyOnDisk (files, pieceLength, pieces /* Buffer n*20 */, gpu = 0, slotBytes = 256 * 1024 * 1024, only = null) {
  const total = files.reduce((s, f) => s + f.length, 0)
  const nPieces = Math.ceil(total / pieceLength)
  const todo = only || Array.from({ length: nPieces }, (_, i) => i)
  const ok = new Array(nPieces).fill(only ? undefined : false)
  const perSlot = Math.max(1, Math.floor(slotBytes / pieceLength))
  const handles = await Promise.all(files.map(f => fsp.open(f.path, 'r').catch(() => null)))
  const starts = []
  files.reduce((pos, f) => { starts.push(pos); return pos + f.length }, 0)
  const pending = []
  const collect = ({ r, batch }) => { for (let k = 0; k < batch.length; k++) ok[batch[k]] = r.ok[k] }
  // stream bytes [lo, hi) of the files laid end to end (as the torrent sees them) -> buffer at `at`; gaps read as zeros
  const readRange = async (buffer, at, lo, hi) => {
    for (let i = 0; i < files.length; i++) {
      const a = Math.max(lo, starts[i])
      const b = Math.min(hi, starts[i] + files[i].length)
      if (b <= a) continue
      if (handles[i]) await th.readFully(handles[i], buffer, at + (a - lo), b - a, a - starts[i], 'zero')
      else buffer.fill(0, at + (a - lo), at + (b - lo))
    }
  }
  try {
    for (let t = 0; t < todo.length; t += perSlot) {
      const batch = todo.slice(t, t + perSlot)
      const lens = batch.map(i => Math.min(pieceLength, total - i * pieceLength))
      const bytes = (batch.length - 1) * pieceLength + lens[lens.length - 1]
      const { slot, buffer } = await th.ringAcquire(gpu, bytes, pending, collect)
      let submitted
      try {
        for (let k = 0; k < batch.length;) { // consecutive pieces are one read
          let e = k + 1
          while (e < batch.length && batch[e] === batch[e - 1] + 1) e++
          await readRange(buffer, k * pieceLength, batch[k] * pieceLength, batch[e - 1] * pieceLength + lens[e - 1])
          k = e
        }
        const msgs = batch.map((_, k) => ({ offset: k * pieceLength, length: lens[k] }))
        const expected = Buffer.concat(batch.map(i => pieces.slice(i * 20, i * 20 + 20)))
        submitted = th.sha1Batch(slot, msgs, expected)
      } catch (err) {
        th.ringRelease(slot) // a slot that is neither submitted nor released would block the ring for good
        throw err
      }
      // keep reading while the slot is hashed: digests are collected a few slots later
      pending.push(submitted.then(r => ({ r, batch })))
      if (pending.length >= IN_FLIGHT) collect(await pending.shift())
    }
    while (pending.length) collect(await pending.shift())
  } finally {
    await Promise.all(handles.map(h => h && h.close()))
  }
  return ok
}

module.exports = async (config, emitter, logger) => {
  const downloadStage = config.downloadStageEnum
  const sourceName = config.sourceTypeName // (enumValue) => 'TORRENT' | 'HTTP' | 'FILE' | 'BUCKET'
  th.init(config.gpus || [0], config.ringBytes || (1 << 30), config.deviceRingBytes || 0) // device ring: 0 = same size as the pinned ring

  const methods = {
    torrent: async (uri, id, downloadPath) => {
      const client = config.torrentClient
      const torrent = await client.fetch(uri, { path: downloadPath }) // metadata + transport, verification off
      const files = torrent.files.map(f => ({ path: path.join(downloadPath, f.path), length: f.length }))
      const pieces = Buffer.concat(torrent.info.pieces)
      // round 0: bulk verify of whatever is on disk; later rounds: only the pieces that were requested again
      const ok = await verifyOnDisk(files, torrent.pieceLength, pieces)
      for (;;) {
        const have = ok.filter(Boolean).length
        await global.telem.emitProgress(id, downloadStage, Math.floor(have / ok.length * 50))
        if (have === ok.length) return
        const missing = ok.map((good, i) => good ? -1 : i).filter(i => i >= 0)
        missing.forEach(i => logger.warn(`Piece ${i} failed verification`))
        await client.request(torrent, missing, STALL_MS)
        const again = await verifyOnDisk(files, torrent.pieceLength, pieces, 0, undefined, missing)
        let repaired = 0
        for (const i of missing) if (again[i]) { ok[i] = true; repaired++ }
        if (repaired === 0) { // a round that repaired nothing: the swarm cannot supply these pieces
          const err = new Error('Download stalled.')
          err.code = 'ERRDLSTALL'
          throw err
        }
      }
    },
    http: async (uri, id, downloadPath, job) => {
      if (path.extname(new URL(uri).pathname) === '.torrent') {
        logger.info('downloading a .torrent, chaining to torrent downloader')
        return methods.torrent(uri, id, downloadPath, job)
      }
      return config.httpDownload(uri, downloadPath)
    },
    file: async (uri, id, downloadPath) => {
      if (process.env.ALLOW_FILE_URLS !== 'true') throw new Error('File URLs are not allowed.')
      const src = new URL(uri).pathname
      return fsp.copyFile(src, path.join(downloadPath, path.basename(src)))
    },
    bucket: async (uri, id, downloadPath) => config.bucketDownload(uri, downloadPath)
  }

  return async job => {
    const fileId = job.media.id
    const root = path.isAbsolute(config.instance.download_path) ? '' : path.join(__dirname, '..')
    const downloadPath = path.join(root, config.instance.download_path, fileId)
    await fsp.mkdir(downloadPath, { recursive: true }).catch(e => logger.error('Failed to create directory', e.message))
    const protocol = sourceName(job.media.source)
    logger.info(`Trying to download with protocol '${protocol}', the URL '${job.media.sourceURI}'`)
    await global.telem.emitProgress(fileId, downloadStage, 0)
    const method = methods[protocol.toLowerCase()]
    if (!method) throw new Error('Protocol not supported.')
    try {
      await method(job.media.sourceURI, fileId, downloadPath, job)
    } catch (err) {
      logger.error('Download error: ', err.message)
      throw err
    }
    logger.info('finished download')
    await global.telem.emitProgress(fileId, downloadStage, 50)
    return { path: downloadPath }
  }
}
module.exports.verifyOnDisk = verifyOnDisk
