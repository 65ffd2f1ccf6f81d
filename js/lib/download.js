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

/** Re-hash a download directory against info.pieces; returns the verified bitfield as boolean[] */
async function verifyOnDisk (files, pieceLength, pieces /* Buffer n*20 */, gpu = 0, slotBytes = 256 * 1024 * 1024) {
  const total = files.reduce((s, f) => s + f.length, 0)
  const nPieces = Math.ceil(total / pieceLength)
  const ok = new Array(nPieces).fill(false)
  const perSlot = Math.max(1, Math.floor(slotBytes / pieceLength))
  const handles = await Promise.all(files.map(f => fsp.open(f.path, 'r').catch(() => null)))
  const pending = []
  const collect = ({ r, first, count }) => { for (let k = 0; k < count; k++) ok[first + k] = r.ok[k] }
  try {
    for (let first = 0; first < nPieces; first += perSlot) {
      const count = Math.min(perSlot, nPieces - first)
      const start = first * pieceLength
      const bytes = Math.min(count * pieceLength, total - start)
      const { slot, buffer } = await th.ringAcquire(gpu, bytes, pending, collect)
      buffer.fill(0, 0, bytes)
      let pos = 0
      try {
        for (let i = 0; i < files.length; i++) { // files laid end to end, as the torrent sees them
          const lo = Math.max(start, pos)
          const hi = Math.min(start + bytes, pos + files[i].length)
          if (hi > lo && handles[i]) await handles[i].read(buffer, lo - start, hi - lo, lo - pos)
          pos += files[i].length
        }
      } catch (err) {
        th.ringRelease(slot) // an unsubmitted slot would block the ring for good
        throw err
      }
      const msgs = []
      for (let k = 0; k < count; k++) msgs.push({ offset: k * pieceLength, length: Math.min(pieceLength, total - (first + k) * pieceLength) })
      // keep reading while the slot is hashed: digests are collected a few slots later
      pending.push(th.sha1Batch(slot, msgs, pieces.slice(first * 20, (first + count) * 20)).then(r => ({ r, first, count })))
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
      let last = -1
      for (;;) {
        const files = torrent.files.map(f => ({ path: path.join(downloadPath, f.path), length: f.length }))
        const ok = await verifyOnDisk(files, torrent.pieceLength, Buffer.concat(torrent.info.pieces))
        const have = ok.filter(Boolean).length
        await global.telem.emitProgress(id, downloadStage, Math.floor(have / ok.length * 50))
        if (have === ok.length) return
        if (have === last) {
          const err = new Error('Download stalled.')
          err.code = 'ERRDLSTALL'
          throw err
        }
        last = have
        ok.forEach((good, i) => { if (!good) logger.warn(`Piece ${i} failed verification`) })
        await client.request(torrent, ok.map((good, i) => good ? -1 : i).filter(i => i >= 0), STALL_MS)
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
