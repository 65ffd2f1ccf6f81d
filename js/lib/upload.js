'use strict'
/**
 * Upload stage with the per-part digests computed on the GPU.  WRITTEN, NOT EXECUTED on this image.
 * Contract of the reference stage (lib/upload.js:14-66): factory `(config, emitter, logger)`; the job's
 * `lastStage` is `{ files, downloadPath }`; objects go to bucket `triton-staging` as
 * `<media.id>/original/<base64(basename)>`; progress 50..100 through `global.telem.emitProgress`;
 * a `done` object is written last; the download directory is removed; the stage returns nothing.
 *
 * minio-js hashes each part on the event-loop thread right before sending it.  Here a whole file's parts
 * are read into pinned ring slots, hashed in one launch per slot, and the client gets header values.
 * `config.s3` must offer putObjectWithDigest / newMultipartUpload / putPart / completeMultipartUpload
 * (a thin subclass of minio's Client that skips its own hash pass).
 */
const fs = require('fs')
const fsp = require('fs').promises
const path = require('path')
const th = require('./tritonhash')

const BUCKET = 'triton-staging'
const MiB = 1024 * 1024
const IN_FLIGHT = 8 // slots submitted and not yet collected (the device ring holds them; th_init_staged)

function partSizeFor (size, base = 64 * MiB, override = false) {
  if (override) return base
  let p = base
  while (p * 10000 <= size) p += 16 * MiB
  return p
}

async function digestParts (file, size, partSize, gpu, slotBytes, mode) {
  const fd = await fsp.open(file, 'r')
  const md5 = Buffer.alloc(Math.max(1, Math.ceil(size / partSize)) * 16)
  const perPart = []
  const pending = []
  const collect = ({ r, first }) => {
    r.md5.copy(md5, first * 16)
    for (let k = 0; k < r.digests.length; k++) perPart[first + k] = r.digests[k]
  }
  try {
    const partsPerSlot = Math.max(1, Math.floor(slotBytes / partSize))
    const nParts = Math.max(1, Math.ceil(size / partSize))
    for (let first = 0; first < nParts; first += partsPerSlot) {
      const count = Math.min(partsPerSlot, nParts - first)
      const start = first * partSize
      const bytes = Math.min(count * partSize, size - start)
      const { slot, buffer } = await th.ringAcquire(gpu, Math.max(bytes, 16), pending, collect)
      const msgs = []
      for (let k = 0; k < count; k++) msgs.push({ offset: k * partSize, length: Math.min(partSize, size - (first + k) * partSize) })
      let submitted
      try {
        // straight into pinned memory; every byte of every part must come from the file (a file that shrank throws)
        if (bytes > 0) await th.readFully(fd, buffer, 0, bytes, start, 'throw')
        submitted = th.partDigests(slot, msgs, { both: true, enableSHA256: mode.enableSHA256 })
      } catch (err) {
        th.ringRelease(slot) // a slot that is neither submitted nor released would block the ring for good
        throw err
      }
      // a slot's digests arrive one chain latency after its copy (0.25 s for 16 MiB parts): keep reading while they run
      pending.push(submitted.then(r => ({ r, first })))
      if (pending.length >= IN_FLIGHT) collect(await pending.shift())
    }
    while (pending.length) collect(await pending.shift())
  } finally {
    await fd.close()
  }
  return { md5, perPart }
}

module.exports = async (config, emitter, logger) => {
  const s3 = config.s3
  const mode = { enableSHA256: !config.anonymous && !config.useSSL }
  const downloadStage = config.downloadStageEnum
  th.init(config.gpus || [0], config.ringBytes || (1 << 30), config.deviceRingBytes || 0) // device ring: 0 = same size as the pinned ring

  return async job => {
    const { files, downloadPath } = job.lastStage
    if (!Array.isArray(files)) throw new Error(`Invalid files data type, expected array, got '${typeof files}'`)
    logger.info('starting file upload')
    const fileId = job.media.id
    if (!await s3.bucketExists(BUCKET)) await s3.makeBucket(BUCKET)

    let done = 0
    for (const file of files) {
      done++
      logger.info('upload', path.basename(file))
      if (!fs.existsSync(file)) {
        logger.error('failed to upload file, not found')
        throw new Error(`${file} not found.`)
      }
      const objectName = path.join(fileId, 'original/', Buffer.from(path.basename(file)).toString('base64'))
      const { size } = await fsp.stat(file)
      const partSize = partSizeFor(size, config.partSize, config.overridePartSize)
      const { md5, perPart } = await digestParts(file, size, partSize, 0, config.slotBytes || (256 * MiB), mode)
      const header = i => mode.enableSHA256 ? { sha256sum: perPart[i].sha256sum } : { md5sum: perPart[i].md5sum }
      if (size <= partSize) {
        await s3.putObjectWithDigest(BUCKET, objectName, fs.createReadStream(file), size, header(0))
      } else {
        const uploadId = await s3.newMultipartUpload(BUCKET, objectName)
        const etags = []
        for (let i = 0; i < perPart.length; i++) {
          const start = i * partSize
          const end = Math.min(size, start + partSize) - 1
          etags.push(await s3.putPart(BUCKET, objectName, uploadId, i + 1, fs.createReadStream(file, { start, end }), end - start + 1, header(i)))
        }
        const etag = await s3.completeMultipartUpload(BUCKET, objectName, uploadId, etags)
        if (String(etag).replace(/"/g, '') !== th.multipartEtag(md5)) throw new Error(`multipart ETag mismatch for ${objectName}`)
      }
      await global.telem.emitProgress(fileId, downloadStage, Math.floor((done / files.length * 50) + 50))
    }

    await s3.putObject(BUCKET, path.join(fileId, 'original/', 'done'), 'true')
    logger.info('finished uploading all files')
    try {
      await fsp.rm(downloadPath, { recursive: true, force: true })
    } catch (err) {
      logger.warn('err', `failed to clean up directory: ${err.message}`)
    }
  }
}
module.exports.partSizeFor = partSizeFor
