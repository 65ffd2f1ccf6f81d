'use strict'
/**
 * Job runner.  WRITTEN, NOT EXECUTED on this image.
 * Keeps what the stages depend on from the reference orchestrator (lib/main.js:28-32,99-151,197-204):
 * `module.exports = async config => termHandler`; stages are loaded by name from this directory, must export
 * a factory, run in order, and each receives the previous result as `job.lastStage`; an error whose code is
 * 'ERRDLSTALL' drops the message (ack), any other error marks the job ERRORED and requeues it (nack).
 * A job whose `<id>/original/done` object already exists in `triton-staging` skips the stages (lib/main.js:117-124).
 * Transport, telemetry, the S3 client and the protobuf codec are injected (`config.queue`, `config.telem`,
 * `config.s3`, `config.codec`) —
 * the AMQP / Prometheus / health-check plumbing of the reference is out of scope here (SURVEY.md §2).
 */
const path = require('path')
const th = require('./tritonhash')

const STAGES = ['download', 'process', 'upload']
const STATUS_DOWNLOADING = 2
const STATUS_ERRORED = 6

async function loadStages (config, emitter, logger) {
  const table = {}
  for (const name of STAGES) {
    const factory = require(path.join(__dirname, name + '.js'))
    if (typeof factory !== 'function') throw new Error(`Invalid stage '${name}': not a function`)
    table[name] = await factory(config, emitter, logger.child ? logger.child({ stage: name }) : logger)
  }
  return table
}

module.exports = async config => {
  const logger = config.logger || console
  const telem = config.telem
  global.telem = telem
  let active = 0

  async function handle (message) {
    const job = config.codec.decodeDownload(message.content)
    const jobId = job.media.id
    active++
    await telem.emitStatus(jobId, STATUS_DOWNLOADING)
    try {
      const table = await loadStages(config, null, logger)
      // A job whose `done` marker is already in staging was finished by an earlier delivery: skip every stage and
      // only publish the convert job again (lib/main.js:117-124,153-155) — that is what upload.js writes `done` for.
      let staged = false
      try {
        await config.s3.getObject('triton-staging', path.join(jobId, 'original/', 'done'))
        staged = true
      } catch (err) {
        logger.info('failed to find done file in staging', err && err.message)
      }
      if (!staged) {
        let lastStage = {}
        for (const name of STAGES) lastStage = await table[name](Object.assign(Object.create(job), { lastStage }))
      } else {
        logger.warn('skipping download due to files existing in triton-staging')
      }
    } catch (err) {
      active--
      if (err.code === 'ERRDLSTALL') return message.ack() // dropped, as in the reference
      await telem.emitStatus(jobId, STATUS_ERRORED)
      return message.nack()
    }
    // Outside the stage try/catch, as in the reference (lib/main.js:157-166): a failed publish is logged and the
    // message is left neither acked nor nacked — it is NOT turned into ERRORED + nack.
    try {
      await config.queue.publish('v1.convert', config.codec.encodeConvert({ createdAt: new Date().toISOString(), media: job.media }))
    } catch (err) {
      active--
      return logger.error('failed to create job:', err.message || err)
    }
    message.ack()
    active--
  }

  await config.queue.listen('v1.download', handle)
  return async function termHandler () {
    await config.queue.close()
    th.shutdown()
    return active
  }
}
module.exports.STAGES = STAGES
