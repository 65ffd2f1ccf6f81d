'use strict'
/**
 * Job runner.  WRITTEN, NOT EXECUTED on this image.
 * Keeps what the stages depend on from the reference orchestrator (lib/main.js:28-32,99-151,197-204):
 * `module.exports = async config => termHandler`; stages are loaded by name from this directory, must export
 * a factory, run in order, and each receives the previous result as `job.lastStage`; an error whose code is
 * 'ERRDLSTALL' drops the message (ack), any other error marks the job ERRORED and requeues it (nack).
 * Transport, telemetry and the protobuf codec are injected (`config.queue`, `config.telem`, `config.codec`) —
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
      let lastStage
      for (const name of STAGES) lastStage = await table[name](Object.assign(Object.create(job), { lastStage }))
      await config.queue.publish('v1.convert', config.codec.encodeConvert({ createdAt: new Date().toISOString(), media: job.media }))
      message.ack()
    } catch (err) {
      if (err.code === 'ERRDLSTALL') {
        message.ack() // dropped, as in the reference
      } else {
        await telem.emitStatus(jobId, STATUS_ERRORED)
        message.nack()
      }
    } finally {
      active--
    }
  }

  await config.queue.listen('v1.download', handle)
  return async function termHandler () {
    await config.queue.close()
    th.shutdown()
    return active
  }
}
module.exports.STAGES = STAGES
