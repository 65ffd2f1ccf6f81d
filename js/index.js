'use strict'
/**
 * Process entry.  WRITTEN, NOT EXECUTED on this image.  Same role as the reference's index.js:17-40: build
 * the configuration, start the job runner, and on SIGINT / unhandledRejection drain it before exiting.
 */
const main = require('./lib/main')

async function init (config) {
  const stop = await main(config)
  const exit = async code => {
    try { await stop() } finally { process.exit(code) }
  }
  process.on('SIGINT', () => exit(0))
  process.on('unhandledRejection', err => {
    (config.logger || console).error('unhandled rejection', err)
    exit(1)
  })
}

module.exports = init
if (require.main === module) init(require(process.env.DOWNLOADER_CONFIG || './config'))
