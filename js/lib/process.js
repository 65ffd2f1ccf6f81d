'use strict'
/**
 * Media filter stage.  WRITTEN, NOT EXECUTED on this image (no Node.js).
 * Keeps the reference stage's contract (lib/process.js:101-122): factory `(config, emitter, logger)`,
 * job in, `{ files, downloadPath }` out, and its rules: only media extensions; in TV mode only
 * season-like directories, never extras / commentary; the only top-level directory is always entered;
 * in MOVIE mode every directory.  Walks the tree itself (breadth first, sorted) instead of using klaw.
 */
const fs = require('fs')
const path = require('path')

const WANTED = new Set(['.mp4', '.mkv', '.mov', '.webm'])
const MOVIE = 'MOVIE'

function isMovie (media, enums) {
  if (enums && typeof enums.MOVIE !== 'undefined') return media.type === enums.MOVIE
  return media.type === MOVIE || media.type === 1
}

function wantDirectory (root, full, media, enums) {
  const base = path.basename(full)
  if (path.dirname(full) === root) {
    const top = fs.readdirSync(root)
    if (top.length === 1 && top[0] === base) return true
  }
  if (isMovie(media, enums)) return true
  if (/\/extras|\/commentary/i.test(full)) return false
  return /s\d+|season/i.test(base)
}

function collect (root, media, logger, enums) {
  const found = []
  const pending = [root]
  while (pending.length > 0) {
    const dir = pending.shift()
    for (const entry of fs.readdirSync(dir).sort()) {
      const full = path.join(dir, entry)
      const rel = path.relative(root, full)
      if (fs.statSync(full).isDirectory()) {
        if (wantDirectory(root, full, media, enums)) {
          logger.info(`including directory '${rel}'`)
          pending.push(full)
        } else {
          logger.warn(`skipping directory '${rel}'`)
        }
      } else if (WANTED.has(path.extname(entry))) {
        logger.info(`including file '${rel}'`)
        found.push(full)
      } else {
        logger.warn(`skipping file '${rel}'`)
      }
    }
  }
  return found
}

module.exports = async (config, emitter, logger) => {
  const enums = config && config.mediaTypes
  return async job => {
    const previous = job.lastStage
    logger.info('processing directory', previous.path)
    const files = collect(previous.path, job.media, logger, enums)
    if (files.length === 0) throw new Error('Failed to find any suitable media files')
    logger.info('found', files.length, 'media files')
    return { files, downloadPath: previous.path }
  }
}
