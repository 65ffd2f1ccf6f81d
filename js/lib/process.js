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
// One table for this file and for its Python mirror (downloader_b200/stages/process.py), which is the copy that
// runs against the reference's fixture cases where Node is absent: a rule changed here changes there too.
const RULES = require('./process_rules.json')

const WANTED = new Set(RULES.mediaExts)
const SKIP_DIR = new RegExp(RULES.skipDirectory.source, RULES.skipDirectory.flags)
const SEASON_DIR = new RegExp(RULES.seasonDirectory.source, RULES.seasonDirectory.flags)
const MOVIE = 'MOVIE'

function isMovie (media, enums) {
  if (enums && typeof enums.MOVIE !== 'undefined') return media.type === enums.MOVIE
  return media.type === MOVIE || media.type === 1
}

// lib/process.js:36-67, in the reference's order: only-top-level-directory, movie mode, skip, season
function wantDirectory (root, full, media, enums) {
  const base = path.basename(full)
  if (RULES.onlyTopLevelDirectoryIsAlwaysEntered && fs.existsSync(path.join(root, base))) {
    const top = fs.readdirSync(root)
    if (top.length === 1 && top[0] === base) return true
  }
  if (RULES.movieModeEntersEveryDirectory && isMovie(media, enums)) return true
  if (SKIP_DIR.test(full)) return false
  return SEASON_DIR.test(base)
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
    if (files.length === 0) throw new Error(RULES.noMediaError)
    logger.info('found', files.length, 'media files')
    return { files, downloadPath: previous.path }
  }
}
