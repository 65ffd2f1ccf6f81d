"""Media post-processor — mirror of lib/process.js (the only behaviour the reference's tests cover:
test/process/filter_dirs.js:17-82).  Pure directory walk, no hashing; kept so the stage chain
download -> process -> upload hands the same objects around (`{path}` -> `{files, downloadPath}`)."""
from __future__ import annotations

import json
import os
import re

from . import MediaType

# One rule table for this mirror and for js/lib/process.js (which cannot be executed here): both read
# js/lib/process_rules.json, so the copy that is tested and the copy that ships cannot drift apart.
_RULES_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "js", "lib", "process_rules.json")
with open(_RULES_PATH) as _f:
    RULES = json.load(_f)


def _js_regex(rule) -> "re.Pattern":
    """The JSON holds JavaScript regex sources; the ones used here (alternation, \\d, an escaped '/') mean the same in Python."""
    flags = re.I if "i" in rule.get("flags", "") else 0
    return re.compile(rule["source"].replace("\\/", "/"), flags)


MEDIA_EXTS = tuple(RULES["mediaExts"])                   # lib/process.js:15-20
_RE_SKIP = _js_regex(RULES["skipDirectory"])              # lib/process.js:59
_RE_SEASON = _js_regex(RULES["seasonDirectory"])          # lib/process.js:64


async def find_media_files(abs_path: str, media, logger) -> list:
    """lib/process.js:29-99.  klaw walks breadth-first in fs.readdir order; entries are sorted here so the
    result is deterministic (the reference's first test relies on `S1` preceding `Season 1`)."""
    files = []

    def keep(item: str) -> bool:
        name = os.path.basename(item)
        if os.path.isdir(item):
            try:  # the only top-level directory is always entered (lib/process.js:40-51)
                if os.path.exists(os.path.join(abs_path, name)):
                    top = os.listdir(abs_path)
                    if len(top) == 1 and top[0] == name:
                        logger.info(item, "is allowed because its the only top level directory")
                        return True
            except OSError:
                pass
            if media.get("type") == MediaType.MOVIE:   # in movie mode, assume the best
                return True
            if _RE_SKIP.search(item):
                return False
            return bool(_RE_SEASON.search(name))
        return os.path.splitext(name)[1] in MEDIA_EXTS

    queue = [abs_path]
    while queue:
        cur = queue.pop(0)
        if not os.path.isdir(cur):
            files.append(cur)
            continue
        for name in sorted(os.listdir(cur)):
            item = os.path.join(cur, name)
            rel = os.path.relpath(item, abs_path)
            kind = "directory" if os.path.isdir(item) else "file"
            if keep(item):
                logger.info(f"including {kind} '{rel}'")
                queue.append(item)
            else:
                logger.warn(f"skipping {kind} '{rel}'")
    return files


async def factory(config, emitter, logger):
    async def stage(job):
        file = job["lastStage"]
        logger.info("processing directory", file["path"])
        found = await find_media_files(file["path"], job["media"], logger)
        if len(found) == 0:
            raise RuntimeError(RULES["noMediaError"])
        logger.info("found", len(found), "media files")
        return {"files": found, "downloadPath": file["path"]}
    return stage
