"""Python mirror of the reference's stage plug-ins (lib/download.js, lib/process.js, lib/upload.js).

Same factory contract as the reference (lib/main.js:99-140): `stage = await factory(config, emitter, logger)`
then `result = await stage(job)`, with `job.lastStage` threading results between stages and the module-level
`telem` standing in for `global.telem` (emitProgress(id, stage, percent)).  Node.js is not available on this
image, so these mirrors are what exercises the C ABI the way the stages would.
"""
import enum


class MediaType(enum.IntEnum):       # triton-core proto api.Media MediaType (names as used by lib/process.js:53)
    TV = 0
    MOVIE = 1


class SourceType(enum.IntEnum):      # api.Media SourceType; lower-cased name selects the downloader (lib/download.js:243,256)
    HTTP = 0
    TORRENT = 1
    FILE = 2
    BUCKET = 3


class TelemetryStatus(enum.IntEnum):  # TelemetryStatusEntry (lib/main.js:68,149; lib/download.js:32)
    QUEUED = 0
    DOWNLOADING = 2
    ERRORED = 6


class _NullTelem:
    def __init__(self):
        self.events = []

    async def emitProgress(self, file_id, stage, percent):
        self.events.append((file_id, int(stage), int(percent)))

    async def emitStatus(self, job_id, status):
        self.events.append((job_id, "status", int(status)))


telem = _NullTelem()  # stands in for `global.telem` (lib/main.js:49-53)


class Job(dict):
    """Attribute access over a dict, like the decoded proto message the reference passes around."""
    __getattr__ = dict.get

    def __setattr__(self, k, v):
        self[k] = v
