"""Python mirror of the reference's stage plug-ins (lib/download.js, lib/process.js, lib/upload.js).

Same factory contract as the reference (lib/main.js:99-140): `stage = await factory(config, emitter, logger)`
then `result = await stage(job)`, with `job.lastStage` threading results between stages and the module-level
`telem` standing in for `global.telem` (emitProgress(id, stage, percent)).  Node.js is not available on this
image, so these mirrors are what exercises the C ABI the way the stages would.
"""
import enum


# NUMERIC VALUES UNVERIFIED: the enums live in triton-core's protobuf schema (api.Media, api.TelemetryStatusEntry),
# an un-vendored dependency that is absent from the reference tree.  The reference only ever uses the NAMES
# (`proto.enumToString(..., 'SourceType', media.source)` lower-cased picks the downloader, lib/download.js:243;
# `MediaType` MOVIE vs TV picks the filter, lib/process.js:53), and so does this package: the numbers below are
# placeholders in declaration order, not the wire values.  A real deployment must take them from triton-core.
class MediaType(enum.IntEnum):       # triton-core proto api.Media MediaType (names as used by lib/process.js:53)
    TV = 0       # unverified value
    MOVIE = 1    # unverified value


class SourceType(enum.IntEnum):      # api.Media SourceType; lower-cased name selects the downloader (lib/download.js:243,256)
    HTTP = 0     # unverified value
    TORRENT = 1  # unverified value
    FILE = 2     # unverified value
    BUCKET = 3   # unverified value


class TelemetryStatus(enum.IntEnum):  # TelemetryStatusEntry (lib/main.js:68,149; lib/download.js:32)
    QUEUED = 0        # unverified value
    DOWNLOADING = 2   # the literal lib/main.js:68 passes to emitStatus when a job starts
    ERRORED = 6       # the literal lib/main.js:149 passes to emitStatus on a failed stage


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
