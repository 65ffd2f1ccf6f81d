"""Download stage — mirror of lib/download.js for the part that touches media bytes: the SHA-1 verification
of torrent pieces.  The network side (peer wire, DHT, trackers, HTTP, S3 GET) is out of scope;
a `fetcher` callable stands in for it.  What is mirrored:

  * protocol dispatch on `media.source` -> methods[protocol.lower()] and 'Protocol not supported.'
    (lib/download.js:243-263); `.torrent` URLs chained from http to torrent (lib/download.js:144-155);
    ALLOW_FILE_URLS gate of file URLs (lib/download.js:177-189);
  * progress 0 at start, 50 at the end, result `{path: downloadPath}` (lib/download.js:255,272-275);
  * [dep] webtorrent: metainfo -> `_hashes`; bulk `_verifyPieces` of whatever is on disk; a piece whose
    digest differs is re-requested, never an error; 'done' once every bit is set — all digests in batched
    GPU launches with the compare fused into the kernel (reader.verify_pieces).
"""
from __future__ import annotations

import os
import shutil
import urllib.parse

import numpy as np

from .. import torrent as tor
from ..reader import verify_pieces
from . import SourceType, TelemetryStatus
from . import telem as _default_telem

MAX_REFETCH_ROUNDS = 3


class DownloadStalled(RuntimeError):
    code = "ERRDLSTALL"  # lib/download.js:94-97: acked and dropped by lib/main.js:144-146


async def factory(config, emitter, logger):
    engine = config["engine"]
    fetcher = config.get("fetcher")  # async (metainfo, missing piece indices, download_path) -> None
    tel = config.get("telem", _default_telem)

    async def torrent(source, file_id, download_path, job):
        data = source if isinstance(source, (bytes, bytearray)) else open(urllib.parse.urlparse(source).path or source, "rb").read()
        meta = tor.parse_metainfo(bytes(data), engine)
        logger.debug("hash", meta.info_hash)
        logger.debug("files", len(meta.files))
        files = meta.on_disk(download_path)
        # Round 0 is webtorrent's bulk `_verifyPieces` of whatever is on disk.  After a refetch only the pieces that
        # were requested again are read, copied over PCIe and hashed (the per-piece `onChunk` shape,
        # lib/download.js:64-66): O(missing bytes) per round, not O(torrent).
        res = verify_pieces(engine, files, meta.piece_length, meta.pieces)
        bitfield = res["ok"].copy()
        for _ in range(MAX_REFETCH_ROUNDS + 1):
            await tel.emitProgress(file_id, TelemetryStatus.DOWNLOADING, int(bitfield.mean() * 100) // 2)
            if bitfield.all():
                return {"infoHash": meta.info_hash, "bitfield": bitfield}
            if fetcher is None:
                raise DownloadStalled("Download stalled.")
            missing = [int(i) for i in (~bitfield).nonzero()[0]]
            for i in missing:
                logger.warn(f"Piece {i} failed verification")
            await fetcher(meta, missing, download_path)
            res = verify_pieces(engine, files, meta.piece_length, meta.pieces, indices=missing)
            if not res["ok"].any():  # a round that repaired nothing: the swarm cannot supply these pieces
                raise DownloadStalled("Download stalled.")
            bitfield[np.asarray(missing)[res["ok"]]] = True
        raise DownloadStalled("Download stalled.")

    async def http(resource_url, file_id, download_path, job):
        parsed = urllib.parse.urlparse(resource_url)
        if os.path.splitext(parsed.path)[1] == ".torrent":
            logger.info("downloading a .torrent, chaining to torrent downloader")
            return await torrent(resource_url, file_id, download_path, job)
        raise NotImplementedError("plain HTTP download is network plumbing (out of scope)")

    async def file(resource_url, file_id, download_path, job):
        if os.environ.get("ALLOW_FILE_URLS") != "true":
            raise PermissionError("File URLs are not allowed.")
        src = urllib.parse.unquote(urllib.parse.urlparse(resource_url).path)
        shutil.copyfile(src, os.path.join(download_path, os.path.basename(src)))

    methods = {"torrent": torrent, "http": http, "file": file}

    async def stage(job):
        media = job["media"]
        file_id = media["id"]
        base = config["instance"]["download_path"]
        if not os.path.isabs(base):
            base = os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), base)
        download_path = os.path.join(base, file_id)
        os.makedirs(download_path, exist_ok=True)
        protocol = SourceType(media["source"]).name
        await tel.emitProgress(file_id, TelemetryStatus.DOWNLOADING, 0)
        method = methods.get(protocol.lower())
        if not method:
            raise RuntimeError("Protocol not supported.")
        try:
            await method(media["sourceURI"], file_id, download_path, job)
        except Exception as err:
            logger.error("Download error: ", str(err))
            raise
        logger.info("finished download")
        await tel.emitProgress(file_id, TelemetryStatus.DOWNLOADING, 50)
        return {"path": download_path}
    return stage
