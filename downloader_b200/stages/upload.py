"""Upload stage — mirror of lib/upload.js with the digest pass on the GPU.

Reference: for each file `s3Client.fPutObject('triton-staging', <id>/original/<base64(basename)>, file)`
(lib/upload.js:43-45), progress 50..100 (lib/upload.js:48-51), then the `done` marker object
(lib/upload.js:55) and removal of the download directory (lib/upload.js:60-64).  [dep] minio@7.0.12 hashes
every part on the event-loop thread before sending it; here all parts of a file are hashed in a few GPU
launches (reader.hash_file_parts) and the S3 client only receives ready-made header values.

`config` keys used: `s3` (a client with the minio-js method names bucketExists / makeBucket / putObject /
newMultipartUpload / putPart / completeMultipartUpload / listParts — e.g. oracle.s3mock.S3Mock in tests),
`engine` (downloader_b200.Engine), `part_size` (+ `override_part_size`), `anonymous`, `use_ssl`.
"""
from __future__ import annotations

import base64
import math
import os
import shutil

import numpy as np

from .. import s3 as s3f
from .._capi import ALG_MD5, ALG_SHA256
from ..reader import hash_file_parts
from . import TelemetryStatus
from . import telem as _default_telem

BUCKET = "triton-staging"


async def f_put_object(client, engine, bucket: str, object_name: str, path: str, part_size: int, override: bool,
                       want: int, listed_parts=None) -> str:
    """minio fPutObject with the first (hash) pass replaced by one batched GPU pass.  Returns the ETag."""
    size = os.path.getsize(path)
    plan = s3f.plan_object(size, part_size, override)
    dig = hash_file_parts(engine, path, plan.ranges, ALG_MD5 | ALG_SHA256)
    md5s, shas = dig[ALG_MD5], dig[ALG_SHA256]

    def headers(i):
        h = {}
        if want & ALG_MD5:
            h["Content-MD5"] = s3f.content_md5(md5s[i].tobytes())
        if want & ALG_SHA256:
            h["x-amz-content-sha256"] = s3f.sha256_hex(shas[i].tobytes())
        else:
            h["x-amz-content-sha256"] = "UNSIGNED-PAYLOAD"
        return h

    with open(path, "rb") as f:
        if not plan.multipart:
            return client.putObject(bucket, object_name, f.read(), headers(0))
        upload_id = client.newMultipartUpload(bucket, object_name)
        skip = set(s3f.parts_to_skip(md5s, listed_parts or {}))  # resume: already uploaded and identical
        etags = []
        for i, (s, e) in enumerate(plan.ranges):
            if i + 1 in skip:
                etags.append(s3f.md5_hex(md5s[i].tobytes()))
                continue
            f.seek(s)
            etags.append(client.putPart(bucket, object_name, upload_id, i + 1, f.read(e - s), headers(i)).strip('"'))
        etag = client.completeMultipartUpload(bucket, object_name, upload_id, etags)
        # what the server answered must equal the ETag computed on the GPU from the part digests
        if etag.strip('"') != s3f.multipart_etag(engine, md5s)[0]:
            raise RuntimeError(f"multipart ETag mismatch for {object_name}")
        return etag


async def factory(config, emitter, logger):
    client = config["s3"]
    engine = config["engine"]
    tel = config.get("telem", _default_telem)
    want = s3f.select_digests(config.get("anonymous", False), config.get("use_ssl", True))

    async def stage(job):
        last = job["lastStage"]
        files, download_path = last.get("files"), last.get("downloadPath")
        if not isinstance(files, list):
            raise TypeError(f"Invalid files data type, expected array, got '{type(files).__name__}'")
        logger.info("starting file upload")
        file_id = job["media"]["id"]
        if not client.bucketExists(BUCKET):
            client.makeBucket(BUCKET)
        for i, file in enumerate(files, 1):
            logger.info("upload", os.path.basename(file))
            if not os.path.exists(file):
                logger.error("failed to upload file, not found")
                raise FileNotFoundError(f"{file} not found.")
            filename = base64.b64encode(os.path.basename(file).encode()).decode()
            object_name = os.path.join(file_id, "original/", filename)
            await f_put_object(client, engine, BUCKET, object_name, file, config.get("part_size", s3f.DEFAULT_PART_SIZE),
                               config.get("override_part_size", False), want)
            percent = (i / len(files) * 50) + 50
            await tel.emitProgress(file_id, TelemetryStatus.DOWNLOADING, math.floor(percent))
        # the 4-byte 'done' body goes through the same kernels (degenerate length)
        body = b"true"
        d = engine.hash_batch([body], ALG_MD5 | ALG_SHA256)
        hdr = {"Content-MD5": s3f.content_md5(d[ALG_MD5][0].tobytes())} if want & ALG_MD5 else \
              {"x-amz-content-sha256": s3f.sha256_hex(d[ALG_SHA256][0].tobytes())}
        client.putObject(BUCKET, os.path.join(file_id, "original/", "done"), body, hdr)
        logger.info("finished uploading all files")
        try:
            shutil.rmtree(download_path)
        except OSError as err:
            logger.warn("err", f"failed to clean up directory: {err}")
    return stage
