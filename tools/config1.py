#!/usr/bin/env python
"""BASELINE config 1: hash-and-stage one synthetic 1 GiB file as 64 x 16 MiB multipart parts (MD5 + SHA-256)
into the in-process S3 mock — the reference's CPU semantics timed next to the GPU stage.

  reference arm  parts strictly in order on ONE thread (Node's event loop): read [k*16 MiB, (k+1)*16 MiB) in 64 KiB
                 chunks -> EVP_DigestUpdate for MD5 and SHA-256 -> finalise -> base64 / hex -> re-read the part and
                 PUT it (minio reads every part twice) -> complete.  libcrypto = the code Node crypto runs.
  b200 arm       downloader_b200.stages.upload.f_put_object: all parts through the pinned ring in a few launches,
                 then the same PUTs.
The mock re-hashes every body (it plays the server), which costs both arms the same; `hash_s` isolates the pass
this repo replaces.  Prints one JSON line.
"""
import asyncio
import hashlib
import json
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import oracle
from oracle.s3mock import S3Mock

MiB = 1 << 20


def reference_arm(path, size, part):
    mock = S3Mock()
    mock.makeBucket("triton-staging")
    t0 = time.perf_counter()
    hash_s = 0.0
    uid = mock.newMultipartUpload("triton-staging", "o")
    etags = []
    with open(path, "rb") as f:
        for k, start in enumerate(range(0, size, part)):
            h0 = time.perf_counter()
            m5, s256 = hashlib.md5(), hashlib.sha256()   # hashlib = OpenSSL EVP
            f.seek(start)
            left = min(part, size - start)
            while left > 0:                              # fs.createReadStream highWaterMark = 64 KiB
                chunk = f.read(min(65536, left))
                m5.update(chunk)
                s256.update(chunk)
                left -= len(chunk)
            hdr = {"Content-MD5": oracle.base64_(m5.digest()), "x-amz-content-sha256": s256.hexdigest()}
            hash_s += time.perf_counter() - h0
            f.seek(start)                                # pass 2: re-read and send
            etags.append(mock.putPart("triton-staging", "o", uid, k + 1, f.read(min(part, size - start)), hdr))
    etag = mock.completeMultipartUpload("triton-staging", "o", uid, etags)
    return time.perf_counter() - t0, hash_s, etag


def b200_arm(path, size, part):
    from downloader_b200 import Engine
    from downloader_b200.reader import hash_file_parts
    from downloader_b200.stages.upload import f_put_object
    from downloader_b200 import s3 as s3f
    mock = S3Mock()
    mock.makeBucket("triton-staging")
    with Engine([0], ring_bytes=2 * 512 * MiB) as eng:
        hash_file_parts(eng, path, s3f.part_ranges(64 * MiB, part), batch_bytes=512 * MiB)  # warm-up (allocations, clocks)
        h0 = time.perf_counter()
        hash_file_parts(eng, path, s3f.part_ranges(size, part), batch_bytes=512 * MiB)
        hash_s = time.perf_counter() - h0
        t0 = time.perf_counter()
        etag = asyncio.new_event_loop().run_until_complete(f_put_object(mock, eng, "triton-staging", "o", path, part, True, 2 | 4))
        total = time.perf_counter() - t0
    return total, hash_s, etag


def main():
    size, part = 1 << 30, 16 * MiB
    d = tempfile.mkdtemp(dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    path = os.path.join(d, "media.mkv")
    oracle.synth_batch(0, 1, size).tofile(path)
    try:
        rt, rh, retag = reference_arm(path, size, part)
        bt, bh, betag = b200_arm(path, size, part)
        assert retag == betag and retag.endswith("-64")
        print(json.dumps({
            "config": "1 GiB file, 64 x 16 MiB parts, MD5 + SHA-256, in-process S3 mock (tmpfs)",
            "etag": retag,
            "reference": {"total_s": round(rt, 3), "hash_s": round(rh, 3), "hash_GBps": round(size / rh / 1e9, 3), "threads": 1,
                          "what": "hashlib/libcrypto, 64 KiB updates, parts in order, every part read twice"},
            "b200": {"total_s": round(bt, 3), "hash_s": round(bh, 3), "hash_GBps": round(size / bh / 1e9, 3),
                     "what": "file -> pinned ring (preadv) -> H2D -> MD5 + SHA-256 kernels -> digests; then the same PUTs"},
            "hash_speedup": round(rh / bh, 2)}))
    finally:
        os.remove(path)
        os.rmdir(d)


if __name__ == "__main__":
    main()
