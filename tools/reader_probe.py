#!/usr/bin/env python
"""File -> pinned ring -> GPU digests, end to end: one staged media file hashed in one pass as
torrent pieces (SHA-1) and S3 parts (MD5 + SHA-256) by reader.hash_media_file.

  --dir /dev/shm      tmpfs: the file is in memory (page-cache-warm by construction; O_DIRECT is refused there)
  --dir <disk path>   a real file system: each timed pass can start COLD (the file's pages are dropped from the page
                      cache with posix_fadvise(DONTNEED) after an fsync) and can read buffered or with O_DIRECT
  --gpus N            N files of --gib each, dealt to N GPUs of one context, one feeder thread + reader pool per GPU
                      (reader.hash_media_files)

One JSON line per (threads, mode) with GB/s and how many bytes came through O_DIRECT.
Usage: python tools/reader_probe.py [--gib 8] [--dir /dev/shm] [--direct] [--cold] [--gpus 1]"""
import argparse
import json
import os
import sys
import time

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import oracle
from downloader_b200 import ALG_MD5, ALG_SHA1, ALG_SHA256, Engine, reader, s3

MiB, GiB = 1 << 20, 1 << 30


def drop_cache(path):
    """Evict a (clean) file's pages from the page cache; True if the call was accepted."""
    fd = os.open(path, os.O_RDONLY)
    try:
        os.fsync(fd)
        os.posix_fadvise(fd, 0, 0, os.POSIX_FADV_DONTNEED)
        return True
    except OSError:
        return False
    finally:
        os.close(fd)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gib", type=int, default=8)
    ap.add_argument("--dir", default="/dev/shm")
    ap.add_argument("--piece-mib", type=int, default=1)
    ap.add_argument("--part-mib", type=int, default=16)
    ap.add_argument("--batch-mib", type=int, default=1024)
    ap.add_argument("--direct", action="store_true", help="also time O_DIRECT reads")
    ap.add_argument("--cold", action="store_true", help="drop the file from the page cache before every timed pass")
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--threads", default="", help="comma-separated reader-pool sizes (default: 1 and the pool default)")
    a = ap.parse_args()
    size = a.gib * GiB + 12345
    block = oracle.synth_batch(9, 64, MiB, threads=8).tobytes()  # 64 MiB pattern, repeated
    paths = [os.path.join(a.dir, f"tritonhash_reader_probe_{k}.bin") for k in range(a.gpus)]
    for path in paths:
        with open(path, "wb") as f:
            left = size
            while left > 0:
                f.write(block[:min(left, len(block))])
                left -= len(block)
    try:
        ranges = s3.part_ranges(size, a.part_mib * MiB)
        eng = Engine(list(range(a.gpus)), ring_bytes=4 * a.batch_mib * MiB, device_bytes=16 * a.batch_mib * MiB)
        jobs = [dict(path=p, piece_length=a.piece_mib * MiB, expected_sha1=None, part_ranges=ranges, batch_bytes=a.batch_mib * MiB)
                for p in paths]
        ref = None
        thread_counts = [int(x) for x in a.threads.split(",")] if a.threads else [1, reader.READ_THREADS]
        for threads in thread_counts:
            reader.READ_THREADS = threads
            reader._POOLS.clear()
            for direct in ([False, True] if a.direct else [False]):
                best, read = None, None
                for _ in range(2):
                    cold = a.cold and all(drop_cache(p) for p in paths)
                    t0 = time.perf_counter()
                    if a.gpus == 1:
                        rs = [reader.hash_media_file(eng, direct=direct, **jobs[0])]
                    else:
                        rs = reader.hash_media_files(eng, [dict(j, direct=direct) for j in jobs])
                    dt = time.perf_counter() - t0
                    if best is None or dt < best:
                        best, read = dt, {k: sum(r["read"][k] for r in rs) for k in ("direct_bytes", "buffered_bytes")}
                r = rs[0]
                if ref is None:
                    ref = r
                    first = np.frombuffer(block[:a.piece_mib * MiB], np.uint8)
                    assert (oracle.batch(oracle.SHA1, first, np.array([0], np.uint64), np.array([first.size], np.uint64))[0]
                            == r["pieces"][ALG_SHA1][0]).all()
                for r2 in rs:  # every file holds the same bytes: every pass and every GPU must give the same digests
                    assert (ref["pieces"][ALG_SHA1] == r2["pieces"][ALG_SHA1]).all()
                    assert (ref["parts"][ALG_MD5] == r2["parts"][ALG_MD5]).all() and (ref["parts"][ALG_SHA256] == r2["parts"][ALG_SHA256]).all()
                print(json.dumps({"dir": a.dir, "files": len(paths), "file_GiB": a.gib, "gpus": a.gpus, "read_threads_per_gpu": threads,
                                  "o_direct": direct, "cold_page_cache": bool(cold), "pieces": len(r["pieces"][ALG_SHA1]), "parts": len(ranges),
                                  "seconds": round(best, 3), "GBps": round(len(paths) * size / best / 1e9, 2), **read,
                                  "verified": "piece 0 vs the oracle; all passes and GPUs give identical digests"}), flush=True)
        eng.close()
    finally:
        for p in paths:
            os.unlink(p)


if __name__ == "__main__":
    main()
