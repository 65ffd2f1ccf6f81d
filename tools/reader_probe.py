#!/usr/bin/env python
"""File -> pinned ring -> GPU digests, end to end: one staged media file on tmpfs hashed in one
pass as torrent pieces (SHA-1) and S3 parts (MD5 + SHA-256) by reader.hash_media_file, with 1 reader thread and
with the default pool.  Usage: python tools/reader_probe.py [--gib 8] [--dir /dev/shm]"""
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gib", type=int, default=8)
    ap.add_argument("--dir", default="/dev/shm")
    ap.add_argument("--piece-mib", type=int, default=1)
    ap.add_argument("--part-mib", type=int, default=16)
    ap.add_argument("--batch-mib", type=int, default=1024)
    a = ap.parse_args()
    path = os.path.join(a.dir, "tritonhash_reader_probe.bin")
    size = a.gib * GiB + 12345
    block = oracle.synth_batch(9, 64, MiB, threads=8).tobytes()  # 64 MiB pattern, repeated
    with open(path, "wb") as f:
        left = size
        while left > 0:
            f.write(block[:min(left, len(block))])
            left -= len(block)
    try:
        ranges = s3.part_ranges(size, a.part_mib * MiB)
        eng = Engine([0], ring_bytes=4 * a.batch_mib * MiB, device_bytes=16 * a.batch_mib * MiB)
        ref = None
        for threads in (1, reader.READ_THREADS):
            reader.READ_THREADS = threads
            best = None
            for _ in range(2):
                t0 = time.perf_counter()
                r = reader.hash_media_file(eng, path, a.piece_mib * MiB, None, ranges, batch_bytes=a.batch_mib * MiB)
                dt = time.perf_counter() - t0
                best = dt if best is None else min(best, dt)
            if ref is None:
                ref = r
                first = np.frombuffer(block[:a.piece_mib * MiB], np.uint8)
                assert (oracle.batch(oracle.SHA1, first, np.array([0], np.uint64), np.array([first.size], np.uint64))[0]
                        == r["pieces"][ALG_SHA1][0]).all()
            else:
                assert (ref["pieces"][ALG_SHA1] == r["pieces"][ALG_SHA1]).all()
                assert (ref["parts"][ALG_MD5] == r["parts"][ALG_MD5]).all() and (ref["parts"][ALG_SHA256] == r["parts"][ALG_SHA256]).all()
            print(json.dumps({"file_GiB": a.gib, "read_threads": threads, "pieces": len(r["pieces"][ALG_SHA1]), "parts": len(ranges),
                              "seconds": round(best, 3), "GBps": round(size / best / 1e9, 2)}), flush=True)
        eng.close()
    finally:
        os.unlink(path)


if __name__ == "__main__":
    main()
