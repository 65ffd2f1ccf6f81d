#!/usr/bin/env python
"""Device-resident throughput of RAGGED batches (unequal lengths) through th_submit_device, per digest.
Distributions: small files (uniform 1..64 KiB), log-normal media sidecars, and near-equal pieces (256 KiB +- jitter).
Kernel time comes from th_ticket_timing (device events); one JSON line per (distribution, digest)."""
import json
import os
import sys
import time

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from downloader_b200 import ALG_MD5, ALG_SHA1, ALG_SHA256, Engine, _capi

IDX = {ALG_SHA1: 1, ALG_MD5: 3, ALG_SHA256: 5}
NAME = {ALG_SHA1: "sha1", ALG_MD5: "md5", ALG_SHA256: "sha256"}


def main():
    for kv in sys.argv[1:]:  # th_set_option pairs, e.g. ragged_rows=0 rows_variant=3
        k, v = kv.split("=")
        assert _capi.lib().th_set_option(k.encode(), int(v)) == 0, kv
    rng = np.random.default_rng(5)
    dists = {
        "uniform 1..64 KiB x 200k": rng.integers(1, 65536, 200_000),
        "lognormal median 8 KiB x 300k": np.minimum(np.exp(rng.normal(np.log(8192), 1.2, 300_000)).astype(np.int64) + 1, 8 << 20),
        "256 KiB +- 4 KiB x 32k": rng.integers(258_048, 266_240, 32_768),
        "sorted uniform 1..64 KiB x 200k": np.sort(rng.integers(1, 65536, 200_000)),
    }
    L = _capi.lib()
    torch.cuda.set_device(0)
    eng = Engine([0], ring_bytes=1 << 20)
    timed = os.environ.get("PROBE_TIMING", "1") != "0"  # 0: wall clock around submit..result, ticket timing off
    eng.set_timing(timed)
    for name, lens in dists.items():
        lens = lens.astype(np.uint64)
        offs = np.zeros(lens.size, np.uint64)
        offs[1:] = np.cumsum((lens[:-1] + 15) // 16 * 16)
        total = int(offs[-1] + lens[-1])
        data = torch.empty((total + 255) // 256 * 256, dtype=torch.uint8, device="cuda")
        L.th_synth_fill(data.data_ptr(), 0, 1, data.numel(), data.numel(), 3, None)
        torch.cuda.synchronize()
        for alg in (ALG_SHA1, ALG_MD5, ALG_SHA256):
            best, host = None, None
            for _ in range(3):
                h0 = time.perf_counter()
                t = eng.submit_device(data.data_ptr(), offs, lens, alg)
                h1 = time.perf_counter() - h0
                host = h1 if host is None else min(host, h1)
                tm = eng.timing(t)
                eng.result(t, lens.size, alg)
                ms = (tm[IDX[alg] + 1] - tm[IDX[alg]]) if timed else (time.perf_counter() - h0) * 1e3
                best = ms if best is None else min(best, ms)
            print(json.dumps({"dist": name, "alg": NAME[alg], "n": int(lens.size), "GiB": round(int(lens.sum()) / 2**30, 2),
                              "ms": round(best, 3), "submit_host_ms": round(host * 1e3, 3), "GBps": round(int(lens.sum()) / best / 1e6, 1)}), flush=True)
        del data
    eng.close()


if __name__ == "__main__":
    main()
