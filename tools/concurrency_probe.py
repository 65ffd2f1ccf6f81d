#!/usr/bin/env python
"""Do long digest kernels of different submits overlap?  N device-resident submits of the same 64 x 64 MiB parts
(MD5 | SHA-256): wall time for N = 1, 2, 4, 8 should stay near one chain latency if they run concurrently."""
import os
import sys
import time

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from downloader_b200 import ALG_MD5, ALG_SHA256, ALG_SHA1, Engine, _capi

L = _capi.lib()
n, ln = 64, 64 << 20
d = torch.empty(n * ln, dtype=torch.uint8, device="cuda")
L.th_synth_fill(d.data_ptr(), 0, n, ln, ln, 3, None)
torch.cuda.synchronize()
offs = np.arange(n, dtype=np.uint64) * ln
lens = np.full(n, ln, np.uint64)
po = np.arange(n * 32, dtype=np.uint64) * (2 << 20)
pl = np.full(n * 32, 2 << 20, np.uint64)
with Engine([0], ring_bytes=0) as eng:
    for algs, name in ((ALG_SHA256, "sha256"), (ALG_MD5 | ALG_SHA256, "md5+sha256"), (0, "pieces+parts")):
        for N in (1, 2, 4, 8, 12):
            t0 = time.perf_counter()
            ts = []
            for _ in range(N):
                if algs:
                    ts.append((eng.submit_device(d.data_ptr(), offs, lens, algs), n, algs))
                else:
                    ts.append((eng.submit_device(d.data_ptr(), po, pl, ALG_SHA1), n * 32, ALG_SHA1))
                    ts.append((eng.submit_device(d.data_ptr(), offs, lens, ALG_MD5 | ALG_SHA256), n, ALG_MD5 | ALG_SHA256))
            marks = []
            for t, k, a in ts:
                eng.result(t, k, a)
                marks.append(round(time.perf_counter() - t0, 2))
            print(name, "N", N, "seconds", round(time.perf_counter() - t0, 3), marks if name == "pieces+parts" else "", flush=True)
