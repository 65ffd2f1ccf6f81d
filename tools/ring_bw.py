#!/usr/bin/env python
"""Host->device throughput of the pinned ring as a function of ring size (NUMA placement of large pinned
allocations shows up here): cycles 2 GiB slots of 256 KiB SHA-1 pieces through the whole ring twice."""
import json
import os
import sys
import time

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from downloader_b200 import ALG_SHA1, Engine

GiB = 1 << 30
for ring_gib in [int(x) for x in (sys.argv[1:] or ["16", "32", "64", "96"])]:
    slot = 2 * GiB
    n_slots = ring_gib * GiB // slot
    eng = Engine([0], ring_bytes=n_slots * slot)
    offs = np.arange(slot // 262144, dtype=np.uint64) * 262144
    lens = np.full(offs.size, 262144, np.uint64)
    pend = []
    # touch every page once (first pass), then time the second and third passes
    for rep in range(3):
        if rep == 1:
            t0 = time.perf_counter()
        for k in range(n_slots):
            while len(pend) >= n_slots - 1:
                eng.result(pend.pop(0), offs.size, ALG_SHA1)
            sid, view = eng.ring_acquire(slot)
            if rep == 0:
                view[::4096] = 1
            pend.append(eng.submit(sid, offs, lens, ALG_SHA1))
    while pend:
        eng.result(pend.pop(0), offs.size, ALG_SHA1)
    dt = time.perf_counter() - t0
    print(json.dumps({"ring_GiB": ring_gib, "e2e_GBps": round(2 * n_slots * slot / dt / 1e9, 2)}), flush=True)
    eng.close()
