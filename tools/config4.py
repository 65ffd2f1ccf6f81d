#!/usr/bin/env python
"""BASELINE config 4: mixed batch, end to end from pinned host memory.

10 synthetic media files, sizes uniform in [4, 40] GiB (seeded), each hashed twice from the same bytes: as
torrent pieces (piece length from {1, 2, 4} MiB, SHA-1) and as S3 parts (part size from {8, 16, 32, 64} MiB,
MD5 + SHA-256), short last piece / part included.  File bytes stream through the pinned ring in slots of
--slot-mib; every slot crosses PCIe once and is submitted twice.  To avoid regenerating ~220 GiB, file content
is periodic (one 64 MiB synthetic pattern, written into the whole ring once, untimed; slots are 64 MiB-aligned),
so digests of equal (offset, length) ranges repeat — which the run checks — and slot 0 is checked against the CPU oracle.
Timed region = acquire + submit + result for every slot of every file (H2D, kernels, D2H inside).
Prints one JSON line (bytes counted once per file).
"""
import argparse
import json
import os
import sys
import time

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import oracle
from downloader_b200 import ALG_MD5, ALG_SHA1, ALG_SHA256, Engine, _capi

MiB, GiB = 1 << 20, 1 << 30


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ring-gib", type=int, default=32)
    ap.add_argument("--slot-mib", type=int, default=1024)
    ap.add_argument("--device-gib", type=int, default=0, help="device staging ring (0 = same as the pinned ring)")
    ap.add_argument("--timeline", action="store_true", help="per-ticket device timeline (th_set_ticket_timing: limits the run to 30 digest streams)")
    ap.add_argument("--files", type=int, default=10)
    ap.add_argument("--scale", type=float, default=1.0, help="scale file sizes (1.0 = 4..40 GiB)")
    a = ap.parse_args()
    rng = np.random.default_rng(0x747269746F6E)
    sizes = [int(rng.uniform(4, 40) * a.scale * GiB) + int(rng.integers(1, MiB)) for _ in range(a.files)]
    plens = [int(rng.choice([1, 2, 4])) * MiB for _ in range(a.files)]
    psizes = [int(rng.choice([8, 16, 32, 64])) * MiB for _ in range(a.files)]
    slot_bytes = a.slot_mib * MiB
    n_slots = (a.ring_gib * GiB) // slot_bytes
    L = _capi.lib()
    torch.cuda.set_device(0)
    PERIOD = 64 * MiB  # ring content repeats every 64 MiB, so any 64 MiB-aligned slot holds the same stream
    pat = torch.empty(PERIOD, dtype=torch.uint8, device="cuda")
    L.th_synth_fill(pat.data_ptr(), 0, 1, PERIOD, PERIOD, 11, None)
    torch.cuda.synchronize()
    win = pat.repeat(slot_bytes // PERIOD)
    n_dev_slots = max(n_slots, (a.device_gib * GiB) // slot_bytes)
    eng = Engine([0], ring_bytes=n_slots * slot_bytes, device_bytes=n_dev_slots * slot_bytes)
    eng.set_timing(a.timeline)
    held = []
    for _ in range(n_slots):  # every slot gets the same window; contents persist across re-acquisition
        sid, view = eng.ring_acquire(slot_bytes)
        torch.from_numpy(view).copy_(win)
        held.append(sid)
    torch.cuda.synchronize()
    for sid in held:
        eng.ring_release(sid)
    host_win = win.cpu().numpy()

    def msgs(nbytes, unit):
        off = np.arange(0, nbytes, unit, dtype=np.uint64)
        ln = np.minimum(unit, nbytes - off).astype(np.uint64)
        return off, ln

    per_file, first = [], {}
    s0 = eng.stats()
    pending = []

    prof = {"retire": 0.0, "acquire": 0.0, "submit": 0.0}
    gpu_tl = []  # per slot: part MiB, copy done, sha1 start/end, md5 start/end, sha256 start/end (s since th_init)
    timeline = []  # (part MiB, seconds since start) at which each slot's results were in hand

    def retire(k):
        while len(pending) > k:
            tp, np_, tq, nq, key = pending.pop(0)
            tm_p, tm_q = eng.timing(tp), eng.timing(tq)
            gpu_tl.append([key[2] // MiB] + [round(x / 1e3, 2) for x in (tm_p[0], tm_p[1], tm_p[2], tm_q[3], tm_q[4], tm_q[5], tm_q[6])])
            t_ = time.perf_counter()
            rp = eng.result(tp, np_, ALG_SHA1)
            rq = eng.result(tq, nq, ALG_MD5 | ALG_SHA256)
            prof["retire"] += time.perf_counter() - t_
            timeline.append((key[2] // MiB, round(time.perf_counter() - T0, 2)))
            ref = first.setdefault(key, (rp, rq))
            assert (ref[0][ALG_SHA1] == rp[ALG_SHA1]).all() and (ref[1][ALG_SHA256] == rq[ALG_SHA256]).all()
            assert (ref[1][ALG_MD5] == rq[ALG_MD5]).all()

    global T0
    T0 = time.perf_counter()
    for size, plen, psize in zip(sizes, plens, psizes):  # the batch is pipelined ACROSS files: no drain at file ends
        pos = 0
        while pos < size:
            nb = min(slot_bytes, size - pos)
            retire(n_dev_slots - 1)  # results are read only when the device ring is about to fill
            t_ = time.perf_counter()
            sid, _ = eng.ring_acquire(-(-nb // PERIOD) * PERIOD)  # 64 MiB-aligned, so every slot holds the same stream
            prof["acquire"] += time.perf_counter() - t_
            po, pl = msgs(nb, plen)
            qo, ql = msgs(nb, psize)
            t_ = time.perf_counter()
            tp = eng.submit(sid, po, pl, ALG_SHA1)
            tq = eng.submit(sid, qo, ql, ALG_MD5 | ALG_SHA256)
            prof["submit"] += time.perf_counter() - t_
            pending.append((tp, po.size, tq, qo.size, (nb, plen, psize)))
            pos += nb
        per_file.append({"GiB": round(size / GiB, 2), "piece_MiB": plen // MiB, "part_MiB": psize // MiB})
    retire(0)
    T = time.perf_counter() - T0
    s1 = eng.stats()
    # oracle check of one geometry (slot 0 content is the generator's message 0)
    (nb, plen, psize), (rp, rq) = next(iter(first.items()))
    po, pl = msgs(nb, plen)
    qo, ql = msgs(nb, psize)
    pick = [0, 1, po.size - 1]
    assert (oracle.batch(oracle.SHA1, host_win, po[pick], pl[pick]) == rp[ALG_SHA1][pick]).all()
    pick = [0, qo.size - 1]
    assert (oracle.batch(oracle.SHA256, host_win, qo[pick], ql[pick]) == rq[ALG_SHA256][pick]).all()
    assert (oracle.batch(oracle.MD5, host_win, qo[pick], ql[pick]) == rq[ALG_MD5][pick]).all()
    total = sum(sizes)
    print(json.dumps({"config": "mixed batch: %d files, 4-40 GiB x %.2f, pieces 1-4 MiB (SHA-1) + parts 8-64 MiB (MD5+SHA-256), e2e from pinned host" % (a.files, a.scale),
                      "ring_GiB": a.ring_gib, "device_ring_GiB": n_dev_slots * slot_bytes // GiB, "slot_MiB": a.slot_mib, "total_GiB": round(total / GiB, 1), "seconds": round(T, 2),
                      "e2e_GBps": round(total / T / 1e9, 2),
                      "h2d_GBps": round((s1["h2d_bytes"] - s0["h2d_bytes"]) / T / 1e9, 2),
                      "launches": s1["launches"] - s0["launches"], "kernel_ms_sum": round(s1["kernel_ms"] - s0["kernel_ms"], 1),
                      "host_seconds": {k: round(v, 2) for k, v in prof.items()}, "slot_done_timeline": timeline, "gpu_timeline": gpu_tl, "files": per_file}))
    eng.close()


if __name__ == "__main__":
    main()
