#!/usr/bin/env python
"""north_star's pass bar, ">= 70 % of the per-GPU hashing roofline on 16 MiB-part batches", needs enough parts for
every sub-partition to have work: one chain + helper warp per 32 parts fills the chip at 296 groups = 9,472 parts =
148 GiB, the largest 16 MiB-part batch one B200 (180 GB) holds.  Device-resident, each digest alone and MD5 + SHA-256
together, CUDA events, 64 stratified parts per digest checked against hashlib (`verified`).  One JSON line per run.
(bench.py's default line now carries the same measurement as its `parts_fill_point` workload; this tool remains for
ncu captures: --only md5 --iters 1 launches exactly one kernel.)"""
import argparse
import ctypes as C
import hashlib
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from downloader_b200 import _capi

OPS = {"sha1": 613, "md5": 324, "sha256": 1400}
DL = {"sha1": 20, "md5": 16, "sha256": 32}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--parts", type=int, default=9472)
    ap.add_argument("--mib", type=int, default=16)
    ap.add_argument("--only", default="", help="one of sha1, md5, sha256, md5+sha256")
    ap.add_argument("--iters", type=int, default=2)
    ap.add_argument("--no-verify", action="store_true")
    ap.add_argument("--opt", action="append", default=[], help="th_set_option name=value (tuning), e.g. parts_variant=1")
    ap.add_argument("--no-fuse", action="store_true", help="MD5 + SHA-256 as two kernels on two streams instead of the one-launch kernel")
    a = ap.parse_args()
    L = _capi.lib()
    if a.no_fuse:
        assert L.th_set_option(b"fuse_md5_sha256", 0) == 0
    for o in a.opt:
        k, v = o.split("=")
        assert L.th_set_option(k.encode(), int(v)) == 0, o
    n, ln = a.parts, a.mib << 20
    data = torch.empty(n * ln, dtype=torch.uint8, device="cuda")
    L.th_synth_fill(data.data_ptr(), 0, n, ln, ln, 5, None)
    torch.cuda.synchronize()
    outs = {k: torch.zeros(n * DL[k], dtype=torch.uint8, device="cuda") for k in DL}
    side = torch.cuda.Stream()

    def launch(alg, stream):
        sp = C.c_void_p(stream.cuda_stream)
        if alg == "sha1":
            return L.th_sha1_batch(data.data_ptr(), None, None, None, ln, ln, n, outs[alg].data_ptr(), None, None, sp)
        fn = L.th_md5_batch if alg == "md5" else L.th_sha256_batch
        return fn(data.data_ptr(), None, None, None, ln, ln, n, outs[alg].data_ptr(), sp)

    cur = torch.cuda.current_stream()
    for algs in (["sha1"], ["md5"], ["sha256"], ["md5", "sha256"]):
        if a.only and "+".join(algs) != a.only:
            continue
        best = None
        for _ in range(a.iters):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(cur)
            side.wait_stream(cur)
            if len(algs) > 1 and L.th_md5_sha256_fused(ln, ln, n) == 1:  # one launch does both (digest_parts_kernel)
                assert L.th_md5_sha256_batch(data.data_ptr(), ln, ln, n, outs["md5"].data_ptr(), outs["sha256"].data_ptr(),
                                             C.c_void_p(cur.cuda_stream)) == 0
            else:
                assert launch(algs[0], cur) == 0
                if len(algs) > 1:
                    assert launch(algs[1], side) == 0
                    cur.wait_stream(side)
            e1.record(cur)
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            best = ms if best is None else min(best, ms)
        pick = [] if a.no_verify else sorted({int(x) for x in np.linspace(0, n - 1, 64)})
        for alg in algs:
            for i in pick:
                ref = getattr(hashlib, alg)(data[i * ln:(i + 1) * ln].cpu().numpy().tobytes()).digest()
                assert outs[alg][i * DL[alg]:(i + 1) * DL[alg]].cpu().numpy().tobytes() == ref, (alg, i)
        ops = sum(OPS[k] for k in algs) / 64.0
        roof = 148 * 64 * 1.965e9 / ops / 1e9
        gbps = n * ln / best / 1e6
        print(json.dumps({"algs": "+".join(algs), "parts": n, "part_MiB": a.mib, "GiB": n * ln >> 30, "ms": round(best, 1),
                          "GBps": round(gbps, 1), "int_roofline_GBps": round(roof, 1), "frac": round(gbps / roof, 3),
                          "kernel": ("parts (one launch)" if len(algs) > 1 and L.th_md5_sha256_fused(ln, ln, n) == 1 else
                                     _capi.plan({"sha1": 1, "md5": 2, "sha256": 4}[algs[-1]], ln, ln, n).split(":")[0]),
                          "verified": f"{len(pick)} of {n} vs hashlib"}), flush=True)


if __name__ == "__main__":
    main()
