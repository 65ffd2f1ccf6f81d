#!/usr/bin/env python
"""Times every compiled staging variant (chunk bytes x stages) per digest, device-resident.
Usage: python tools/sweep.py [--n 65536 --len 262144] [--algs sha1,md5,sha256]  -> one JSON line per cell."""
import argparse
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from downloader_b200 import _capi

VARIANTS = [(64, 4), (128, 2), (128, 3), (128, 4), (256, 2), (256, 3), (512, 2)]
ALG = {"sha1": 1, "md5": 2, "sha256": 4}
DL = {"sha1": 20, "md5": 16, "sha256": 32}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=65536)
    ap.add_argument("--len", type=int, default=262144)
    ap.add_argument("--algs", default="sha1,md5,sha256")
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--variants", default="")
    ap.add_argument("--path", type=int, default=1, help="1 bulk (variants = CHxNST), 2 tile, 3 sliced (variants = KxNST)")
    ap.add_argument("--workers", default="12")
    ap.add_argument("--rounds", default="32")
    ap.add_argument("--opt", default="", help="extra th_set_option pairs, e.g. assist_variant=2")
    a = ap.parse_args()
    L = _capi.lib()
    n, ln = a.n, a.len
    data = torch.empty(n * ln, dtype=torch.uint8, device="cuda")
    L.th_synth_fill(data.data_ptr(), 0, n, ln, ln, 1, None)
    TILINGS = [(1, 2), (1, 3), (1, 4), (2, 2), (2, 3), (4, 2)]
    variants = (VARIANTS if a.path == 1 else TILINGS) if not a.variants else [tuple(map(int, v.split("x"))) for v in a.variants.split(",")]
    assert L.th_set_option(b"path", a.path) == 0
    for kv in filter(None, a.opt.split(",")):
        k, v = kv.split("=")
        assert L.th_set_option(k.encode(), int(v)) == 0, kv
    combos = [(int(w), int(r)) for w in a.workers.split(",") for r in a.rounds.split(",")] if a.path == 3 else [(12, 32)]
    for alg in a.algs.split(","):
        out = torch.zeros(n * DL[alg], dtype=torch.uint8, device="cuda")
        fn = {"sha1": lambda: L.th_sha1_batch(data.data_ptr(), None, None, None, ln, ln, n, out.data_ptr(), None, None, None),
              "md5": lambda: L.th_md5_batch(data.data_ptr(), None, None, None, ln, ln, n, out.data_ptr(), None),
              "sha256": lambda: L.th_sha256_batch(data.data_ptr(), None, None, None, ln, ln, n, out.data_ptr(), None)}[alg]
        for (ch, st), (wk, rd) in [(v, c) for v in variants for c in combos]:
            if a.path == 1:
                assert L.th_set_staging(ALG[alg], ch, st) == 0
            else:
                assert L.th_set_tiling(ALG[alg], ch, st) == 0
                assert L.th_set_option(b"workers", wk) == 0 and L.th_set_option(b"rounds", rd) == 0
            for _ in range(2):
                assert fn() == 0
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.iters):
                fn()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / a.iters
            print(json.dumps({"alg": alg, "path": a.path, "n": n, "len": ln, "chunk_or_k": ch, "stages": st, "workers": wk, "rounds": rd, "ms": round(ms, 3),
                              "GBps": round(n * ln / ms / 1e6, 1)}), flush=True)


if __name__ == "__main__":
    main()
