#!/usr/bin/env python
"""Small, fast pass over every kernel path (bulk, tile, sliced with slicing engaged, assist, etag, static ragged,
synth) for compute-sanitizer runs:  compute-sanitizer --tool memcheck python tools/sanitize_case.py"""
import ctypes as C
import hashlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from downloader_b200 import Engine, _capi

L = _capi.lib()
NAMES = {1: "sha1", 2: "md5", 4: "sha256"}
DL = {1: 20, 2: 16, 4: 32}


def check_uniform(path, n, ln, opts=()):
    assert L.th_set_option(b"path", path) == 0
    for k, v in opts:
        assert L.th_set_option(k, v) == 0
    stride = (ln + 15) & ~15
    d = torch.zeros(n * stride, dtype=torch.uint8, device="cuda")
    assert L.th_synth_fill(d.data_ptr(), 0, n, ln, stride, 5, None) == 0
    host = None
    for a in (1, 2, 4):
        o = torch.zeros(n * DL[a], dtype=torch.uint8, device="cuda")
        fn = {1: lambda: L.th_sha1_batch(d.data_ptr(), None, None, None, stride, ln, n, o.data_ptr(), None, None, None),
              2: lambda: L.th_md5_batch(d.data_ptr(), None, None, None, stride, ln, n, o.data_ptr(), None),
              4: lambda: L.th_sha256_batch(d.data_ptr(), None, None, None, stride, ln, n, o.data_ptr(), None)}[a]
        assert fn() == 0
        torch.cuda.synchronize()
        if host is None:
            host = d.cpu().numpy()
        got = o.cpu().numpy().reshape(n, -1)
        for i in (0, n // 2, n - 1):
            assert got[i].tobytes() == hashlib.new(NAMES[a], host[i * stride:i * stride + ln]).digest(), (path, a, i)
    L.th_set_option(b"path", 0)


torch.cuda.set_device(0)
check_uniform(1, 70, 1024 + 57)
check_uniform(2, 70, 2048 + 56)
check_uniform(3, 148 * 14 * 32 - 5, 1024 + 64, [(b"workers", 12), (b"rounds", 4)])   # time slicing engaged
check_uniform(4, 40, 4096 + 63)
with Engine([0], ring_bytes=8 << 20) as e:
    bufs = [bytes([i]) * n for i, n in enumerate([0, 1, 55, 56, 64, 65, 1000, 70000])]
    exp = np.stack([np.frombuffer(hashlib.sha1(b).digest(), np.uint8) for b in bufs])
    r = e.hash_batch(bufs, 7, expected_sha1=exp)
    assert r["ok"].all() and (r[2][3].tobytes() == hashlib.md5(bufs[3]).digest())
    assert e.etag_combine(r[2], [len(bufs)])[0].tobytes() == hashlib.md5(r[2].tobytes()).digest()
    lens = [5000, 3, 0, 70000, 64]
    offs, pos = [], 0
    for n in lens:
        offs.append(pos)
        pos = (pos + n + 15) & ~15
    data = torch.arange(pos, dtype=torch.int64, device="cuda").to(torch.uint8)
    d_off = torch.tensor(offs, dtype=torch.int64, device="cuda")
    d_len = torch.tensor(lens, dtype=torch.int64, device="cuda")
    o = torch.zeros(len(lens) * 32, dtype=torch.uint8, device="cuda")
    assert L.th_digest_ragged_static(4, data.data_ptr(), d_off.data_ptr(), d_len.data_ptr(), len(lens), o.data_ptr(), None, None, None) == 0
    torch.cuda.synchronize()
    h = data.cpu().numpy()
    for i, (a, n) in enumerate(zip(offs, lens)):
        assert o.cpu().numpy().reshape(-1, 32)[i].tobytes() == hashlib.sha256(h[a:a + n]).digest()
print("sanitize_case ok")
