"""N>1 path on CPU: two gloo ranks shard messages by index, hash their shards (hashlib stands in for the
GPU here — this test is about the host-side partition / gather logic, not the kernels), rank 0 gathers
the digests by index and checks the whole set against the oracle."""
import hashlib
import os
import socket
import subprocess
import sys

import numpy as np

from downloader_b200.sharding import place_by_index, shard_by_bytes, shard_range

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import hashlib, os, sys
import numpy as np
import torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
import oracle
from downloader_b200.sharding import shard_range, shard_by_bytes, place_by_index
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
# equal-length messages: contiguous ranges (configs 2/3)
n, ln = 37, 4096 + 17
lo, hi = shard_range(n, world, rank)
mine = b"".join(hashlib.sha1(oracle.synth(i, ln).tobytes()).digest() for i in range(lo, hi))
t = torch.frombuffer(bytearray(mine.ljust(20 * ((n + world - 1) // world), b"\0")), dtype=torch.uint8)
gathered = [torch.empty_like(t) for _ in range(world)] if rank == 0 else None
dist.gather(t, gathered, dst=0)
# mixed lengths: byte-balanced assignment (config 4)
lens = [((i * 7919) % 50000) + (i % 3 == 0) * 200000 for i in range(53)]
assign = shard_by_bytes(lens, world)
mine2 = b"".join(hashlib.md5(oracle.synth(i, lens[i]).tobytes()).digest() for i in assign[rank])
t2 = torch.frombuffer(bytearray(mine2.ljust(16 * max(len(a) for a in assign), b"\0")), dtype=torch.uint8)
g2 = [torch.empty_like(t2) for _ in range(world)] if rank == 0 else None
dist.gather(t2, g2, dst=0)
if rank == 0:
    parts = []
    for r in range(world):
        a, b = shard_range(n, world, r)
        parts.append((list(range(a, b)), bytes(gathered[r].numpy().tobytes()[:20 * (b - a)])))
    got = place_by_index(n, 20, parts)
    flat = np.concatenate([oracle.synth(i, ln) for i in range(n)])
    exp = oracle.batch(oracle.SHA1, flat, np.arange(n, dtype=np.uint64) * ln, np.full(n, ln, np.uint64))
    assert bytes(got) == exp.tobytes(), "sharded sha1 gather differs from the oracle"
    parts2 = [(assign[r], bytes(g2[r].numpy().tobytes()[:16 * len(assign[r])])) for r in range(world)]
    got2 = place_by_index(len(lens), 16, parts2)
    exp2 = b"".join(oracle.digest(oracle.MD5, oracle.synth(i, lens[i])) for i in range(len(lens)))
    assert bytes(got2) == exp2, "byte-balanced md5 gather differs from the oracle"
    loads = [sum(lens[i] for i in a) for a in assign]
    assert max(loads) - min(loads) <= max(lens)
    print("OK")
dist.barrier()
dist.destroy_process_group()
'''


def test_shard_range_and_balance():
    for n in (0, 1, 7, 65536, 4096):
        for w in (1, 2, 4, 8):
            rs = [shard_range(n, w, r) for r in range(w)]
            assert rs[0][0] == 0 and rs[-1][1] == n
            assert all(rs[i][1] == rs[i + 1][0] for i in range(w - 1))
            assert max(b - a for a, b in rs) - min(b - a for a, b in rs) <= 1
    lens = [10, 10, 10, 1000, 5, 995]
    a = shard_by_bytes(lens, 2)
    assert sorted(a[0] + a[1]) == list(range(6))
    assert abs(sum(lens[i] for i in a[0]) - sum(lens[i] for i in a[1])) <= 30
    assert bytes(place_by_index(3, 2, [([2, 0], b"ccaa"), ([1], b"bb")])) == b"aabbcc"


def test_two_rank_gloo_shard_and_gather(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), str(script), ROOT]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "OK" in r.stdout
