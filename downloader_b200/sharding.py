"""Index sharding of pieces / parts across GPUs: no exchange on the data path, only the
16..32-byte digests are gathered on the host.  Pure host logic; used by bench.py and the stages."""
from __future__ import annotations

from typing import List, Sequence, Tuple


def shard_range(n: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous index range [lo, hi) of rank `rank` for n equal-length messages (configs 2, 3)."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    return (n * rank) // world, (n * (rank + 1)) // world


def shard_by_bytes(lengths: Sequence[int], world: int) -> List[List[int]]:
    """Greedy longest-first assignment of mixed-length messages so every GPU gets about the same number
    of BYTES (config 4).  Returns, per rank, the message indices it hashes (ascending)."""
    if world <= 0:
        raise ValueError("bad world")
    load = [0] * world
    out: List[List[int]] = [[] for _ in range(world)]
    for i in sorted(range(len(lengths)), key=lambda k: (-lengths[k], k)):
        r = min(range(world), key=lambda k: (load[k], k))
        out[r].append(i)
        load[r] += lengths[i]
    for lst in out:
        lst.sort()
    return out


def place_by_index(n: int, dlen: int, parts) -> bytearray:
    """Host-side gather: parts = iterable of (indices, digests bytes) per rank -> n*dlen bytes by index."""
    buf = bytearray(n * dlen)
    for idx, dig in parts:
        assert len(dig) == len(idx) * dlen
        for k, i in enumerate(idx):
            buf[i * dlen:(i + 1) * dlen] = dig[k * dlen:(k + 1) * dlen]
    return buf
