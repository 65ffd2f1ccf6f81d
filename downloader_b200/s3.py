"""S3 multipart framing around the part digests.

Mirrors what [dep] minio@7.0.12 does around `getHashSummer` inside `fPutObject` / `putObject`
(reference call sites lib/upload.js:45 and :55): part planning (`calculatePartSize`, [start,end] ranges),
the header encodings — `Content-MD5` = base64(md5) over TLS, `x-amz-content-sha256` = lower-case hex when
`enableSHA256 = !anonymous && !useSSL` — the resume rule (a listed part whose ETag equals hex(md5) is
skipped) and the object ETag `hex(md5(md5_1 || ... || md5_N)) + "-" + N`.  The digests themselves come
from the GPU (Engine); nothing here hashes on the CPU.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import _capi

MiB = 1 << 20
DEFAULT_PART_SIZE = 64 * MiB          # minio-js default partSize
MAX_PARTS = 10000
MIN_PART_SIZE = 5 * MiB               # S3 limit (all parts but the last)
MAX_PART_SIZE = 5 * 1024 * MiB        # 5 GiB
MAX_OBJECT_SIZE = 5 * 1024 * 1024 * MiB  # 5 TiB


def calculate_part_size(size: int, part_size: int = DEFAULT_PART_SIZE, override: bool = False) -> int:
    """minio-js calculatePartSize: `part_size` (64 MiB) then +16 MiB steps until part_size*10000 > size.
    override=True is minio's `overRidePartSize` (BASELINE configs use a 16 MiB override)."""
    if size < 0 or size > MAX_OBJECT_SIZE:
        raise ValueError(f"size should not be more than {MAX_OBJECT_SIZE}")
    if override:
        return part_size
    p = part_size
    while p * MAX_PARTS <= size:
        p += 16 * MiB
    return p


def part_ranges(size: int, part_size: int) -> List[Tuple[int, int]]:
    """[start, end) of every part; the last one is short.  A zero-length object has one empty part
    (minio sends a single PUT for size <= partSize)."""
    if size == 0:
        return [(0, 0)]
    return [(s, min(s + part_size, size)) for s in range(0, size, part_size)]


def content_md5(md5: bytes) -> str:
    """`Content-MD5` header value: base64 of the 16 raw bytes (getHashSummer's md5sum)."""
    return _capi.base64_(bytes(md5))


def sha256_hex(sha256: bytes) -> str:
    """`x-amz-content-sha256` header value: 64 lower-case hex characters (getHashSummer's sha256sum)."""
    return _capi.hex_(bytes(sha256))


def md5_hex(md5: bytes) -> str:
    """Part ETag as S3 returns it (unquoted) — what minio's resume path compares against."""
    return _capi.hex_(bytes(md5))


def select_digests(anonymous: bool, use_ssl: bool) -> int:
    """Which digest the connection mode needs (minio 7.0.x computes only that one):
    enableSHA256 = !anonymous && !useSSL -> SHA-256, else MD5."""
    return _capi.ALG_SHA256 if (not anonymous and not use_ssl) else _capi.ALG_MD5


@dataclass
class PartPlan:
    size: int
    part_size: int
    ranges: List[Tuple[int, int]]

    @property
    def multipart(self) -> bool:
        return self.size > self.part_size


def plan_object(size: int, part_size: int = DEFAULT_PART_SIZE, override: bool = False) -> PartPlan:
    ps = calculate_part_size(size, part_size, override)
    if size <= ps:
        return PartPlan(size, ps, [(0, size)])
    return PartPlan(size, ps, part_ranges(size, ps))


def multipart_etag(engine, md5s, counts: Optional[Sequence[int]] = None) -> List[str]:
    """Object ETag(s) from part MD5s on the GPU (th_etag_combine): hex(md5(concat)) + '-' + N."""
    m = np.ascontiguousarray(md5s, np.uint8).reshape(-1, 16)
    cnt = [m.shape[0]] if counts is None else list(counts)
    out = engine.etag_combine(m, cnt)
    return [_capi.hex_(out[i].tobytes()) + "-" + str(c) for i, c in enumerate(cnt)]


def parts_to_skip(md5s, listed_etags: dict) -> List[int]:
    """minio's resume check: part numbers (1-based) whose previously listed ETag equals hex(md5)."""
    m = np.ascontiguousarray(md5s, np.uint8).reshape(-1, 16)
    return [i + 1 for i in range(m.shape[0]) if listed_etags.get(i + 1, "").strip('"') == md5_hex(m[i].tobytes())]
