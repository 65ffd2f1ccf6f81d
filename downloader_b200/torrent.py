""".torrent metainfo -> expected-digest table, piece geometry and info-hash.

What [dep] parse-torrent@7.0.0 / create-torrent@4.0.0 give webtorrent (reference call site
lib/download.js:64-66, and lib/download.js:144-155 for `.torrent` URLs): `info.pieces` split into 20-byte
SHA-1 digests (webtorrent's `_hashes`), piece length, the file list laid end to end, the short last piece,
and infoHash = sha1(bencode(info)).  Hashing (info-hash, piece digests when creating) runs on the GPU.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Any, List, Sequence, Tuple

import numpy as np

from ._capi import ALG_SHA1, hex_


# ---- bencode (BEP 3) ---------------------------------------------------------------------------
def bencode(x: Any) -> bytes:
    if isinstance(x, bool):
        raise TypeError("bool is not bencodable")
    if isinstance(x, int):
        return b"i%de" % x
    if isinstance(x, str):
        x = x.encode()
    if isinstance(x, (bytes, bytearray, memoryview)):
        b = bytes(x)
        return b"%d:" % len(b) + b
    if isinstance(x, (list, tuple)):
        return b"l" + b"".join(bencode(v) for v in x) + b"e"
    if isinstance(x, dict):
        items = sorted(((k.encode() if isinstance(k, str) else bytes(k)), v) for k, v in x.items())
        return b"d" + b"".join(bencode(k) + bencode(v) for k, v in items) + b"e"
    raise TypeError(f"cannot bencode {type(x).__name__}")


def _bdecode_at(data: bytes, i: int, spans: dict, path: tuple):
    c = data[i:i + 1]
    if c == b"i":
        j = data.index(b"e", i)
        return int(data[i + 1:j]), j + 1
    if c == b"l":
        out, i = [], i + 1
        while data[i:i + 1] != b"e":
            v, i = _bdecode_at(data, i, spans, path + (len(out),))
            out.append(v)
        return out, i + 1
    if c == b"d":
        out, i = {}, i + 1
        while data[i:i + 1] != b"e":
            k, i = _bdecode_at(data, i, spans, path)
            s = i
            v, i = _bdecode_at(data, i, spans, path + (k,))
            out[k] = v
            spans[path + (k,)] = (s, i)
        return out, i + 1
    if c.isdigit():
        j = data.index(b":", i)
        n = int(data[i:j])
        if j + 1 + n > len(data):
            raise ValueError("truncated bencoded string")
        return data[j + 1:j + 1 + n], j + 1 + n
    raise ValueError(f"bad bencode at offset {i}")


def bdecode(data: bytes, spans: dict | None = None):
    """Decodes one bencoded value; dictionary keys stay bytes.  `spans` (optional) receives the raw byte
    span of every dictionary value, keyed by its key path — used to hash `info` exactly as transmitted."""
    v, end = _bdecode_at(bytes(data), 0, spans if spans is not None else {}, ())
    if end != len(data):
        raise ValueError("trailing bytes after bencoded value")
    return v


# ---- metainfo --------------------------------------------------------------------------------------
@dataclass
class Metainfo:
    name: str
    piece_length: int
    pieces: np.ndarray              # (n, 20) uint8 — webtorrent's _hashes, raw
    files: List[Tuple[str, int]]    # (relative path, length), in torrent order
    length: int
    info_hash: str                  # 40 lower-case hex characters

    @property
    def num_pieces(self) -> int:
        return self.pieces.shape[0]

    @property
    def last_piece_length(self) -> int:
        return self.length % self.piece_length or self.piece_length

    def piece_hex(self, i: int) -> str:
        return hex_(self.pieces[i].tobytes())

    def on_disk(self, root: str) -> List[Tuple[str, int]]:
        """Absolute paths under the download directory ([dep] fs-chunk-store layout: <path>/<name>/...)."""
        return [(os.path.join(root, p), n) for p, n in self.files]


def parse_metainfo(data: bytes, engine) -> Metainfo:
    spans: dict = {}
    top = bdecode(data, spans)
    if not isinstance(top, dict) or b"info" not in top:
        raise ValueError("Torrent is missing required field: info")
    info = top[b"info"]
    for k in (b"name", b"piece length", b"pieces"):
        if k not in info:
            raise ValueError(f"Torrent is missing required field: info.{k.decode()}")
    raw = info[b"pieces"]
    if len(raw) % 20:
        raise ValueError("info.pieces is not a multiple of 20 bytes")
    name = info[b"name"].decode("utf-8", "replace")
    if b"files" in info:
        files = [(os.path.join(name, *[c.decode("utf-8", "replace") for c in f[b"path"]]), int(f[b"length"])) for f in info[b"files"]]
    else:
        files = [(name, int(info[b"length"]))]
    length = sum(n for _, n in files)
    plen = int(info[b"piece length"])
    pieces = np.frombuffer(raw, np.uint8).reshape(-1, 20).copy()
    if pieces.shape[0] != (length + plen - 1) // plen:
        raise ValueError("piece count does not match the file lengths")
    s, e = spans[(b"info",)]
    ih = engine.hash_batch([bytes(data[s:e])], ALG_SHA1)[ALG_SHA1][0].tobytes()  # infoHash = sha1(bencode(info))
    return Metainfo(name, plen, pieces, files, length, hex_(ih))


def create_metainfo(engine, root: str, rel_files: Sequence[str], piece_length: int, name: str) -> bytes:
    """Builds a .torrent for files under root/name (piece digests on the GPU) — the create-torrent side,
    used by tests and tools to make fixtures."""
    from .reader import _ConcatFiles, _stream_messages, piece_bounds
    files = [(os.path.join(root, name, p), os.path.getsize(os.path.join(root, name, p))) for p in rel_files]
    src = _ConcatFiles(files)
    try:
        res = _stream_messages(engine, src, piece_bounds(src.total, piece_length), ALG_SHA1, None, 0, 64 << 20, 2)
    finally:
        src.close()
    info: dict = {"name": name, "piece length": piece_length, "pieces": res[ALG_SHA1].tobytes()}
    if len(rel_files) == 1 and rel_files[0] == "":
        info["length"] = files[0][1]
    else:
        info["files"] = [{"length": n, "path": p.split("/")} for p, (_, n) in zip(rel_files, files)]
    return bencode({"announce": "", "info": info})
