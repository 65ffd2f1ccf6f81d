"""File -> pinned-ring streaming reader.

Reads large sequential ranges of the staged media files straight into pinned ring slots (readinto on the
slot's buffer: no intermediate copy), keeps several slots in flight so the H2D copy of slot k+1 overlaps the
kernels of slot k, and cuts the same byte stream either into torrent pieces (SHA-1 + verify bitfield,
reference path lib/download.js:64-66 -> webtorrent `_verifyPieces`) or into S3 parts (MD5 / SHA-256,
reference path lib/upload.js:45 -> minio `fPutObject`).  Every digest comes from the GPU.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from ._capi import ALG_MD5, ALG_SHA1, ALG_SHA256, DIGEST_LEN


READ_CHUNK = 16 << 20
READ_THREADS = max(1, min(16, int(os.environ.get("TRITONHASH_READ_THREADS", "0")) or (os.cpu_count() or 1)))
DIRECT_ALIGN = 4096   # O_DIRECT wants memory address, file offset and length aligned to the logical block size
_POOLS: Dict[int, object] = {}


def _pool(gpu: int = 0):
    """One reader pool per GPU ring: with several GPUs every ring has its own readers, so a feeder that is waiting
    for a slow read does not hold up the reads that fill another GPU's slots."""
    pool = _POOLS.get(gpu)
    if pool is None:
        from concurrent.futures import ThreadPoolExecutor
        pool = _POOLS[gpu] = ThreadPoolExecutor(max_workers=READ_THREADS, thread_name_prefix=f"tritonhash-read-gpu{gpu}")
    return pool


def _align(x: int, a: int) -> int:
    return (x + a - 1) // a * a


def _direct_default() -> bool:
    return os.environ.get("TRITONHASH_O_DIRECT", "0") == "1"


class _ConcatFiles:
    """The torrent's view of a download directory: files laid end to end ([dep] fs-chunk-store).

    direct=True reads with O_DIRECT where it can: a staged media file is read once, front to back, and is usually
    larger than what is worth keeping in the page cache — buffered reads copy every byte twice (disk -> page cache ->
    pinned slot) and evict everything else.  With O_DIRECT the device DMA lands in the pinned ring slot itself.  The
    aligned body of each chunk (4 KiB-aligned slot address, file offset and length; ring slots are 4 KiB-aligned)
    goes through the O_DIRECT descriptor, the unaligned head / tail bytes and file systems that refuse O_DIRECT
    (tmpfs) through the ordinary one."""

    def __init__(self, files: Sequence[Tuple[str, int]], direct: Optional[bool] = None):
        self.files = [(p, int(n)) for p, n in files]
        self.starts = np.cumsum([0] + [n for _, n in self.files]).tolist()
        self.total = self.starts[-1]
        self.direct = _direct_default() if direct is None else bool(direct)
        self._fds: Dict[int, int] = {}
        self._dfds: Dict[int, int] = {}
        self.direct_bytes = 0      # bytes that came through O_DIRECT descriptors
        self.buffered_bytes = 0    # bytes that came through the page cache

    def _fd(self, i: int) -> int:
        fd = self._fds.get(i)
        if fd is None:
            try:
                fd = os.open(self.files[i][0], os.O_RDONLY)
            except OSError:
                fd = -1
            self._fds[i] = fd
        return fd

    def _dfd(self, i: int) -> int:
        if not self.direct:
            return -1
        fd = self._dfds.get(i)
        if fd is None:
            try:
                fd = os.open(self.files[i][0], os.O_RDONLY | os.O_DIRECT)
            except (OSError, AttributeError):  # tmpfs and some network file systems refuse O_DIRECT
                fd = -1
            self._dfds[i] = fd
        return fd

    @staticmethod
    def _pread_all(fd: int, mv: memoryview, off: int) -> int:
        got = 0
        while got < len(mv):
            r = os.preadv(fd, [mv[got:]], off + got)
            if r <= 0:
                break
            got += r
        return got

    def _pread_chunk(self, fd: int, dfd: int, view: np.ndarray, off: int) -> Tuple[int, int]:
        """preadv straight into the (pinned) view; what the file does not supply reads as zeros.
        Returns (bytes read with O_DIRECT, bytes read through the page cache)."""
        take, got, direct = view.size, 0, 0
        if fd >= 0:
            mv = memoryview(view)
            body = take // DIRECT_ALIGN * DIRECT_ALIGN
            if dfd >= 0 and body and off % DIRECT_ALIGN == 0 and view.ctypes.data % DIRECT_ALIGN == 0:
                try:
                    direct = self._pread_all(dfd, mv[:body], off)  # short only at the end of the file
                except OSError:
                    direct = 0                                      # e.g. EINVAL from a stacked file system
                got = direct
            if got < take and (direct == 0 or got == body):
                got += self._pread_all(fd, mv[got:], off + got)
        if got < take:
            view[got:] = 0
        return direct, got - direct

    def readinto(self, view: np.ndarray, pos: int, gpu: int = 0) -> None:
        """Fill `view` with stream bytes [pos, pos+len(view)); missing / short files read as zeros.
        Large ranges are cut into READ_CHUNK pieces read by that GPU's thread pool (preadv releases the GIL): one
        thread copies from the page cache at a few GB/s, an order of magnitude below what the H2D link takes."""
        n = view.size
        done = 0
        jobs = []
        i = max(0, int(np.searchsorted(self.starts, pos, side="right")) - 1)
        while done < n and i < len(self.files):
            flen = self.files[i][1]
            off = pos + done - self.starts[i]
            take = min(n - done, flen - off)
            if take > 0:
                fd, dfd = self._fd(i), self._dfd(i)
                for c in range(0, take, READ_CHUNK):
                    m = min(READ_CHUNK, take - c)
                    jobs.append((fd, dfd, view[done + c:done + c + m], off + c))
                done += take
            i += 1
        if done < n:
            view[done:] = 0
        if len(jobs) > 1 and READ_THREADS > 1:
            counts = [f.result() for f in [_pool(gpu).submit(self._pread_chunk, *j) for j in jobs]]
        else:
            counts = [self._pread_chunk(*j) for j in jobs]
        self.direct_bytes += sum(c[0] for c in counts)
        self.buffered_bytes += sum(c[1] for c in counts)

    def close(self) -> None:
        for fd in list(self._fds.values()) + list(self._dfds.values()):
            if fd >= 0:
                os.close(fd)
        self._fds.clear()
        self._dfds.clear()


def _slots_in_flight(engine, batch_bytes: int, in_flight: Optional[int]) -> int:
    """How many submitted slots to leave uncollected: as many as the device staging ring holds (a chain over a
    large part runs for much longer than its slot's copy; th_ring_acquire does the waiting when a ring fills)."""
    if in_flight is not None:
        return max(1, in_flight)
    return max(2, int(getattr(engine, "device_bytes", 0)) // max(batch_bytes, 1) - 1)


def _stream_messages(engine, src: _ConcatFiles, bounds: Sequence[Tuple[int, int]], algs: int,
                     expected_sha1: Optional[np.ndarray], gpu: int, batch_bytes: int, in_flight: int) -> dict:
    """Hashes messages given as [start,end) ranges of the concatenated stream (ascending, non-overlapping)."""
    n = len(bounds)
    out = {a: np.zeros((n, DIGEST_LEN[a]), np.uint8) for a in (ALG_SHA1, ALG_MD5, ALG_SHA256) if algs & a}
    ok = np.zeros(n, bool) if expected_sha1 is not None else None
    pending: List[Tuple[int, int, int]] = []  # (ticket, first message, count)
    in_flight = _slots_in_flight(engine, batch_bytes, in_flight)

    def drain(k: int) -> None:
        while len(pending) > k:
            t, first, cnt = pending.pop(0)
            r = engine.result(t, cnt, algs, want_ok=ok is not None)
            for a in out:
                out[a][first:first + cnt] = r[a]
            if ok is not None:
                ok[first:first + cnt] = r["ok"]

    i = 0
    while i < n:
        # a batch = consecutive messages whose 16-byte-aligned images fit in one slot
        offs, lens, pos = [], [], 0
        j = i
        while j < n:
            ln = bounds[j][1] - bounds[j][0]
            if offs and pos + ln > batch_bytes:
                break
            offs.append(pos)
            lens.append(ln)
            pos = _align(pos + ln, 16)
            j += 1
        drain(in_flight - 1)
        slot, view = engine.ring_acquire(max(pos, 16), gpu)
        contiguous = all(bounds[k + 1][0] == bounds[k][1] and (bounds[k][1] - bounds[k][0]) % 16 == 0 for k in range(i, j - 1))
        try:  # an acquired slot that is neither submitted nor released would block the FIFO ring for good
            if contiguous:  # pieces / parts back to back: one sequential read for the whole slot
                src.readinto(view[:bounds[j - 1][1] - bounds[i][0]], bounds[i][0], gpu)
            else:
                for k in range(i, j):
                    src.readinto(view[offs[k - i]:offs[k - i] + lens[k - i]], bounds[k][0], gpu)
            exp = expected_sha1[i:j] if expected_sha1 is not None else None
            ticket = engine.submit(slot, np.array(offs, np.uint64), np.array(lens, np.uint64), algs, exp)
        except BaseException:
            engine.ring_release(slot)
            raise
        pending.append((ticket, i, j - i))
        i = j
    drain(0)
    if ok is not None:
        out["ok"] = ok
    return out


def hash_file_parts(engine, path: str, ranges: Sequence[Tuple[int, int]], algs: int = ALG_MD5 | ALG_SHA256, gpu: int = 0,
                    batch_bytes: int = 256 << 20, in_flight: Optional[int] = None, direct: Optional[bool] = None) -> dict:
    """Per-part digests of one file (minio's first pass over every part, all parts in a few launches).
    direct: read with O_DIRECT (default: TRITONHASH_O_DIRECT=1), see _ConcatFiles."""
    size = os.path.getsize(path)
    src = _ConcatFiles([(path, size)], direct)
    try:
        return _stream_messages(engine, src, list(ranges), algs, None, gpu, batch_bytes, in_flight)
    finally:
        src.close()


def piece_bounds(total: int, piece_length: int) -> List[Tuple[int, int]]:
    """Piece i = bytes [i*pieceLength, min((i+1)*pieceLength, total)) of the concatenated file list."""
    return [(s, min(s + piece_length, total)) for s in range(0, total, piece_length)]


def verify_pieces(engine, files: Sequence[Tuple[str, int]], piece_length: int, expected_sha1, gpu: int = 0,
                  batch_bytes: int = 256 << 20, in_flight: Optional[int] = None, indices: Optional[Sequence[int]] = None,
                  direct: Optional[bool] = None) -> dict:
    """Bulk re-verify of an on-disk download ([dep] webtorrent `_verifyPieces`): returns {'ok': bool[n]
    (the bitfield), ALG_SHA1: digests}.  Missing files hash as zeros and simply fail verification.

    indices: verify only these pieces (ascending) — webtorrent's per-piece `onChunk` shape (call site
    lib/download.js:64-66): after a refetch only the pieces that were requested again are read, copied and hashed,
    not the whole download.  expected_sha1 stays the full info.pieces table; 'ok' and the digests then have one row
    per requested index, in that order ('indices' echoes them)."""
    src = _ConcatFiles(files, direct)
    bounds = piece_bounds(src.total, piece_length)
    exp = np.ascontiguousarray(expected_sha1, np.uint8).reshape(-1, 20)
    if exp.shape[0] != len(bounds):
        raise ValueError(f"expected {len(bounds)} piece digests, got {exp.shape[0]}")
    idx = None
    if indices is not None:
        idx = np.asarray(list(indices), np.int64)
        if idx.size and (idx.min() < 0 or idx.max() >= len(bounds) or (np.diff(idx) <= 0).any()):
            raise ValueError("piece indices must be ascending, unique and within the torrent")
        bounds = [bounds[int(i)] for i in idx]
        exp = exp[idx]
    try:
        if not bounds:
            return {ALG_SHA1: np.zeros((0, 20), np.uint8), "ok": np.zeros(0, bool), "indices": idx}
        out = _stream_messages(engine, src, bounds, ALG_SHA1, exp, gpu, batch_bytes, in_flight)
        if idx is not None:
            out["indices"] = idx
        return out
    finally:
        src.close()


def hash_media_file(engine, path: str, piece_length: int, expected_sha1, part_ranges: Sequence[Tuple[int, int]],
                    part_algs: int = ALG_MD5 | ALG_SHA256, gpu: int = 0, batch_bytes: int = 512 << 20, in_flight: Optional[int] = None,
                    direct: Optional[bool] = None) -> dict:
    """One pass over a staged media file feeding BOTH consumers of its bytes: SHA-1 by torrent piece (with
    the fused verify) and MD5 / SHA-256 by S3 part.  Each ring slot is read once, crosses PCIe once, and is
    submitted twice (th_submit allows that).  Slots hold whole parts, and parts are multiples of the piece
    length, so both message lists tile every slot.  Returns {'pieces': {...}, 'parts': {...}}."""
    size = os.path.getsize(path)
    part_ranges = list(part_ranges)
    pb = piece_bounds(size, piece_length)
    exp = None if expected_sha1 is None else np.ascontiguousarray(expected_sha1, np.uint8).reshape(-1, 20)
    pieces = {ALG_SHA1: np.zeros((len(pb), 20), np.uint8)}
    if exp is not None:
        pieces["ok"] = np.zeros(len(pb), bool)
    parts = {a: np.zeros((len(part_ranges), DIGEST_LEN[a]), np.uint8) for a in (ALG_MD5, ALG_SHA256) if part_algs & a}
    for s0, e0 in part_ranges:  # checked before any slot is taken: a slot holds whole pieces and whole parts
        if s0 % piece_length or (e0 % piece_length and e0 != size):
            raise ValueError("part boundaries must be multiples of the piece length")
    src = _ConcatFiles([(path, size)], direct)
    pending: List[tuple] = []
    in_flight = _slots_in_flight(engine, batch_bytes, in_flight)

    def drain(k):
        while len(pending) > k:
            tp, p0, pn, tq, q0, qn = pending.pop(0)
            r = engine.result(tp, pn, ALG_SHA1, want_ok=exp is not None)
            pieces[ALG_SHA1][p0:p0 + pn] = r[ALG_SHA1]
            if exp is not None:
                pieces["ok"][p0:p0 + pn] = r["ok"]
            r = engine.result(tq, qn, part_algs)
            for a in parts:
                parts[a][q0:q0 + qn] = r[a]

    try:
        q = 0
        while q < len(part_ranges):
            start = part_ranges[q][0]
            q1 = q
            while q1 < len(part_ranges) and part_ranges[q1][1] - start <= batch_bytes:
                q1 += 1
            q1 = max(q1, q + 1)
            end = part_ranges[q1 - 1][1]
            drain(in_flight - 1)
            slot, view = engine.ring_acquire(max(end - start, 16), gpu)
            tp = None
            try:
                src.readinto(view[:end - start], start, gpu)
                p0, p1 = start // piece_length, -(-end // piece_length)
                offs = np.array([pb[i][0] - start for i in range(p0, p1)], np.uint64)
                lens = np.array([pb[i][1] - pb[i][0] for i in range(p0, p1)], np.uint64)
                tp = engine.submit(slot, offs, lens, ALG_SHA1, exp[p0:p1] if exp is not None else None)
                offs = np.array([part_ranges[i][0] - start for i in range(q, q1)], np.uint64)
                lens = np.array([part_ranges[i][1] - part_ranges[i][0] for i in range(q, q1)], np.uint64)
                tq = engine.submit(slot, offs, lens, part_algs)   # same slot, second message list: no second copy
            except BaseException:
                if tp is None:
                    engine.ring_release(slot)  # an unsubmitted slot would block the ring for good
                else:
                    engine.result(tp, p1 - p0, ALG_SHA1)  # submitted once: retire that ticket, the slot then frees itself
                raise
            pending.append((tp, p0, p1 - p0, tq, q, q1 - q))
            q = q1
        drain(0)
    finally:
        src.close()
    return {"pieces": pieces, "parts": parts, "read": {"direct_bytes": src.direct_bytes, "buffered_bytes": src.buffered_bytes}}


def hash_media_files(engine, jobs: Sequence[dict], gpus: Optional[Sequence[int]] = None) -> List[dict]:
    """Several staged media files over several GPUs of ONE context — the files loop of lib/upload.js:34-52, which the
    reference runs strictly one file after the other.  Files are dealt to the GPUs by bytes (sharding.shard_by_bytes);
    every GPU has its own feeder thread, its own ring and its own reader pool, so the GPUs' reads, copies and chains
    overlap and no feeder waits on another GPU's ring (the library takes one lock per GPU).
    jobs: dicts of hash_media_file's arguments (path, piece_length, expected_sha1, part_ranges, ...).
    Returns hash_media_file's results in the order of `jobs`, each with the 'gpu' that hashed it."""
    import threading

    from .sharding import shard_by_bytes
    gpus = list(gpus) if gpus is not None else list(range(len(getattr(engine, "gpu_ids", [0]))))
    sizes = [os.path.getsize(j["path"]) for j in jobs]
    shards = shard_by_bytes(sizes, len(gpus))
    out: List[Optional[dict]] = [None] * len(jobs)
    errors: List[BaseException] = []

    def feeder(k: int) -> None:
        try:
            for i in shards[k]:
                r = hash_media_file(engine, gpu=gpus[k], **jobs[i])
                r["gpu"] = gpus[k]
                out[i] = r
        except BaseException as e:  # noqa: BLE001 — re-raised in the caller's thread
            errors.append(e)

    threads = [threading.Thread(target=feeder, args=(k,), name=f"tritonhash-feed-gpu{gpus[k]}") for k in range(len(gpus))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    if errors:
        raise errors[0]
    return out  # type: ignore[return-value]
