"""Host-side mirror of the reference's digest call shapes, batched, over the C ABI.

Reference shapes:
  [dep] simple-sha1@2.1.2   sha1(buf, cb) / sha1.sync(buf) -> lower-case hex          (lib/download.js:64-66)
  [dep] minio@7.0.12        getHashSummer(enableSHA256) -> {md5sum: base64, sha256sum: hex}  (lib/upload.js:45,55)
Here they are batched: Engine.hash_batch(list of buffers, alg mask, expected?) -> digests, ok bits.
"""
from __future__ import annotations

import ctypes as C
from typing import Iterable, Sequence

import numpy as np

from . import _capi
from ._capi import ALG_MD5, ALG_SHA1, ALG_SHA256, DIGEST_LEN, ThMsg, ThStats, check

_NULL_CB = C.cast(None, _capi.DONE_CB)


def _align16(x: int) -> int:
    return (x + 15) & ~15


class Engine:
    """One context = per-GPU pinned ring + device staging ring + streams (th_init_staged / th_shutdown).

    device_bytes sizes the device ring (default: same as the pinned ring).  Make it larger for big parts:
    pinned bytes are reusable once copied, device bytes only once hashed (~0.25 s per 16 MiB chain)."""

    def __init__(self, gpu_ids: Sequence[int] | None = None, ring_bytes: int = 256 << 20, device_bytes: int = 0):
        self._L = _capi.lib()
        ids = list(gpu_ids) if gpu_ids is not None else [0]
        arr = (C.c_int * len(ids))(*ids)
        self._ctx = C.c_void_p()
        check(self._L.th_init_staged(arr, len(ids), ring_bytes, device_bytes, C.byref(self._ctx)), "th_init_staged")
        self.gpu_ids = ids
        self.ring_bytes = ring_bytes
        self.device_bytes = device_bytes or ring_bytes
        self._cbs = {}
        self._retired_cbs = []  # thunks of retired tickets: result() may run INSIDE the callback, whose thunk must outlive it

    # -- lifecycle ------------------------------------------------------------------------------
    def close(self) -> None:
        if self._ctx:
            self._L.th_shutdown(self._ctx)
            self._ctx = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- ring -----------------------------------------------------------------------------------
    def ring_acquire(self, nbytes: int, gpu: int = 0, wait: bool = True):
        """-> (slot id, writable uint8 numpy view of the pinned slot).  wait=False never blocks: returns None when
        the bytes do not fit right now (th_ring_try_acquire)."""
        ptr, slot = C.c_void_p(), C.c_uint64()
        if not wait:
            rc = self._L.th_ring_try_acquire(self._ctx, gpu, nbytes, C.byref(ptr), C.byref(slot))
            if rc == -6:  # TH_EBUSY
                return None
            check(rc, "th_ring_try_acquire", self._ctx)
        else:
            check(self._L.th_ring_acquire(self._ctx, gpu, nbytes, C.byref(ptr), C.byref(slot)), "th_ring_acquire", self._ctx)
        view = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(nbytes,))
        return slot.value, view

    def ring_release(self, slot: int) -> None:
        check(self._L.th_ring_release(self._ctx, slot), "th_ring_release", self._ctx)

    @staticmethod
    def _msgs(offsets, lengths):
        n = len(offsets)
        arr = (ThMsg * n)()
        a = np.ctypeslib.as_array(arr).view(np.uint64).reshape(n, 2)
        a[:, 0] = offsets
        a[:, 1] = lengths
        return arr

    def submit(self, slot: int, offsets, lengths, algs: int, expected_sha1=None, callback=None) -> int:
        """th_submit.  callback(ticket, status), if given, is delivered once on the library's completion thread
        with the ticket's real status (TH_OK = 0, or -3 when the device faulted); it may call result() / wait() /
        submit() of this engine, and the ticket must still be retired with result() whatever the status."""
        msgs = self._msgs(offsets, lengths)
        exp = None
        if expected_sha1 is not None:
            exp = np.ascontiguousarray(expected_sha1, np.uint8).reshape(-1)
            assert exp.size == 20 * len(offsets)
        t = C.c_uint64()
        cb = _NULL_CB
        if callback is not None:
            cb = _capi.DONE_CB(lambda user, ticket, status: callback(ticket, status))
        rc = self._L.th_submit(self._ctx, slot, msgs, len(offsets), algs,
                               exp.ctypes.data if exp is not None else None, cb, None, C.byref(t))
        check(rc, "th_submit", self._ctx)
        if callback is not None:
            self._cbs[t.value] = cb
        return t.value

    def submit_device(self, d_base: int, offsets, lengths, algs: int, expected_sha1=None, gpu: int = 0) -> int:
        msgs = self._msgs(offsets, lengths)
        exp = None
        if expected_sha1 is not None:
            exp = np.ascontiguousarray(expected_sha1, np.uint8).reshape(-1)
            assert exp.size == 20 * len(offsets)
        t = C.c_uint64()
        rc = self._L.th_submit_device(self._ctx, gpu, d_base, msgs, len(offsets), algs,
                                      exp.ctypes.data if exp is not None else None, _NULL_CB, None, C.byref(t))
        check(rc, "th_submit_device", self._ctx)
        return t.value

    def set_timing(self, on: bool) -> None:
        """Per-ticket device timeline + stats()['kernel_ms'] for later submits (th_set_ticket_timing; a diagnostic:
        it limits the library to 30 concurrent digest streams)."""
        check(self._L.th_set_ticket_timing(self._ctx, 1 if on else 0), "th_set_ticket_timing", self._ctx)

    def timing(self, ticket: int) -> list:
        """[ready, sha1 start, sha1 end, md5 start, md5 end, sha256 start, sha256 end] in ms since th_init."""
        out = (C.c_double * 7)()
        check(self._L.th_ticket_timing(self._ctx, ticket, out), "th_ticket_timing", self._ctx)
        return list(out)

    def wait(self, ticket: int) -> None:
        check(self._L.th_wait(self._ctx, ticket), "th_wait", self._ctx)

    def poll(self, ticket: int) -> bool:
        """th_poll: True when the ticket's digests are in host memory, False while it runs; never blocks."""
        rc = self._L.th_poll(self._ctx, ticket)
        if rc == -6:  # TH_EBUSY
            return False
        check(rc, "th_poll", self._ctx)
        return True

    def result(self, ticket: int, n: int, algs: int, want_ok: bool = False) -> dict:
        out = {}
        bufs = {a: np.empty((n, DIGEST_LEN[a]), np.uint8) for a in (ALG_SHA1, ALG_MD5, ALG_SHA256) if algs & a}
        ok = np.zeros((n + 31) // 32, np.uint32) if want_ok else None

        def p(a):
            return bufs[a].ctypes.data if a in bufs else None

        rc = self._L.th_result(self._ctx, ticket, p(ALG_SHA1), p(ALG_MD5), p(ALG_SHA256),
                               ok.ctypes.data if ok is not None else None)
        cb = self._cbs.pop(ticket, None)
        if cb is not None:  # freed a little later, never while it may still be executing
            self._retired_cbs.append(cb)
            if len(self._retired_cbs) > 256:
                del self._retired_cbs[:128]
        check(rc, "th_result", self._ctx)
        out.update(bufs)
        if ok is not None:
            out["ok_bits"] = ok
            out["ok"] = ((ok[np.arange(n) >> 5] >> (np.arange(n, dtype=np.uint32) & 31)) & 1).astype(bool)
        return out

    # -- the batched call the stages use -----------------------------------------------------------
    def hash_batch(self, buffers: Iterable, algs: int, expected_sha1=None, gpu: int = 0) -> dict:
        """Hashes a list of host buffers (bytes / memoryview / uint8 arrays) in one submit.

        Returns {ALG_*: (n, len) uint8, 'ok': bool[n] when expected_sha1 is given}.
        """
        bufs = [np.frombuffer(b, np.uint8) if not isinstance(b, np.ndarray) else b.reshape(-1).view(np.uint8)
                for b in buffers]
        n = len(bufs)
        if n == 0:
            return {a: np.empty((0, DIGEST_LEN[a]), np.uint8) for a in (ALG_SHA1, ALG_MD5, ALG_SHA256) if algs & a}
        lengths = np.array([b.size for b in bufs], np.uint64)
        offsets = np.zeros(n, np.uint64)
        pos = 0
        for i, b in enumerate(bufs):
            offsets[i] = pos
            pos = _align16(pos + b.size)
        slot, view = self.ring_acquire(max(pos, 16), gpu)
        try:
            for i, b in enumerate(bufs):
                o = int(offsets[i])
                view[o:o + b.size] = b
            t = self.submit(slot, offsets, lengths, algs, expected_sha1)
        except BaseException:
            self.ring_release(slot)  # a slot that is neither submitted nor released blocks the FIFO ring for good
            raise
        return self.result(t, n, algs, want_ok=expected_sha1 is not None and bool(algs & ALG_SHA1))

    def etag_combine(self, md5s, counts: Sequence[int], gpu: int = 0) -> np.ndarray:
        """md5 over each object's concatenated part digests -> (n_objects, 16) uint8."""
        m = np.ascontiguousarray(md5s, np.uint8).reshape(-1)
        cnt = np.ascontiguousarray(counts, np.uint32)
        assert m.size == 16 * int(cnt.sum())
        out = np.zeros((cnt.size, 16), np.uint8)
        rc = self._L.th_etag_combine(self._ctx, gpu, m.ctypes.data if m.size else None, cnt.ctypes.data, cnt.size,
                                     out.ctypes.data)
        check(rc, "th_etag_combine", self._ctx)
        return out

    def stats(self, gpu: int = -1) -> dict:
        s = ThStats()
        check(self._L.th_stats_get(self._ctx, gpu, C.byref(s)), "th_stats_get", self._ctx)
        return {k: getattr(s, k) for k, _ in ThStats._fields_}
