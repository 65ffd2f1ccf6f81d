"""CPU oracle for the byte-integrity path — TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py`` may import this package.  The product (``downloader_b200``) never does.

PARITY UNPINNED by the reference: tritonmedia/downloader holds no digest test or golden vector
(its tests are test/process/filter_dirs.js:17-82 and test/eslint.js:1-30).  The arithmetic lives in
un-vendored dependencies (simple-sha1@2.1.2 yarn.lock:3182, minio@7.0.12 yarn.lock:2152) that
forward to Node crypto = OpenSSL; call sites lib/download.js:64-66 and lib/upload.js:45,55.
The oracle is ``liboracle.so``: ``digest_oracle.c`` (plain-C restatement of RFC 1321 / FIPS 180-4)
and ``evp_ref.c`` (this image's libcrypto, the library Node binds), pinned by tests/golden/kat.json.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

SHA1, MD5, SHA256 = 1, 2, 4
DIGEST_LEN = {SHA1: 20, MD5: 16, SHA256: 32}
ALG_NAME = {SHA1: "sha1", MD5: "md5", SHA256: "sha256"}
DEFAULT_SEED = 0x747269746F6E  # "triton", SURVEY §8d

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile liboracle.so with gcc (needs libcrypto headers, present on this image)."""
    if force or not os.path.exists(_LIB_PATH):
        subprocess.check_call(["make", "-C", _HERE, "-B" if force else "-s", "liboracle.so"])
    return _LIB_PATH


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        u8p, u64p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint64)
        L.orc_batch.argtypes = [C.c_uint32, C.c_void_p, u64p, u64p, C.c_size_t, C.c_void_p, C.c_int]
        L.orc_batch.restype = None
        L.orc_synth_fill.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64]
        L.orc_synth_fill.restype = None
        L.orc_splitmix64.argtypes = [C.c_uint64]
        L.orc_splitmix64.restype = C.c_uint64
        L.orc_etag_multipart.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        L.orc_etag_multipart.restype = None
        L.orc_calc_part_size.argtypes = [C.c_uint64]
        L.orc_calc_part_size.restype = C.c_uint64
        L.orc_hex.argtypes = [C.c_void_p, C.c_size_t, C.c_char_p]
        L.orc_base64.argtypes = [C.c_void_p, C.c_size_t, C.c_char_p]
        L.orc_base64.restype = C.c_size_t
        L.orc_init.argtypes = [C.c_void_p, C.c_uint32]
        L.orc_update.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.orc_update.restype = None
        L.orc_final.argtypes = [C.c_void_p, C.c_void_p]
        L.evp_digest.argtypes = [C.c_uint32, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p]
        L.evp_digest.restype = C.c_int
        L.evp_batch.argtypes = [C.c_uint32, C.c_void_p, u64p, u64p, C.c_size_t, C.c_size_t, C.c_int,
                                C.c_void_p, C.c_void_p, C.c_void_p]
        L.evp_batch.restype = C.c_double
        L.evp_synth_fill_mt.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int]
        L.evp_synth_fill_mt.restype = None
        L.evp_version.restype = C.c_char_p
        _lib = L
    return _lib


def _as_u8(buf) -> np.ndarray:
    a = np.frombuffer(buf, dtype=np.uint8) if not isinstance(buf, np.ndarray) else buf
    return np.ascontiguousarray(a.view(np.uint8).reshape(-1))


def digest(alg: int, data) -> bytes:
    """Plain-C restatement, one message."""
    a = _as_u8(data)
    off = np.zeros(1, np.uint64)
    ln = np.array([a.size], np.uint64)
    out = np.zeros(DIGEST_LEN[alg], np.uint8)
    lib().orc_batch(alg, a.ctypes.data, off.ctypes.data_as(C.POINTER(C.c_uint64)),
                    ln.ctypes.data_as(C.POINTER(C.c_uint64)), 1, out.ctypes.data, 1)
    return out.tobytes()


def digest_chunked(alg: int, data, chunk: int) -> bytes:
    """createHash(alg); update() per `chunk` bytes; digest() — through the C restatement."""
    a = _as_u8(data)
    ctx = C.create_string_buffer(256)
    L = lib()
    assert L.orc_init(ctx, alg) == 0
    for o in range(0, a.size, max(chunk, 1)):
        n = min(chunk, a.size - o)
        L.orc_update(ctx, a.ctypes.data + o, n)
    out = np.zeros(DIGEST_LEN[alg], np.uint8)
    L.orc_final(ctx, out.ctypes.data)
    return out.tobytes()


def evp(alg: int, data, chunk: int = 0) -> bytes:
    """libcrypto EVP (the code Node's crypto runs), one message."""
    a = _as_u8(data)
    out = np.zeros(DIGEST_LEN[alg], np.uint8)
    n = lib().evp_digest(alg, a.ctypes.data, a.size, chunk, out.ctypes.data)
    assert n == DIGEST_LEN[alg], n
    return out.tobytes()


def batch(alg: int, base, offsets, lengths, threads: int = 1) -> np.ndarray:
    """Plain-C restatement over a flat buffer -> (n, digest_len) uint8."""
    a = _as_u8(base)
    off = np.ascontiguousarray(offsets, np.uint64)
    ln = np.ascontiguousarray(lengths, np.uint64)
    out = np.zeros((off.size, DIGEST_LEN[alg]), np.uint8)
    u64p = C.POINTER(C.c_uint64)
    lib().orc_batch(alg, a.ctypes.data, off.ctypes.data_as(u64p), ln.ctypes.data_as(u64p), off.size,
                    out.ctypes.data, threads)
    return out


def evp_batch(alg_mask: int, base, offsets, lengths, threads: int = 1, chunk: int = 0):
    """libcrypto over a flat buffer. Returns (seconds, {alg: (n, len) uint8})."""
    a = _as_u8(base)
    off = np.ascontiguousarray(offsets, np.uint64)
    ln = np.ascontiguousarray(lengths, np.uint64)
    outs = {g: np.zeros((off.size, DIGEST_LEN[g]), np.uint8) for g in (SHA1, MD5, SHA256) if alg_mask & g}
    u64p = C.POINTER(C.c_uint64)

    def p(g):
        return outs[g].ctypes.data if g in outs else None

    dt = lib().evp_batch(alg_mask, a.ctypes.data, off.ctypes.data_as(u64p), ln.ctypes.data_as(u64p),
                         off.size, chunk, threads, p(SHA1), p(MD5), p(SHA256))
    if dt < 0:
        raise RuntimeError("libcrypto EVP digest failed")
    return dt, outs


def synth(msg_index: int, nbytes: int, seed: int = DEFAULT_SEED, byte_off: int = 0) -> np.ndarray:
    """Synthetic message bytes (SURVEY §8d generator)."""
    out = np.zeros(nbytes, np.uint8)
    if nbytes:
        lib().orc_synth_fill(out.ctypes.data, msg_index, byte_off, nbytes, seed)
    return out


def synth_batch(first_msg: int, n: int, length: int, seed: int = DEFAULT_SEED, threads: int = 0,
                out: np.ndarray | None = None) -> np.ndarray:
    """n equal-length synthetic messages back to back (threaded)."""
    if out is None:
        out = np.empty(n * length, np.uint8)
    assert out.size >= n * length
    if n * length:
        lib().evp_synth_fill_mt(out.ctypes.data, first_msg, n, length, seed, threads or (os.cpu_count() or 1))
    return out


def etag_multipart(md5s) -> str:
    """S3 multipart ETag: hex(md5(md5_1||...||md5_N)) + '-' + N."""
    a = _as_u8(md5s)
    assert a.size % 16 == 0
    out = np.zeros(16, np.uint8)
    lib().orc_etag_multipart(a.ctypes.data, a.size // 16, out.ctypes.data)
    return out.tobytes().hex() + "-" + str(a.size // 16)


def calc_part_size(size: int) -> int:
    return int(lib().orc_calc_part_size(size))


def hex_(b: bytes) -> str:
    out = C.create_string_buffer(2 * len(b) + 1)
    lib().orc_hex(b, len(b), out)
    return out.value.decode()


def base64_(b: bytes) -> str:
    out = C.create_string_buffer(4 * ((len(b) + 2) // 3) + 1)
    lib().orc_base64(b, len(b), out)
    return out.value.decode()


def openssl_version() -> str:
    return lib().evp_version().decode()
