"""ctypes binding of include/tritonhash.h — the same C ABI the N-API shim binds."""
from __future__ import annotations

import ctypes as C
import os
import re

ALG_SHA1, ALG_MD5, ALG_SHA256 = 1, 2, 4
DIGEST_LEN = {ALG_SHA1: 20, ALG_MD5: 16, ALG_SHA256: 32}

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
_LIB = os.path.join(_HERE, "libtritonhash.so")
_HEADER = os.path.join(_ROOT, "include", "tritonhash.h")


class TritonHashError(RuntimeError):
    def __init__(self, code: int, where: str, detail: str = ""):
        self.code = code
        msg = f"{where}: {_strerror(code)} ({code})"
        if detail:
            msg += f" — {detail}"
        super().__init__(msg)


class ThMsg(C.Structure):
    _fields_ = [("offset", C.c_uint64), ("length", C.c_uint64)]


class ThStats(C.Structure):
    _fields_ = [("messages", C.c_uint64), ("bytes", C.c_uint64), ("launches", C.c_uint64),
                ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64), ("submits", C.c_uint64),
                ("kernel_ms", C.c_double), ("overflow_allocs", C.c_uint64)]


DONE_CB = C.CFUNCTYPE(None, C.c_void_p, C.c_uint64, C.c_int)

_lib = None


def library_path() -> str:
    return _LIB


def declared_symbols() -> list[str]:
    """Every function include/tritonhash.h declares (used by the export test)."""
    with open(_HEADER) as f:
        src = re.sub(r"/\*.*?\*/", "", f.read(), flags=re.S)
    return sorted(set(re.findall(r"\b(th_[a-z0-9_]+)\s*\(", src)))


def _strerror(code: int) -> str:
    try:
        return lib().th_strerror(code).decode()
    except Exception:  # library not loadable: still give the code
        return "error"


def lib() -> C.CDLL:
    """Loads libtritonhash.so.  Raises if it has not been built: there is no fallback path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB):
        raise ImportError(
            f"{_LIB} is missing — build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). downloader_b200 has no CPU fallback.")
    L = C.CDLL(_LIB)
    vp, u64, u32, i32 = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int
    L.th_init.argtypes = [C.POINTER(i32), i32, C.c_size_t, C.POINTER(vp)]
    L.th_init_staged.argtypes = [C.POINTER(i32), i32, C.c_size_t, C.c_size_t, C.POINTER(vp)]
    L.th_shutdown.argtypes = [vp]
    L.th_gpu_count.argtypes = [vp]
    L.th_ring_acquire.argtypes = [vp, i32, C.c_size_t, C.POINTER(vp), C.POINTER(u64)]
    L.th_ring_try_acquire.argtypes = [vp, i32, C.c_size_t, C.POINTER(vp), C.POINTER(u64)]
    L.th_ring_release.argtypes = [vp, u64]
    L.th_submit.argtypes = [vp, u64, C.POINTER(ThMsg), C.c_size_t, u32, vp, DONE_CB, vp, C.POINTER(u64)]
    L.th_submit_device.argtypes = [vp, i32, vp, C.POINTER(ThMsg), C.c_size_t, u32, vp, DONE_CB, vp, C.POINTER(u64)]
    L.th_wait.argtypes = [vp, u64]
    L.th_poll.argtypes = [vp, u64]
    L.th_ticket_timing.argtypes = [vp, u64, C.POINTER(C.c_double)]
    L.th_result.argtypes = [vp, u64, vp, vp, vp, vp]
    L.th_etag_combine.argtypes = [vp, i32, vp, vp, C.c_size_t, vp]
    L.th_stats_get.argtypes = [vp, i32, C.POINTER(ThStats)]
    L.th_strerror.argtypes = [i32]
    L.th_strerror.restype = C.c_char_p
    L.th_last_error.argtypes = [vp]
    L.th_last_error.restype = C.c_char_p
    L.th_hex.argtypes = [vp, C.c_size_t, C.c_char_p]
    L.th_hex.restype = None
    L.th_base64.argtypes = [vp, C.c_size_t, C.c_char_p]
    L.th_base64.restype = C.c_size_t
    L.th_sha1_batch.argtypes = [vp, vp, vp, vp, u64, u64, u32, vp, vp, vp, vp]
    L.th_md5_batch.argtypes = [vp, vp, vp, vp, u64, u64, u32, vp, vp]
    L.th_sha256_batch.argtypes = [vp, vp, vp, vp, u64, u64, u32, vp, vp]
    L.th_md5_sha256_batch.argtypes = [vp, u64, u64, u32, vp, vp, vp]
    L.th_md5_sha256_fused.argtypes = [u64, u64, u32]
    L.th_etag_combine_device.argtypes = [vp, vp, u32, vp, vp]
    L.th_synth_fill.argtypes = [vp, u64, u64, u64, u64, u64, vp]
    L.th_digest_ragged_static.argtypes = [u32, vp, vp, vp, u32, vp, vp, vp, vp]
    L.th_set_staging.argtypes = [u32, i32, i32]
    L.th_get_staging.argtypes = [u32, C.POINTER(i32), C.POINTER(i32)]
    L.th_set_tiling.argtypes = [u32, i32, i32]
    L.th_set_option.argtypes = [C.c_char_p, C.c_longlong]
    L.th_set_ticket_timing.argtypes = [vp, C.c_int]
    L.th_plan.argtypes = [u32, u64, u64, u32, C.c_char_p, C.c_size_t]
    _lib = L
    return L


def check(code: int, where: str, ctx=None) -> None:
    if code != 0:
        detail = ""
        if ctx is not None and code == -3:
            detail = lib().th_last_error(ctx).decode()
        raise TritonHashError(code, where, detail)


def plan(alg: int, length: int, stride: int, n: int) -> str:
    buf = C.create_string_buffer(512)
    check(lib().th_plan(alg, length, stride, n, buf, 512), "th_plan")
    return buf.value.decode()


def hex_(b: bytes) -> str:
    out = C.create_string_buffer(2 * len(b) + 1)
    lib().th_hex(b, len(b), out)
    return out.value.decode()


def base64_(b: bytes) -> str:
    out = C.create_string_buffer(4 * ((len(b) + 2) // 3) + 1)
    lib().th_base64(b, len(b), out)
    return out.value.decode()
