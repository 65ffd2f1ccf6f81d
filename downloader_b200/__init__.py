"""downloader_b200 — B200-native byte-integrity engine for the tritonmedia/downloader worker.

Only the hot path of the reference is here: SHA-1 verification of torrent pieces
(reference call site lib/download.js:64-66) and MD5 / SHA-256 of S3 parts (lib/upload.js:45,55),
as hand-written sm_100a kernels behind the C ABI of include/tritonhash.h
(downloader_b200/libtritonhash.so).  There is no CPU fallback: importing the binding without the
built library raises.
"""
from ._capi import ALG_MD5, ALG_SHA1, ALG_SHA256, DIGEST_LEN, TritonHashError, lib, library_path  # noqa: F401
from .engine import Engine  # noqa: F401

__all__ = ["Engine", "ALG_SHA1", "ALG_MD5", "ALG_SHA256", "DIGEST_LEN", "TritonHashError", "lib", "library_path"]
