"""CPU-side checks of the boundary: the library loads, exports every declared symbol, validates
arguments, and refuses to run without a GPU (no CPU fallback).  No compute calls here."""
import ctypes as C
import os
import subprocess

import pytest

from downloader_b200 import _capi


def test_library_is_built_in_tree():
    assert os.path.exists(_capi.library_path()), "run __graft_entry__.build() first"


def test_exports_every_declared_symbol():
    L = _capi.lib()
    syms = _capi.declared_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/tritonhash.h but not exported"
    out = subprocess.run(["nm", "-D", "--defined-only", _capi.library_path()], capture_output=True, text=True).stdout
    exported = {ln.split()[-1] for ln in out.splitlines() if " T " in ln}
    assert set(syms) <= exported


def test_library_does_not_link_the_oracle():
    out = subprocess.run(["ldd", _capi.library_path()], capture_output=True, text=True).stdout
    assert "liboracle" not in out and "libcrypto" not in out
    nm = subprocess.run(["nm", "-D", _capi.library_path()], capture_output=True, text=True).stdout
    assert "orc_" not in nm and "EVP_" not in nm


def test_strerror_and_encoders():
    L = _capi.lib()
    assert L.th_strerror(0) == b"ok"
    assert b"16-byte" in L.th_strerror(-5)
    assert b"no CPU fallback" in L.th_strerror(-2)
    assert _capi.hex_(b"\x00\xab\xff") == "00abff"
    import base64
    for n in range(0, 40):
        b = bytes((7 * i + 3) & 255 for i in range(n))
        assert _capi.base64_(b) == base64.b64encode(b).decode()


def test_staging_variants_selectable():
    L = _capi.lib()
    ch, st = C.c_int(), C.c_int()
    assert L.th_get_staging(1, C.byref(ch), C.byref(st)) == 0
    default = (ch.value, st.value)
    assert L.th_set_staging(7, 256, 2) == 0
    assert L.th_get_staging(4, C.byref(ch), C.byref(st)) == 0 and (ch.value, st.value) == (256, 2)
    assert L.th_set_staging(7, 100, 2) == -1  # TH_EINVAL: no such variant
    assert L.th_set_staging(7, *default) == 0


def test_argument_validation_without_gpu():
    L = _capi.lib()
    # misaligned base -> TH_EALIGN before anything touches CUDA
    assert L.th_sha1_batch(8, None, None, None, 64, 64, 4, 16, None, None, None) == -5
    assert L.th_sha1_batch(16, None, None, None, 24, 64, 4, 16, None, None, None) == -5  # stride not 16-aligned
    assert L.th_md5_batch(16, None, None, None, 64, 64, 4, None, None) == -1            # no output buffer
    assert L.th_sha256_batch(16, None, None, None, 64, 1 << 40, 4, 16, None) == -7      # TH_ETOOBIG
    assert L.th_init(None, 0, 0, None) == -1


@pytest.mark.skipif(os.path.exists("/dev/nvidiactl"), reason="GPU present")
def test_init_fails_loudly_without_gpu():
    import downloader_b200
    with pytest.raises(downloader_b200.TritonHashError) as ei:
        downloader_b200.Engine()
    assert ei.value.code == -2
