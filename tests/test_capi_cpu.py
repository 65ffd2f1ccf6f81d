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


def test_napi_shim_type_checks_and_links_against_the_c_abi(tmp_path):
    """js/addon/tritonhash_napi.cc cannot run here (no Node), but it must at least compile against N-API's documented
    prototypes (tests/stubs/node_api.h) and resolve every th_* symbol it uses in libtritonhash.so."""
    import shutil
    import subprocess
    if not shutil.which("g++"):
        pytest.skip("g++ not available")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib = os.path.join(root, "downloader_b200", "libtritonhash.so")
    if not os.path.exists(lib):
        pytest.skip("libtritonhash.so not built")
    out = tmp_path / "tritonhash.node"
    cmd = ["g++", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-shared", "-fPIC", "-I", os.path.join(root, "tests", "stubs"),
           "-I", os.path.join(root, "include"), os.path.join(root, "js", "addon", "tritonhash_napi.cc"), "-o", str(out),
           "-L", os.path.dirname(lib), "-ltritonhash"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    syms = subprocess.run(["nm", "-D", str(out)], capture_output=True, text=True).stdout
    assert " T napi_register_module_v1" in syms
    used = sorted({ln.split()[-1] for ln in syms.splitlines() if " U th_" in ln})
    assert "th_init_staged" in used and "th_submit" in used and "th_result" in used
    exported = subprocess.run(["nm", "-D", "--defined-only", lib], capture_output=True, text=True).stdout
    for s in used:
        assert f" T {s}" in exported, s
