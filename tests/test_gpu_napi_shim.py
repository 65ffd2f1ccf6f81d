"""Runs js/addon/tritonhash_napi.cc — the reference-side N-API binding of INTEGRATION.md — on the GPU without Node:
tests/napi_mock/napi_mock.cc is a miniature N-API runtime plus a driver that plays js/lib/tritonhash.js
(init -> ringAcquire -> hashBatch Promise -> etagCombine -> stats -> shutdown) and checks every digest, the verify
bitfield and the ETag against libcrypto, and that bad arguments become JS exceptions."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_napi_shim_executes_against_a_mock_runtime(tmp_path):
    if not shutil.which("g++"):
        pytest.skip("g++ not available")
    libdir = os.path.join(ROOT, "downloader_b200")
    exe = tmp_path / "napi_mock"
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-Wextra", "-I", os.path.join(ROOT, "tests", "stubs"), "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "napi_mock", "napi_mock.cc"), os.path.join(ROOT, "js", "addon", "tritonhash_napi.cc"),
           "-L", libdir, "-ltritonhash", "-lcrypto", "-lpthread", f"-Wl,-rpath,{libdir}", "-o", str(exe)]
    b = subprocess.run(cmd, capture_output=True, text=True)
    assert b.returncode == 0, b.stderr[-3000:]
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "NAPI MOCK OK" in r.stdout, (r.stdout[-1000:], r.stderr[-2000:])
