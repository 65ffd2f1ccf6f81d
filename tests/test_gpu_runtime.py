"""GPU tests of the host runtime around the kernels (downloader_b200/csrc/runtime.cu): completion delivery,
non-blocking poll, allocation-free ETag combine, overflow accounting, per-GPU feeding.  Digests are checked
against hashlib / the oracle as everywhere else."""
import hashlib
import os
import subprocess
import sys
import threading
import time

import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHA1, MD5, SHA256 = 1, 2, 4


@pytest.fixture(scope="module")
def torch():
    import torch
    assert torch.cuda.is_available(), "-m gpu tests need a GPU"
    torch.cuda.set_device(0)
    return torch


def test_callback_runs_off_the_cuda_thread_and_may_retire_its_own_ticket(torch):
    """The callback is delivered by the library's completion thread with no lock held (tritonhash.h): it may call
    th_result on its own ticket — which a CUDA host function could not — and later submits still work."""
    from downloader_b200 import Engine
    got, done = {}, threading.Event()
    with Engine([0], ring_bytes=8 << 20) as e:
        def cb(ticket, status):
            got["status"] = status
            got["thread"] = threading.get_ident()
            got["md5"] = e.result(ticket, 2, MD5)[MD5].copy()  # th_result from inside the callback
            done.set()
        slot, view = e.ring_acquire(8192)
        view[:] = np.arange(8192, dtype=np.uint8)
        t = e.submit(slot, [0, 4096], [4096, 4096], MD5, callback=cb)
        assert done.wait(30)
        assert got["status"] == 0 and got["thread"] != threading.get_ident()
        ref = [hashlib.md5(bytes(view[o:o + 4096])).digest() for o in (0, 4096)]
        assert [bytes(x) for x in got["md5"]] == ref
        with pytest.raises(Exception) as ei:  # the ticket is gone: retired by the callback
            e.wait(t)
        assert getattr(ei.value, "code", 0) == -8
        # a second round on the same context: result() from the caller's thread waits for the callback to return
        seen = []
        slot, view = e.ring_acquire(4096)
        view[:] = 7
        t2 = e.submit(slot, [0], [4096], SHA1, callback=lambda tk, st: (time.sleep(0.05), seen.append((tk, st))))
        r = e.result(t2, 1, SHA1)
        assert seen == [(t2, 0)], "th_result returned before the ticket's callback had"
        assert bytes(r[SHA1][0]) == hashlib.sha1(bytes([7]) * 4096).digest()


def test_poll_never_blocks_and_turns_true(torch):
    from downloader_b200 import Engine, _capi
    L = _capi.lib()
    n, ln = 32, 8 << 20  # one group of 8 MiB SHA-256 chains: ~0.12 s
    d = torch.empty(n * ln, dtype=torch.uint8, device="cuda")
    assert L.th_synth_fill(d.data_ptr(), 0, n, ln, ln, oracle.DEFAULT_SEED, None) == 0
    torch.cuda.synchronize()
    with Engine([0], ring_bytes=0) as e:
        offs = np.arange(n, dtype=np.uint64) * ln
        t0 = time.perf_counter()
        t = e.submit_device(d.data_ptr(), offs, np.full(n, ln, np.uint64), SHA256)
        first = e.poll(t)
        dt_first = time.perf_counter() - t0
        assert first is False and dt_first < 0.05, (first, dt_first)
        deadline = time.perf_counter() + 30
        while not e.poll(t):
            assert time.perf_counter() < deadline
            time.sleep(0.002)
        r = e.result(t, n, SHA256)
        host = d[:2 * ln].cpu().numpy()
        for i in range(2):
            assert bytes(r[SHA256][i]) == hashlib.sha256(host[i * ln:(i + 1) * ln].tobytes()).digest()


def test_etag_combine_allocates_nothing_and_does_not_stall_running_chains(torch):
    """th_etag_combine used cudaMalloc/cudaFree per call, and an allocation synchronises the device.
    Now its buffers come from the staging arena: 1,000 combines issued while a batch of 16 MiB parts is being
    hashed leave free device memory unchanged and do not lengthen that ticket's SHA-256 chain (th_ticket_timing)."""
    from downloader_b200 import Engine, _capi
    L = _capi.lib()
    n, ln = 64, 16 << 20
    d = torch.empty(n * ln, dtype=torch.uint8, device="cuda")
    assert L.th_synth_fill(d.data_ptr(), 0, n, ln, ln, oracle.DEFAULT_SEED, None) == 0
    torch.cuda.synchronize()
    offs = np.arange(n, dtype=np.uint64) * ln
    lens = np.full(n, ln, np.uint64)
    rng = np.random.default_rng(7)
    md5s = rng.integers(0, 256, (64, 16), dtype=np.uint8)
    want = hashlib.md5(md5s.tobytes()).digest()
    with Engine([0], ring_bytes=0) as e:
        e.set_timing(True)
        e.etag_combine(md5s, [64])  # warm: module load, first launch
        tk = e.submit_device(d.data_ptr(), offs, lens, SHA256)
        alone = e.timing(tk)
        base = e.result(tk, n, SHA256)[SHA256].copy()
        alone_ms = alone[6] - alone[5]
        free0, _ = torch.cuda.mem_get_info()
        tk = e.submit_device(d.data_ptr(), offs, lens, SHA256)
        t0 = time.perf_counter()
        calls = 0
        while calls < 1000:
            out = e.etag_combine(md5s, [64])
            calls += 1
        spent = time.perf_counter() - t0
        assert bytes(out[0]) == want
        free1, _ = torch.cuda.mem_get_info()
        busy = e.timing(tk)
        r = e.result(tk, n, SHA256)[SHA256]
        busy_ms = busy[6] - busy[5]
        st = e.stats()
    assert (r == base).all()
    assert free1 == free0, "device memory moved during th_etag_combine"
    assert st["overflow_allocs"] == 0
    # the combines ran while the chain did (otherwise the test shows nothing) and the chain is not slower for it
    assert spent < alone_ms / 1e3 * 1.5 or spent < 1.0
    assert busy_ms < alone_ms * 1.10 + 2.0, (alone_ms, busy_ms, spent)


def test_overflow_allocations_are_counted(torch):
    """A submit whose descriptors + digests do not fit the staging arena (64 MiB here: > ~620 k messages) falls
    back to allocating; th_stats.overflow_allocs says so, and the digests are still right."""
    from downloader_b200 import Engine
    n = 700_000
    data = torch.from_numpy(np.random.default_rng(3).integers(0, 256, n * 16, dtype=np.uint8)).cuda()
    offs = np.arange(n, dtype=np.uint64) * 16
    lens = np.full(n, 16, np.uint64)
    with Engine([0], ring_bytes=0) as e:
        assert e.stats()["overflow_allocs"] == 0
        t = e.submit_device(data.data_ptr(), offs, lens, MD5)
        r = e.result(t, n, MD5)[MD5]
        assert e.stats()["overflow_allocs"] == 1
        t = e.submit_device(data.data_ptr(), offs, lens, MD5)  # the overflow buffer is kept and reused
        e.result(t, n, MD5)
        assert e.stats()["overflow_allocs"] == 1
    host = data.cpu().numpy().reshape(n, 16)
    for i in (0, 1, 31, 32, 4095, n // 2, n - 1):
        assert bytes(r[i]) == hashlib.md5(host[i].tobytes()).digest()


def test_device_fault_settles_callbacks_with_an_error():
    """after a device fault CUDA never runs the host functions queued behind it, so a caller waiting
    for its callback (the N-API Promise) would hang.  The completion thread's watchdog must deliver TH_ECUDA.
    Run in a child process: the fault poisons that process's CUDA context."""
    code = r"""
import sys, threading, numpy as np
sys.path.insert(0, ROOT_DIR)
from downloader_b200 import Engine
ev, seen = threading.Event(), {}
e = Engine([0], ring_bytes=0)
bad = 0x10  # not a device address: the kernel faults on its first read
n = 64
import ctypes as C
from downloader_b200 import _capi
# unlike lengths -> the ragged kernel (plain 16-byte async copies): an ordinary illegal-address fault
msgs = e._msgs(np.arange(n, dtype=np.uint64) * 4096, np.where(np.arange(n) % 2 == 0, 4096, 1000).astype(np.uint64))
tk = C.c_uint64()
cb = _capi.DONE_CB(lambda user, ticket, status: (seen.update(ticket=ticket, status=status), ev.set()))
rc = e._L.th_submit_device(e._ctx, 0, bad, msgs, n, 1, None, cb, None, C.byref(tk))
assert rc == 0, rc
assert ev.wait(20), "no callback after a device fault"
assert seen["status"] == -3 and seen["ticket"] == tk.value, seen
rc = e._L.th_result(e._ctx, tk.value, None, None, None, None)
assert rc == -3, rc
assert e._L.th_last_error(e._ctx), "no error text"
print("FAULT SETTLED", seen["status"])
import os
os._exit(0)
""".replace("ROOT_DIR", repr(ROOT))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=180)
    assert "FAULT SETTLED -3" in r.stdout, (r.stdout[-500:], r.stderr[-1500:])


def test_threads_feed_the_gpus_of_one_context_independently(torch):
    """One context, one feeder thread per GPU (the add-on's shape with several GPUs): each thread acquires, fills,
    submits and collects on its own GPU; the per-GPU locks mean none of them waits for another's blocking call."""
    from downloader_b200 import Engine
    ng = min(torch.cuda.device_count(), 4)
    if ng < 2:
        pytest.skip("needs 2+ GPUs (run on a multi-GPU gpurun; log under profiles/)")
    errors = []
    with Engine(list(range(ng)), ring_bytes=64 << 20) as e:
        def feeder(g):
            try:
                rng = np.random.default_rng(40 + g)
                for it in range(8):
                    data = rng.integers(0, 256, 16 << 20, dtype=np.uint8)
                    slot, view = e.ring_acquire(data.size, gpu=g)
                    view[:] = data
                    offs = np.arange(0, data.size, 1 << 20, dtype=np.uint64)
                    lens = np.full(offs.size, 1 << 20, np.uint64)
                    t = e.submit(slot, offs, lens, SHA1 | MD5)
                    r = e.result(t, offs.size, SHA1 | MD5)
                    for i in (0, offs.size - 1):
                        b = data[int(offs[i]):int(offs[i]) + (1 << 20)].tobytes()
                        if bytes(r[SHA1][i]) != hashlib.sha1(b).digest() or bytes(r[MD5][i]) != hashlib.md5(b).digest():
                            errors.append((g, it, i))
            except Exception as ex:  # noqa: BLE001
                errors.append((g, repr(ex)))
        th = [threading.Thread(target=feeder, args=(g,)) for g in range(ng)]
        for x in th:
            x.start()
        for x in th:
            x.join()
        assert not errors, errors[:4]
        for g in range(ng):
            assert e.stats(g)["submits"] == 8
