"""th_digest_ragged_static — the entry point for callers that must not allocate.

A plug-in called from a media pipeline's callback, or the add-on once its rings fill the GPU, cannot
afford cudaMalloc (it synchronises the device and may simply fail).  This entry point keeps its
length-bucketing table inside the library image instead.  The test calls it through the C ABI twice
in a row — the second batch smaller and differently shaped, so anything left in the shared table by
the first call would show — and requires bit-exact agreement with the existing launch-level entry
points (th_sha1_batch / th_md5_batch / th_sha256_batch) on the same device buffers, and, first-hand, with
hashlib on the same bytes.  A third case fills the table to exactly its stated maximum.
Collected last on purpose (file name), so a failure here cannot hide the other tests' results.
"""
import hashlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SHA1, MD5, SHA256 = 1, 2, 4
DLEN = {SHA1: 20, MD5: 16, SHA256: 32}
MAX_MSGS = 65536
HASHLIB = {SHA1: "sha1", MD5: "md5", SHA256: "sha256"}


def _batch(torch, rng, lens):
    offs, pos = [], 0
    for n in lens:
        offs.append(pos)
        pos = (pos + n + 15) & ~15
    data = torch.from_numpy(rng.integers(0, 256, max(pos, 16), dtype=np.uint8)).cuda()
    d_off = torch.tensor(offs, dtype=torch.int64, device="cuda")
    d_len = torch.tensor(lens, dtype=torch.int64, device="cuda")
    return data, d_off, d_len


def _existing(L, torch, alg, data, d_off, d_len, n, expected=None):
    out = torch.zeros(n * DLEN[alg], dtype=torch.uint8, device="cuda")
    ok = torch.zeros((n + 31) // 32, dtype=torch.int32, device="cuda")
    if alg == SHA1:
        rc = L.th_sha1_batch(data.data_ptr(), d_off.data_ptr(), d_len.data_ptr(), None, 0, 0, n, out.data_ptr(),
                             expected.data_ptr() if expected is not None else None,
                             ok.data_ptr() if expected is not None else None, None)
    elif alg == MD5:
        rc = L.th_md5_batch(data.data_ptr(), d_off.data_ptr(), d_len.data_ptr(), None, 0, 0, n, out.data_ptr(), None)
    else:
        rc = L.th_sha256_batch(data.data_ptr(), d_off.data_ptr(), d_len.data_ptr(), None, 0, 0, n, out.data_ptr(), None)
    assert rc == 0
    torch.cuda.synchronize()
    return out.cpu().numpy(), ok.cpu().numpy()


def test_no_alloc_entry_point_matches_existing_on_two_calls_in_a_row():
    import torch

    from downloader_b200 import _capi
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    L = _capi.lib()
    rng = np.random.default_rng(2026)
    first = [262144] * 40 + [0, 1, 55, 56, 63, 64, 65, 119, 120, 4096, 100000, 77] + rng.integers(0, 70000, 900).tolist()
    second = [5000, 3, 0, 1 << 20, 64, 64, 64, 129] + rng.integers(0, 3000, 150).tolist()
    for lens in (first, second):  # two calls in a row on the same device; the table is shared
        n = len(lens)
        data, d_off, d_len = _batch(torch, rng, lens)
        for alg in (SHA1, MD5, SHA256):
            ref, _ = _existing(L, torch, alg, data, d_off, d_len, n)
            out = torch.zeros(n * DLEN[alg], dtype=torch.uint8, device="cuda")
            rc = L.th_digest_ragged_static(alg, data.data_ptr(), d_off.data_ptr(), d_len.data_ptr(), n,
                                           out.data_ptr(), None, None, None)
            assert rc == 0
            torch.cuda.synchronize()
            assert (out.cpu().numpy() == ref).all(), (alg, n)
            # first-hand: the same bytes through hashlib (OpenSSL), not only CUDA against CUDA
            host, got = data.cpu().numpy(), out.cpu().numpy().reshape(n, DLEN[alg])
            offs = d_off.cpu().numpy()
            for i in range(n):
                m = host[int(offs[i]):int(offs[i]) + lens[i]].tobytes()
                assert bytes(got[i]) == hashlib.new(HASHLIB[alg], m).digest(), (alg, i, lens[i])
        # SHA-1 with the fused verify: same bitfield as the existing entry point
        exp = torch.from_numpy(_existing(L, torch, SHA1, data, d_off, d_len, n)[0].copy()).cuda()
        bad = sorted({0, n // 2, n - 1})
        for b in bad:
            exp.view(n, 20)[b, 11] ^= 4
        ref, ref_ok = _existing(L, torch, SHA1, data, d_off, d_len, n, expected=exp)
        out = torch.zeros(n * 20, dtype=torch.uint8, device="cuda")
        ok = torch.full(((n + 31) // 32,), -1, dtype=torch.int32, device="cuda")
        rc = L.th_digest_ragged_static(SHA1, data.data_ptr(), d_off.data_ptr(), d_len.data_ptr(), n, out.data_ptr(),
                                       exp.data_ptr(), ok.data_ptr(), None)
        assert rc == 0
        torch.cuda.synchronize()
        assert (out.cpu().numpy() == ref).all()
        assert (ok.cpu().numpy() == ref_ok).all()
        words = ok.cpu().numpy().view(np.uint32)
        bits = ((words[np.arange(n) >> 5] >> (np.arange(n) & 31)) & 1).astype(bool)
        assert sorted(np.nonzero(~bits)[0].tolist()) == bad


def test_no_alloc_entry_point_at_exactly_its_maximum():
    """n == TH_STATIC_MAX_MSGS fills every slot of the built-in table: accepted, and every digest right."""
    import torch

    from downloader_b200 import _capi
    torch.cuda.set_device(0)
    L = _capi.lib()
    rng = np.random.default_rng(65536)
    lens = rng.integers(0, 700, MAX_MSGS).tolist()
    lens[0], lens[-1], lens[12345] = 70000, 0, 64
    data, d_off, d_len = _batch(torch, rng, lens)
    out = torch.zeros(MAX_MSGS * 20, dtype=torch.uint8, device="cuda")
    rc = L.th_digest_ragged_static(SHA1, data.data_ptr(), d_off.data_ptr(), d_len.data_ptr(), MAX_MSGS, out.data_ptr(),
                                   None, None, None)
    assert rc == 0
    torch.cuda.synchronize()
    ref, _ = _existing(L, torch, SHA1, data, d_off, d_len, MAX_MSGS)
    got = out.cpu().numpy()
    assert (got == ref).all()
    host, offs, got = data.cpu().numpy(), d_off.cpu().numpy(), got.reshape(MAX_MSGS, 20)
    for i in list(range(0, MAX_MSGS, 257)) + [0, 12345, MAX_MSGS - 1]:
        assert bytes(got[i]) == hashlib.sha1(host[int(offs[i]):int(offs[i]) + lens[i]].tobytes()).digest(), i


def test_no_alloc_entry_point_refuses_more_than_its_table_holds():
    from downloader_b200 import _capi
    L = _capi.lib()
    assert L.th_digest_ragged_static(SHA1, 16, 16, 16, MAX_MSGS + 1, 16, None, None, None) == -7  # TH_ETOOBIG
    assert b"limit" in L.th_strerror(-7)
