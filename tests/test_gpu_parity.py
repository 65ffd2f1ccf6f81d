"""GPU parity tests: every digest that leaves the CUDA path, through the C ABI, must be bit-exact
with the CPU oracle (oracle/: plain-C restatement + libcrypto, pinned by tests/golden/kat.json)."""
import ctypes as C
import hashlib

import numpy as np
import pytest

import oracle
from conftest import kat_message

pytestmark = pytest.mark.gpu

SHA1, MD5, SHA256 = 1, 2, 4
ALL = SHA1 | MD5 | SHA256
NAMES = {SHA1: "sha1", MD5: "md5", SHA256: "sha256"}
VARIANTS = [(64, 4), (128, 2), (128, 3), (128, 4), (256, 2), (256, 3), (512, 2)]


@pytest.fixture(scope="module")
def torch():
    import torch
    assert torch.cuda.is_available(), "-m gpu tests need a GPU"
    torch.cuda.set_device(0)
    return torch


@pytest.fixture(scope="module")
def eng(torch):
    from downloader_b200 import Engine
    e = Engine([0], ring_bytes=192 << 20)
    yield e
    e.close()


@pytest.fixture(scope="module")
def L():
    from downloader_b200 import _capi
    return _capi.lib()


def ref_digests(alg, bufs):
    return np.stack([np.frombuffer(hashlib.new(NAMES[alg], bytes(b)).digest(), np.uint8) for b in bufs])


def test_kat_vectors(eng, kat):
    msgs = [kat_message(v) for v in kat["vectors"]]
    out = eng.hash_batch(msgs, ALL)
    for i, v in enumerate(kat["vectors"]):
        for alg in (SHA1, MD5, SHA256):
            assert out[alg][i].tobytes().hex() == v[NAMES[alg]], (NAMES[alg], i)


def boundary_lengths():
    lens = list(range(0, 258))
    for k in range(8, 21):
        lens += [(1 << k) - 1, 1 << k, (1 << k) + 1]
    lens += [55, 56, 63, 64, 65, 119, 120, 4096 - 9, 4096 - 8, 65536 + 55, 65536 + 56]
    return lens


@pytest.mark.parametrize("alg", [SHA1, MD5, SHA256])
def test_every_boundary_length_vs_oracle(eng, alg):
    lens = boundary_lengths()
    rng = np.random.default_rng(5)
    bufs = [rng.integers(0, 256, n, dtype=np.uint8) for n in lens]
    out = eng.hash_batch(bufs, alg)[alg]
    # oracle = plain-C restatement over the same bytes; hashlib (OpenSSL) as a second witness
    flat = np.concatenate(bufs) if bufs else np.zeros(0, np.uint8)
    off = np.cumsum([0] + lens[:-1]).astype(np.uint64)
    exp = oracle.batch(alg, flat, off, np.array(lens, np.uint64), threads=4)
    bad = np.nonzero((out != exp).any(axis=1))[0]
    assert bad.size == 0, f"{NAMES[alg]} mismatch at lengths {[lens[i] for i in bad[:10]]}"
    assert (exp == ref_digests(alg, bufs)).all()


@pytest.mark.parametrize("ch,nst", VARIANTS)
def test_all_staging_variants(eng, L, ch, nst):
    ch0, st0 = C.c_int(), C.c_int()
    L.th_get_staging(1, C.byref(ch0), C.byref(st0))
    try:
        assert L.th_set_option(b"ragged_rows", 0) == 0  # the per-lane bulk-copy engine (no longer the default)
        assert L.th_set_staging(7, ch, nst) == 0
        lens = [0, 1, 15, 16, 17, 55, 56, 63, 64, 65, ch - 1, ch, ch + 1, ch + 55, ch + 56, 2 * ch, 3 * ch + 63,
                nst * ch, nst * ch + 1, (nst + 1) * ch - 8, 5000, 70001] + list(range(1000, 1040))
        bufs = [oracle.synth(i, n) for i, n in enumerate(lens)]
        out = eng.hash_batch(bufs, ALL)
        for alg in (SHA1, MD5, SHA256):
            assert (out[alg] == ref_digests(alg, bufs)).all(), (NAMES[alg], ch, nst)
    finally:
        assert L.th_set_staging(7, ch0.value, st0.value) == 0
        assert L.th_set_option(b"ragged_rows", 1) == 0


@pytest.mark.parametrize("variant", [0, 1, 2, 3, 4])
def test_all_row_staging_variants(eng, L, variant):
    """The ragged default: cooperative cp.async into swizzled row tiles, every compiled (tiles, stages) geometry."""
    k, nst = [(1, 3), (1, 4), (2, 2), (2, 3), (4, 2)][variant]
    ch = 128 * k
    try:
        assert L.th_set_option(b"rows_variant", variant) == 0
        lens = [0, 1, 15, 16, 17, 55, 56, 63, 64, 65, 119, 120, 127, 128, 129, ch - 1, ch, ch + 1, ch + 55, ch + 56,
                2 * ch, 3 * ch + 63, nst * ch, nst * ch + 1, (nst + 1) * ch - 8, 5000, 70001, 300001] + list(range(1000, 1040))
        bufs = [oracle.synth(i, n) for i, n in enumerate(lens)]
        out = eng.hash_batch(bufs, ALL)
        for alg in (SHA1, MD5, SHA256):
            wrong = np.nonzero((out[alg] != ref_digests(alg, bufs)).any(axis=1))[0]
            assert wrong.size == 0, (NAMES[alg], k, nst, [lens[i] for i in wrong[:6]])
    finally:
        assert L.th_set_option(b"rows_variant", 2) == 0


def test_synth_fill_matches_oracle_generator(torch, L):
    n, ln, stride = 37, 1003, 1008
    d = torch.zeros(n * stride, dtype=torch.uint8, device="cuda")
    assert L.th_synth_fill(d.data_ptr(), 5, n, ln, stride, oracle.DEFAULT_SEED, None) == 0
    torch.cuda.synchronize()
    h = d.cpu().numpy()
    for k in range(n):
        assert (h[k * stride:k * stride + ln] == oracle.synth(5 + k, ln)).all()
        assert (h[k * stride + ln:(k + 1) * stride] == 0).all()


@pytest.mark.parametrize("n,ln", [(1, 64), (33, 4096), (1000, 16384), (4096, 262144), (7, 16 << 20)])
def test_launch_level_uniform_batches(torch, L, n, ln):
    """Device-resident, stride-addressed launch (the shape of BASELINE configs 2 and 3)."""
    d = torch.empty(n * ln, dtype=torch.uint8, device="cuda")
    assert L.th_synth_fill(d.data_ptr(), 0, n, ln, ln, oracle.DEFAULT_SEED, None) == 0
    outs = {a: torch.zeros(n * oracle.DIGEST_LEN[a], dtype=torch.uint8, device="cuda") for a in (SHA1, MD5, SHA256)}
    assert L.th_sha1_batch(d.data_ptr(), None, None, None, ln, ln, n, outs[SHA1].data_ptr(), None, None, None) == 0
    assert L.th_md5_batch(d.data_ptr(), None, None, None, ln, ln, n, outs[MD5].data_ptr(), None) == 0
    assert L.th_sha256_batch(d.data_ptr(), None, None, None, ln, ln, n, outs[SHA256].data_ptr(), None) == 0
    torch.cuda.synchronize()
    # oracle on a bounded sample of messages, regenerated on the host from the generator definition
    pick = sorted(set(np.linspace(0, n - 1, min(n, 48)).astype(int).tolist()))
    host = np.concatenate([oracle.synth(i, ln) for i in pick])
    off = (np.arange(len(pick)) * ln).astype(np.uint64)
    lens = np.full(len(pick), ln, np.uint64)
    for a in (SHA1, MD5, SHA256):
        got = outs[a].cpu().numpy().reshape(n, -1)[pick]
        _, exp = oracle.evp_batch(a, host, off, lens, threads=8)
        assert (got == exp[a]).all(), NAMES[a]
        assert (oracle.batch(a, host, off, lens, threads=8) == exp[a]).all()


TILINGS = [(1, 2), (1, 3), (1, 4), (2, 2), (2, 3), (4, 2)]


@pytest.mark.parametrize("path", [1, 2, 3])
@pytest.mark.parametrize("k,nst", TILINGS)
def test_uniform_batch_paths_and_tilings(torch, L, path, k, nst):
    """Equal-length batches: per-lane bulk copies (1), tensor-map tiles with one-warp CTAs (2) and the
    persistent time-sliced kernel (3) must all give the oracle's digests, for every tiling, including
    lengths that end mid-tile / mid-block and group counts that do not divide by the worker count."""
    try:
        assert L.th_set_option(b"path", path) == 0
        assert L.th_set_tiling(7, k, nst) == 0
        cases = [(70, 1024), (33, 128 * k * nst + 57), (1500, 4096 + 119), (2100, 2048 + 64), (64, 65536 + 56)]
        if path == 3:
            assert L.th_set_option(b"workers", 8 if k * nst > 4 else 12) == 0
        for n, ln in cases:
            stride = (ln + 15) & ~15
            d = torch.zeros(n * stride, dtype=torch.uint8, device="cuda")
            assert L.th_synth_fill(d.data_ptr(), 0, n, ln, stride, 99, None) == 0
            outs = {a: torch.zeros(n * oracle.DIGEST_LEN[a], dtype=torch.uint8, device="cuda") for a in (SHA1, MD5, SHA256)}
            assert L.th_sha1_batch(d.data_ptr(), None, None, None, stride, ln, n, outs[SHA1].data_ptr(), None, None, None) == 0
            assert L.th_md5_batch(d.data_ptr(), None, None, None, stride, ln, n, outs[MD5].data_ptr(), None) == 0
            assert L.th_sha256_batch(d.data_ptr(), None, None, None, stride, ln, n, outs[SHA256].data_ptr(), None) == 0
            torch.cuda.synchronize()
            host = d.cpu().numpy()
            off = (np.arange(n) * stride).astype(np.uint64)
            lens = np.full(n, ln, np.uint64)
            for a in (SHA1, MD5, SHA256):
                exp = oracle.batch(a, host, off, lens, threads=8)
                got = outs[a].cpu().numpy().reshape(n, -1)
                bad = np.nonzero((got != exp).any(axis=1))[0]
                assert bad.size == 0, (NAMES[a], path, k, nst, n, ln, bad[:8])
    finally:
        L.th_set_option(b"path", 0)
        L.th_set_option(b"workers", 12)
        L.th_set_tiling(7, 1, 3)


@pytest.mark.parametrize("variant", [0, 1, 2, 3, 4, 5, 7])
def test_assist_kernel_schedule_offload(torch, L, variant):
    """Few large messages: chain warp + schedule-helper warp per group (path 4, SHA-1 / SHA-256; MD5 has no
    schedule and takes the tile kernel), in every compiled geometry (tiles, stages, ring slots; the default and variant
    5 hand over two blocks per slot, 7 is round 1's four slots of one block).  Lengths cover: fewer blocks than ring slots, odd block counts, exact block / tile /
    stage multiples, both padding-block cases, a ragged last group."""
    try:
        assert L.th_set_option(b"path", 4) == 0
        assert L.th_set_option(b"assist_variant", variant) == 0
        cases = [(5, 1024), (33, 1024 + 55), (40, 1024 + 56), (64, 2048), (9, 2048 + 64), (7, 4096 + 63), (100, 65536 + 64), (3, (1 << 20) + 9)]
        for n, ln in cases:
            stride = (ln + 15) & ~15
            d = torch.zeros(n * stride, dtype=torch.uint8, device="cuda")
            assert L.th_synth_fill(d.data_ptr(), 0, n, ln, stride, 4242, None) == 0
            outs = {a: torch.zeros(n * oracle.DIGEST_LEN[a], dtype=torch.uint8, device="cuda") for a in (SHA1, MD5, SHA256)}
            ok = torch.full(((n + 31) // 32,), -1, dtype=torch.int32, device="cuda")
            assert L.th_sha1_batch(d.data_ptr(), None, None, None, stride, ln, n, outs[SHA1].data_ptr(), None, None, None) == 0
            exp = outs[SHA1].clone()
            exp.view(n, 20)[n - 1, 2] ^= 8
            assert L.th_sha1_batch(d.data_ptr(), None, None, None, stride, ln, n, outs[SHA1].data_ptr(), exp.data_ptr(), ok.data_ptr(), None) == 0
            assert L.th_md5_batch(d.data_ptr(), None, None, None, stride, ln, n, outs[MD5].data_ptr(), None) == 0
            assert L.th_sha256_batch(d.data_ptr(), None, None, None, stride, ln, n, outs[SHA256].data_ptr(), None) == 0
            torch.cuda.synchronize()
            host = d.cpu().numpy()
            off = (np.arange(n) * stride).astype(np.uint64)
            lens = np.full(n, ln, np.uint64)
            for a in (SHA1, MD5, SHA256):
                e = oracle.batch(a, host, off, lens, threads=8)
                assert (outs[a].cpu().numpy().reshape(n, -1) == e).all(), (NAMES[a], n, ln)
            words = ok.cpu().numpy().view(np.uint32)
            bits = ((words[np.arange(n) >> 5] >> (np.arange(n) & 31)) & 1).astype(bool)
            assert np.nonzero(~bits)[0].tolist() == [n - 1], (n, ln)
    finally:
        L.th_set_option(b"path", 0)
        L.th_set_option(b"assist_variant", 0)


@pytest.mark.parametrize("n,ln", [(1, 1024), (5, 4096 + 55), (33, 100000 + 56), (64, (1 << 20) + 63), (40, 262144), (3, 1 << 20), (7, 16 << 20)])
def test_md5_sha256_one_launch_kernel_vs_hashlib(torch, L, n, ln):
    """th_md5_sha256_batch: MD5 + SHA-256 of few large equal-length messages in ONE launch (digest_parts_kernel: SHA-256
    chain warp, schedule helper warp and an MD5 chain warp per group, each on a sub-partition of its own) — and, with
    the option off, as the two separate kernels.  Tail lengths cover one and two padding blocks, where MD5 and SHA-256
    pad differently (little- vs big-endian bit length); rows past n are zero-filled by the TMA unit."""
    stride = (ln + 15) & ~15
    rng = np.random.default_rng(n * 7919 + ln)
    host = rng.integers(0, 256, n * stride, dtype=np.uint8)
    d = torch.from_numpy(host).cuda()
    bufs = [host[i * stride:i * stride + ln] for i in range(n)]
    try:
        for fuse in (1, 0):
            assert L.th_set_option(b"fuse_md5_sha256", fuse) == 0
            assert L.th_md5_sha256_fused(stride, ln, n) == fuse
            o5 = torch.full((n * 16,), 0xEE, dtype=torch.uint8, device="cuda")
            o256 = torch.full((n * 32,), 0xEE, dtype=torch.uint8, device="cuda")
            assert L.th_md5_sha256_batch(d.data_ptr(), stride, ln, n, o5.data_ptr(), o256.data_ptr(), None) == 0
            torch.cuda.synchronize()
            assert (o5.cpu().numpy().reshape(n, 16) == ref_digests(MD5, bufs)).all(), fuse
            assert (o256.cpu().numpy().reshape(n, 32) == ref_digests(SHA256, bufs)).all(), fuse
    finally:
        assert L.th_set_option(b"fuse_md5_sha256", 1) == 0


def test_md5_sha256_one_launch_kernel_two_groups_per_cta(torch, L):
    """More groups than SMs (149..296): digest_parts_kernel runs two groups per CTA — two SHA-256 chains, their two
    helpers and two MD5 chains in eight warp slots.  5,000 x 64 KiB-ish messages = 157 groups, the last one ragged."""
    n, ln = 5000, 65536 + 4096 + 57
    stride = (ln + 15) & ~15
    host = np.random.default_rng(157).integers(0, 256, n * stride, dtype=np.uint8)
    d = torch.from_numpy(host).cuda()
    assert L.th_md5_sha256_fused(stride, ln, n) == 1
    o5 = torch.zeros(n * 16, dtype=torch.uint8, device="cuda")
    o256 = torch.zeros(n * 32, dtype=torch.uint8, device="cuda")
    assert L.th_md5_sha256_batch(d.data_ptr(), stride, ln, n, o5.data_ptr(), o256.data_ptr(), None) == 0
    torch.cuda.synchronize()
    bufs = [host[i * stride:i * stride + ln] for i in range(n)]
    assert (o5.cpu().numpy().reshape(n, 16) == ref_digests(MD5, bufs)).all()
    assert (o256.cpu().numpy().reshape(n, 32) == ref_digests(SHA256, bufs)).all()


def test_submit_md5_and_sha256_of_parts_share_one_launch(torch, L):
    """th_submit with MD5 | SHA-256 over an object's parts (equal-length prefix + short last part) launches one kernel
    per item — 2 instead of 4; SHA-1 of the same slot still has its own.  With the option off: 4."""
    from downloader_b200 import Engine
    part, last = 4 << 20, (1 << 20) + 12345
    nparts = 6
    total = (nparts - 1) * part + last
    rng = np.random.default_rng(99)
    data = rng.integers(0, 256, total, dtype=np.uint8)
    offs = np.arange(nparts, dtype=np.uint64) * part
    lens = np.array([part] * (nparts - 1) + [last], np.uint64)
    bufs = [data[int(o):int(o + l)] for o, l in zip(offs, lens)]
    with Engine([0], ring_bytes=64 << 20) as e:
        e.set_timing(True)
        slot, view = e.ring_acquire(total)
        view[:] = data
        s0 = e.stats()
        t = e.submit(slot, offs, lens, MD5 | SHA256)
        tm = e.timing(t)
        r = e.result(t, nparts, MD5 | SHA256)
        s1 = e.stats()
        assert (r[MD5] == ref_digests(MD5, bufs)).all() and (r[SHA256] == ref_digests(SHA256, bufs)).all()
        assert s1["launches"] - s0["launches"] == 2
        assert tm[3] >= 0 and tm[4] > tm[3] and tm[6] > tm[5]  # the MD5 timeline is the shared launches'
        slot, view = e.ring_acquire(total)
        view[:] = data
        t = e.submit(slot, offs, lens, SHA1 | MD5 | SHA256)
        r = e.result(t, nparts, SHA1 | MD5 | SHA256)
        assert (r[SHA1] == ref_digests(SHA1, bufs)).all() and (r[MD5] == ref_digests(MD5, bufs)).all()
        assert e.stats()["launches"] - s1["launches"] == 4
    assert L.th_set_option(b"fuse_md5_sha256", 0) == 0
    try:
        with Engine([0], ring_bytes=64 << 20) as e:  # separate MD5 and SHA-256 kernels on separate streams
            slot, view = e.ring_acquire(total)
            view[:] = data
            s0 = e.stats()
            r = e.result(e.submit(slot, offs, lens, MD5 | SHA256), nparts, MD5 | SHA256)
            assert (r[MD5] == ref_digests(MD5, bufs)).all() and (r[SHA256] == ref_digests(SHA256, bufs)).all()
            assert e.stats()["launches"] - s0["launches"] == 4
    finally:
        assert L.th_set_option(b"fuse_md5_sha256", 1) == 0


def test_sliced_kernel_time_slicing_engaged(torch, L):
    """Config-2-like shape scaled down: ~14 groups per SM on 12 workers -> chains are cut into segments
    and handed between warps; result must not change."""
    n, ln = 148 * 14 * 32 - 40, 8192 + 64
    d = torch.empty(n * ln, dtype=torch.uint8, device="cuda")
    L.th_synth_fill(d.data_ptr(), 0, n, ln, ln, 5, None)
    try:
        res = {}
        for path in (2, 3):
            assert L.th_set_option(b"path", path) == 0
            o = torch.zeros(n * 20, dtype=torch.uint8, device="cuda")
            ok = torch.full(((n + 31) // 32,), -1, dtype=torch.int32, device="cuda")
            assert L.th_sha1_batch(d.data_ptr(), None, None, None, ln, ln, n, o.data_ptr(), None, None, None) == 0
            exp = o.clone()
            exp.view(n, 20)[n - 1, 0] ^= 1
            exp.view(n, 20)[12345, 19] ^= 1
            assert L.th_sha1_batch(d.data_ptr(), None, None, None, ln, ln, n, o.data_ptr(), exp.data_ptr(), ok.data_ptr(), None) == 0
            torch.cuda.synchronize()
            res[path] = (o.cpu().numpy(), ok.cpu().numpy().view(np.uint32))
        assert (res[2][0] == res[3][0]).all() and (res[2][1] == res[3][1]).all()
        words = res[3][1]
        bits = ((words[np.arange(n) >> 5] >> (np.arange(n) & 31)) & 1).astype(bool)
        assert sorted(np.nonzero(~bits)[0].tolist()) == [12345, n - 1]
        pick = [0, 1, 31, 32, 5000, n - 1]
        host = np.concatenate([oracle.synth(i, ln, seed=5) for i in pick])
        e = oracle.batch(SHA1, host, (np.arange(len(pick)) * ln).astype(np.uint64), np.full(len(pick), ln, np.uint64))
        assert (res[3][0].reshape(n, 20)[pick] == e).all()
    finally:
        L.th_set_option(b"path", 0)


def test_fused_verify_bitfield_identity_order(torch, L):
    """webtorrent's `hash === _hashes[i]` -> bitfield, fused into the SHA-1 kernel (uniform batch)."""
    n, ln = 2000, 16384
    d = torch.empty(n * ln, dtype=torch.uint8, device="cuda")
    L.th_synth_fill(d.data_ptr(), 0, n, ln, ln, 7, None)
    out = torch.zeros(n * 20, dtype=torch.uint8, device="cuda")
    assert L.th_sha1_batch(d.data_ptr(), None, None, None, ln, ln, n, out.data_ptr(), None, None, None) == 0
    exp = out.clone()
    bad = [0, 31, 32, 777, 1999]
    e2 = exp.view(n, 20)
    for b in bad:
        e2[b, (b * 7) % 20] ^= 0x40
    ok = torch.full(((n + 31) // 32,), -1, dtype=torch.int32, device="cuda")
    assert L.th_sha1_batch(d.data_ptr(), None, None, None, ln, ln, n, out.data_ptr(), exp.data_ptr(), ok.data_ptr(), None) == 0
    torch.cuda.synchronize()
    words = ok.cpu().numpy().view(np.uint32)
    bits = ((words[np.arange(n) >> 5] >> (np.arange(n) & 31)) & 1).astype(bool)
    assert sorted(np.nonzero(~bits)[0].tolist()) == bad
    # bits past n in the last word are zero
    assert (int(words[-1]) >> (n & 31)) == 0 if n & 31 else True
    # and the digests themselves are right
    host = oracle.synth_batch(0, 8, ln, seed=7)
    o8 = oracle.batch(SHA1, host, np.arange(8, dtype=np.uint64) * ln, np.full(8, ln, np.uint64))
    assert (out.cpu().numpy().reshape(n, 20)[:8] == o8).all()


def test_fused_verify_through_submit_mixed_lengths(eng):
    """Same through th_submit with ragged pieces (short last piece) -> length-bucketed order path."""
    rng = np.random.default_rng(3)
    lens = [262144] * 40 + [1000, 0, 64, 262144 - 1, 77777] + [16384] * 30
    bufs = [rng.integers(0, 256, n, dtype=np.uint8) for n in lens]
    exp = ref_digests(SHA1, bufs).copy()
    bad = [2, 41, 44, 74]
    for b in bad:
        exp[b, 19] ^= 1
    out = eng.hash_batch(bufs, SHA1 | MD5, expected_sha1=exp)
    assert sorted(np.nonzero(~out["ok"])[0].tolist()) == bad
    assert (out[SHA1] == ref_digests(SHA1, bufs)).all()
    assert (out[MD5] == ref_digests(MD5, bufs)).all()


def test_submit_uniform_prefix_plus_short_last_piece(eng):
    """A torrent's pieces: all equal but the last.  th_submit sends the equal-length prefix to the
    tensor-map kernels and the remainder to the per-lane path; digests and bitfield must line up."""
    rng = np.random.default_rng(17)
    for n_full, ln, tail in [(100, 8192, [100]), (64, 16384, []), (333, 4096, [4095, 17, 0]), (96, 2048, [2048 + 64])]:
        lens = [ln] * n_full + tail
        bufs = [rng.integers(0, 256, x, dtype=np.uint8) for x in lens]
        exp = ref_digests(SHA1, bufs).copy()
        bad = sorted({0, 31, 32, n_full - 1, len(lens) - 1})
        for b in bad:
            exp[b, 7] ^= 2
        out = eng.hash_batch(bufs, SHA1 | MD5 | SHA256, expected_sha1=exp)
        assert sorted(np.nonzero(~out["ok"])[0].tolist()) == bad, (n_full, ln, tail)
        for a in (SHA1, MD5, SHA256):
            assert (out[a] == ref_digests(a, bufs)).all(), (NAMES[a], n_full, ln, tail)


def test_etag_combine_vs_oracle(eng, kat):
    rng = np.random.default_rng(9)
    counts = [1, 2, 3, 4, 5, 7, 8, 9, 64, 1000, 10000, 0]
    md5s = rng.integers(0, 256, (sum(counts), 16), dtype=np.uint8)
    out = eng.etag_combine(md5s, counts)
    pos = 0
    for j, c in enumerate(counts):
        exp = oracle.etag_multipart(md5s[pos:pos + c])
        assert out[j].tobytes().hex() + f"-{c}" == exp
        assert out[j].tobytes() == hashlib.md5(md5s[pos:pos + c].tobytes()).digest()
        pos += c
    # golden framing vector: parts {5 MiB 'a', 5 MiB 'b', 1 'c'}
    fr = kat["framing"]
    parts = [bytes([p["byte"]]) * p["count"] for p in fr["etag_parts"]]
    d = eng.hash_batch(parts, MD5)[MD5]
    assert eng.etag_combine(d, [3])[0].tobytes().hex() + "-3" == fr["etag"]


def test_ring_many_slots_in_flight_and_wrap(torch):
    from downloader_b200 import Engine
    with Engine([0], ring_bytes=8 << 20) as e:
        e.set_timing(True)  # stats()['kernel_ms'] is only kept with ticket timing on
        ln = 1 << 20
        tickets = []
        for k in range(20):  # 20 MiB through an 8 MiB ring: forces wrap + waits on the oldest slot
            slot, view = e.ring_acquire(ln)
            view[:] = oracle.synth(k, ln)
            offs = np.arange(0, ln, 65536, dtype=np.uint64)
            tickets.append((k, e.submit(slot, offs, np.full(offs.size, 65536, np.uint64), SHA1 | SHA256)))
        for k, t in tickets:
            r = e.result(t, 16, SHA1 | SHA256)
            data = oracle.synth(k, ln)
            bufs = [data[i * 65536:(i + 1) * 65536] for i in range(16)]
            assert (r[SHA1] == ref_digests(SHA1, bufs)).all()
            assert (r[SHA256] == ref_digests(SHA256, bufs)).all()
        st = e.stats()
        assert st["submits"] == 20 and st["launches"] == 40 and st["bytes"] == 20 * ln
        assert st["h2d_bytes"] >= 20 * ln and st["kernel_ms"] > 0


def test_completion_callback(eng):
    import threading
    ev = threading.Event()
    seen = {}
    slot, view = eng.ring_acquire(4096)
    view[:] = 0
    t = eng.submit(slot, [0], [4096], MD5, callback=lambda ticket, status: (seen.update(t=ticket, s=status), ev.set()))
    assert ev.wait(30)
    assert seen == {"t": t, "s": 0}
    r = eng.result(t, 1, MD5)
    assert r[MD5][0].tobytes() == hashlib.md5(bytes(4096)).digest()


def test_errors_are_codes_not_crashes(eng):
    from downloader_b200 import TritonHashError
    slot, view = eng.ring_acquire(1024)
    with pytest.raises(TritonHashError) as ei:
        eng.submit(slot, [8], [16], SHA1)  # misaligned offset
    assert ei.value.code == -5
    with pytest.raises(TritonHashError) as ei:
        eng.submit(slot, [0], [4096], SHA1)  # beyond the slot
    assert ei.value.code == -1
    eng.ring_release(slot)
    with pytest.raises(TritonHashError) as ei:
        eng.wait(123456789)
    assert ei.value.code == -8
    with pytest.raises(TritonHashError):
        eng.ring_acquire(1 << 40)  # TH_ETOOBIG


def test_config1_shape_one_gib_file_as_64_parts(torch, L, eng):
    """BASELINE config 1 shape: 1 GiB file = 64 x 16 MiB parts, MD5 + SHA-256 + multipart ETag,
    every part checked against libcrypto (the code path Node crypto runs)."""
    n, ln = 64, 16 << 20
    d = torch.empty(n * ln, dtype=torch.uint8, device="cuda")
    L.th_synth_fill(d.data_ptr(), 0, n, ln, ln, oracle.DEFAULT_SEED, None)
    offs = np.arange(n, dtype=np.uint64) * ln
    t = eng.submit_device(d.data_ptr(), offs, np.full(n, ln, np.uint64), MD5 | SHA256)
    r = eng.result(t, n, MD5 | SHA256)
    host = d.cpu().numpy()
    _, exp = oracle.evp_batch(MD5 | SHA256, host, offs, np.full(n, ln, np.uint64), threads=16, chunk=65536)
    assert (r[MD5] == exp[MD5]).all() and (r[SHA256] == exp[SHA256]).all()
    assert eng.etag_combine(r[MD5], [n])[0].tobytes().hex() + "-64" == oracle.etag_multipart(exp[MD5])


def _cpu_digests_of_device_batch(torch, d, n, ln, mask, chunk_msgs, update=0):
    """libcrypto digests of every message of a device-resident equal-length batch, staged through pinned memory."""
    stage = torch.empty(chunk_msgs * ln, dtype=torch.uint8).pin_memory()
    outs = {a: np.empty((n, oracle.DIGEST_LEN[a]), np.uint8) for a in (SHA1, MD5, SHA256) if mask & a}
    for i in range(0, n, chunk_msgs):
        k = min(chunk_msgs, n - i)
        stage[:k * ln].copy_(d[i * ln:(i + k) * ln])
        torch.cuda.synchronize()
        _, o = oracle.evp_batch(mask, stage.numpy()[:k * ln], np.arange(k, dtype=np.uint64) * ln, np.full(k, ln, np.uint64),
                                threads=min(32, __import__("os").cpu_count() or 1), chunk=update)
        for a in outs:
            outs[a][i:i + k] = o[a]
    return outs


def test_config2_full_size_every_piece_vs_libcrypto(torch, L):
    """BASELINE config 2 at full size: 65,536 x 256 KiB (16 GiB) SHA-1 + fused verify.  EVERY digest is compared
    with libcrypto, and the expected table of the fused compare is libcrypto's (with K pieces corrupted), not the
    kernel's own output: the bitfield must flag exactly the corrupted pieces."""
    n, ln = 65536, 262144
    free, _ = torch.cuda.mem_get_info()
    if free < n * ln + (1 << 30):
        pytest.skip("not enough free HBM")
    d = torch.empty(n * ln, dtype=torch.uint8, device="cuda")
    assert L.th_synth_fill(d.data_ptr(), 0, n, ln, ln, oracle.DEFAULT_SEED, None) == 0
    torch.cuda.synchronize()
    for i in (0, 1, 4097, 65535):  # the device bytes are the generator's bytes
        assert (d[i * ln:(i + 1) * ln].cpu().numpy() == oracle.synth(i, ln)).all()
    cpu = _cpu_digests_of_device_batch(torch, d, n, ln, SHA1, 4096)[SHA1]
    exp_host = cpu.copy()
    bad = [0, 1, 4097, 32768, 65535]
    for b in bad:
        exp_host[b, 3] ^= 0x80
    exp = torch.from_numpy(exp_host).cuda()
    out1 = torch.zeros(n * 20, dtype=torch.uint8, device="cuda")
    out2 = torch.zeros(n * 20, dtype=torch.uint8, device="cuda")
    ok = torch.zeros(n // 32, dtype=torch.int32, device="cuda")
    assert L.th_sha1_batch(d.data_ptr(), None, None, None, ln, ln, n, out1.data_ptr(), None, None, None) == 0
    assert L.th_sha1_batch(d.data_ptr(), None, None, None, ln, ln, n, out2.data_ptr(), exp.data_ptr(), ok.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert torch.equal(out1, out2)  # determinism; the fused compare does not disturb the digests
    digs = out1.cpu().numpy().reshape(n, 20)
    assert (digs == cpu).all(), "GPU SHA-1 differs from libcrypto on config 2"
    words = ok.cpu().numpy().view(np.uint32)
    bits = ((words[np.arange(n) >> 5] >> (np.arange(n) & 31)) & 1).astype(bool)
    assert sorted(np.nonzero(~bits)[0].tolist()) == bad


def test_config3_full_shape_every_part_vs_libcrypto(torch, L):
    """BASELINE config 3's per-GPU shape at full size: 4,096 x 16 MiB parts (64 GiB), MD5 + SHA-256 on concurrent
    streams as the runtime launches them; all 8,192 digests against libcrypto fed 64 KiB updates (minio's chunks)."""
    n, ln = 4096, 16 << 20
    free, _ = torch.cuda.mem_get_info()
    if free < n * ln + (2 << 30):
        pytest.skip("not enough free HBM")
    d = torch.empty(n * ln, dtype=torch.uint8, device="cuda")
    assert L.th_synth_fill(d.data_ptr(), 0, n, ln, ln, oracle.DEFAULT_SEED, None) == 0
    o5 = torch.zeros(n * 16, dtype=torch.uint8, device="cuda")
    o256 = torch.zeros(n * 32, dtype=torch.uint8, device="cuda")
    s2 = torch.cuda.Stream()
    s2.wait_stream(torch.cuda.current_stream())
    assert L.th_md5_batch(d.data_ptr(), None, None, None, ln, ln, n, o5.data_ptr(), C.c_void_p(s2.cuda_stream)) == 0
    assert L.th_sha256_batch(d.data_ptr(), None, None, None, ln, ln, n, o256.data_ptr(), None) == 0
    cpu = _cpu_digests_of_device_batch(torch, d, n, ln, MD5 | SHA256, 64, update=65536)  # runs while the kernels do
    torch.cuda.synchronize()
    assert (o5.cpu().numpy().reshape(n, 16) == cpu[MD5]).all(), "GPU MD5 differs from libcrypto on config 3"
    assert (o256.cpu().numpy().reshape(n, 32) == cpu[SHA256]).all(), "GPU SHA-256 differs from libcrypto on config 3"


def test_messages_longer_than_512_mib_and_2_gib(torch, L):
    """64-bit length handling: the bit-length field needs its high word from 512 MiB up (tensor-map kernels),
    and from 2 GiB up a message leaves the tensor-map path (per-lane bulk copies, 64-bit byte offsets)."""
    free, _ = torch.cuda.mem_get_info()
    if free < (6 << 30):
        pytest.skip("not enough free HBM")
    ln = (1 << 29) + 64 + 1
    n = 2
    stride = (ln + 15) & ~15
    d = torch.empty(n * stride, dtype=torch.uint8, device="cuda")
    assert L.th_synth_fill(d.data_ptr(), 0, n, ln, stride, 77, None) == 0
    outs = {a: torch.zeros(n * oracle.DIGEST_LEN[a], dtype=torch.uint8, device="cuda") for a in (SHA1, MD5, SHA256)}
    assert L.th_sha1_batch(d.data_ptr(), None, None, None, stride, ln, n, outs[SHA1].data_ptr(), None, None, None) == 0
    assert L.th_md5_batch(d.data_ptr(), None, None, None, stride, ln, n, outs[MD5].data_ptr(), None) == 0
    assert L.th_sha256_batch(d.data_ptr(), None, None, None, stride, ln, n, outs[SHA256].data_ptr(), None) == 0
    torch.cuda.synchronize()
    host = d.cpu().numpy()
    for a in (SHA1, MD5, SHA256):
        got = outs[a].cpu().numpy().reshape(n, -1)
        for i in range(n):
            assert got[i].tobytes() == hashlib.new(NAMES[a], host[i * stride:i * stride + ln]).digest(), (NAMES[a], i)
    del d, host
    ln = (1 << 31) + 16 + 5   # beyond the tensor map's 2^31 coordinate range
    d = torch.empty(ln, dtype=torch.uint8, device="cuda")
    assert L.th_synth_fill(d.data_ptr(), 3, 1, ln, ln, 78, None) == 0
    o = torch.zeros(16, dtype=torch.uint8, device="cuda")
    assert L.th_md5_batch(d.data_ptr(), None, None, None, ln, ln, 1, o.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert o.cpu().numpy().tobytes() == hashlib.md5(d.cpu().numpy()).digest()


def test_ticket_timing_is_a_consistent_device_timeline(eng):
    """th_ticket_timing: inputs ready <= kernel start <= kernel end for requested digests, -1 for the others."""
    ln = 1 << 20
    offs, lens = np.arange(4, dtype=np.uint64) * ln, np.full(4, ln, np.uint64)
    try:
        eng.set_timing(True)
        slot, view = eng.ring_acquire(4 * ln)
        view[:] = 7
        k0 = eng.stats()["kernel_ms"]
        t = eng.submit(slot, offs, lens, SHA1 | SHA256)
        tm = eng.timing(t)
        assert tm[0] >= 0 and tm[0] <= tm[1] <= tm[2] and tm[0] <= tm[5] <= tm[6]
        assert tm[3] == -1 and tm[4] == -1
        r = eng.result(t, 4, SHA1 | SHA256)
        assert r[SHA1][0].tobytes() == hashlib.sha1(bytes([7]) * ln).digest()
        assert eng.stats()["kernel_ms"] > k0
    finally:
        eng.set_timing(False)
    # the default: no timing events on the streams, every entry -1, digests unaffected
    slot, view = eng.ring_acquire(4 * ln)
    view[:] = 9
    t = eng.submit(slot, offs, lens, MD5)
    assert eng.timing(t) == [-1.0] * 7
    assert eng.result(t, 4, MD5)[MD5][3].tobytes() == hashlib.md5(bytes([9]) * ln).digest()


def test_few_large_ragged_messages_take_lone_launches(eng):
    """A handful of large bodies of unlike sizes (whole-file digests of files below the part size): each is launched
    as an equal-length batch of one (TMA + helper-warp path) on its own stream; with a verify table the SHA-1 goes
    through the per-lane kernel instead.  Results must not depend on the route."""
    lens = [300 * 1024 + 5, 1 << 20, 700 * 1024 + 63, (2 << 20) + 1, 256 * 1024, 3 * (1 << 20) + 56]
    bufs = [oracle.synth(40 + i, n) for i, n in enumerate(lens)]
    s0 = eng.stats()
    out = eng.hash_batch(bufs, ALL)
    s1 = eng.stats()
    for a in (SHA1, MD5, SHA256):
        assert (out[a] == ref_digests(a, bufs)).all(), NAMES[a]
    assert s1["launches"] - s0["launches"] == 2 * len(lens)  # per message: SHA-1, and one launch for MD5 + SHA-256
    exp = ref_digests(SHA1, bufs).copy()
    exp[2, 0] ^= 1
    out = eng.hash_batch(bufs, SHA1 | MD5, expected_sha1=exp)
    assert out["ok"].tolist() == [True, True, False, True, True, True]
    assert (out[SHA1] == ref_digests(SHA1, bufs)).all() and (out[MD5] == ref_digests(MD5, bufs)).all()


@pytest.mark.parametrize("seed", [1, 2, 3, 4, 5, 6])
def test_randomised_batches_vs_hashlib(eng, seed):
    """Differential test over random batch shapes: message counts, length mixes (empty, tiny, block-edge, large),
    equal-length runs followed by odd tails, random digest masks and corrupted verify tables."""
    rng = np.random.default_rng(1000 + seed)
    for _ in range(6):
        kind = int(rng.integers(0, 4))
        if kind == 0:      # ragged small
            lens = rng.integers(0, 3000, int(rng.integers(1, 200))).tolist()
        elif kind == 1:    # equal-length run + odd tail(s)
            base = int(rng.choice([1024, 4096, 16384, 65536 + 16, 262144])) + int(rng.integers(0, 2)) * 16
            lens = [base] * int(rng.integers(1, 150)) + rng.integers(0, base, int(rng.integers(0, 4))).tolist()
        elif kind == 2:    # block-edge lengths
            lens = [int(64 * k + d) for k in rng.integers(0, 40, 30) for d in (rng.integers(53, 67),)]
        else:              # a few large bodies of unlike sizes
            lens = rng.integers(256 * 1024, 900 * 1024, int(rng.integers(1, 7))).tolist()
        bufs = [rng.integers(0, 256, int(n), dtype=np.uint8) for n in lens]
        algs = int(rng.integers(1, 8))
        exp = None
        bad = []
        if algs & SHA1 and rng.integers(0, 2):
            exp = ref_digests(SHA1, bufs).copy()
            bad = sorted(set(rng.integers(0, len(lens), int(rng.integers(0, 4))).tolist()))
            for b in bad:
                exp[b, int(rng.integers(0, 20))] ^= 1 << int(rng.integers(0, 8))
        out = eng.hash_batch(bufs, algs, expected_sha1=exp)
        for a in (SHA1, MD5, SHA256):
            if algs & a:
                got, ref = out[a], ref_digests(a, bufs)
                wrong = np.nonzero((got != ref).any(axis=1))[0]
                assert wrong.size == 0, (seed, kind, NAMES[a], [lens[i] for i in wrong[:5]])
        if exp is not None:
            assert sorted(np.nonzero(~out["ok"])[0].tolist()) == bad, (seed, kind, lens[:8])


def test_small_pinned_ring_large_device_ring(torch):
    """th_init_staged: pinned bytes are reusable once copied, device bytes once hashed.  48 slots of unlike,
    odd sizes stream through a 3 MiB pinned ring backed by a 64 MiB device ring with no result read until the
    end; each slot's pinned bytes are overwritten by later slots long before its digests are collected."""
    from downloader_b200 import Engine
    rng = np.random.default_rng(77)
    with Engine([0], ring_bytes=3 << 20, device_bytes=64 << 20) as e:
        assert e.device_bytes == 64 << 20
        jobs = []
        for k in range(48):
            nbytes = int(rng.integers(1, 1 << 20))
            slot, view = e.ring_acquire(nbytes)
            data = oracle.synth(1000 + k, nbytes)
            view[:] = data
            unit = int(rng.choice([4096, 65536, 100000]))
            offs = np.arange(0, nbytes, unit, dtype=np.uint64)
            lens = np.minimum(unit, nbytes - offs).astype(np.uint64)
            t1 = e.submit(slot, offs, lens, SHA1)
            t2 = e.submit(slot, offs, lens, MD5 | SHA256)  # same slot again: no second copy
            jobs.append((data, offs, lens, t1, t2))
        st = e.stats()
        assert st["h2d_bytes"] < sum(j[0].size for j in jobs) + 48 * 2 * (1 << 16)  # one payload copy per slot
        for data, offs, lens, t1, t2 in jobs:
            bufs = [data[int(o):int(o + l)] for o, l in zip(offs, lens)]
            r1 = e.result(t1, offs.size, SHA1)
            r2 = e.result(t2, offs.size, MD5 | SHA256)
            assert (r1[SHA1] == ref_digests(SHA1, bufs)).all()
            assert (r2[MD5] == ref_digests(MD5, bufs)).all()
            assert (r2[SHA256] == ref_digests(SHA256, bufs)).all()


def test_device_ring_smaller_than_the_stream_blocks_and_recovers(torch):
    """Device ring of two slots: the third acquire waits for the oldest slot's tickets; an unsubmitted oldest
    slot makes it TH_EBUSY instead of a deadlock; oversize requests are TH_ETOOBIG for either ring."""
    from downloader_b200 import Engine
    from downloader_b200 import TritonHashError
    ln = 1 << 20
    offs = np.arange(0, ln, 1 << 18, dtype=np.uint64)
    lens = np.full(offs.size, 1 << 18, np.uint64)
    with Engine([0], ring_bytes=4 << 20, device_bytes=2 << 20) as e:
        with pytest.raises(TritonHashError) as ei:
            e.ring_acquire(3 << 20)  # fits the pinned ring, not the device ring
        assert ei.value.code == -7
        out = []
        for k in range(6):
            slot, view = e.ring_acquire(ln)
            view[:] = oracle.synth(50 + k, ln)
            out.append((k, e.submit(slot, offs, lens, SHA256)))
        for k, t in out:
            data = oracle.synth(50 + k, ln)
            assert (e.result(t, 4, SHA256)[SHA256] == ref_digests(SHA256, [data[i << 18:(i + 1) << 18] for i in range(4)])).all()
        a, _ = e.ring_acquire(ln)
        b, vb = e.ring_acquire(ln)
        with pytest.raises(TritonHashError) as ei:
            e.ring_acquire(ln)  # device ring full and slot `a` was never submitted
        assert ei.value.code == -6
        assert e.ring_acquire(ln, wait=False) is None  # th_ring_try_acquire: same condition, reported without waiting
        e.ring_release(a)
        vb[:] = 7
        t = e.submit(b, offs, lens, MD5)
        c, _ = e.ring_acquire(ln)  # fits again
        e.ring_release(c)
        assert (e.result(t, 4, MD5)[MD5] == ref_digests(MD5, [np.full(1 << 18, 7, np.uint8)] * 4)).all()


def test_one_context_many_caller_threads(torch):
    """th_* on one context are serialised by the library (tritonhash.h, threading): four caller threads acquire,
    fill, submit and collect on the same context at once; every digest must match hashlib."""
    import threading
    from downloader_b200 import Engine
    errors = []
    with Engine([0], ring_bytes=8 << 20, device_bytes=64 << 20) as e:
        def worker(tid):
            try:
                rng = np.random.default_rng(500 + tid)
                for it in range(12):
                    nbytes = int(rng.integers(16, 512 << 10))
                    unit = int(rng.choice([1024, 16384, 65536]))
                    data = rng.integers(0, 256, nbytes, dtype=np.uint8)
                    while True:
                        try:
                            slot, view = e.ring_acquire(nbytes)
                            break
                        except Exception as ex:  # TH_EBUSY: the oldest slot belongs to a thread that has not submitted yet
                            if getattr(ex, "code", 0) != -6:
                                raise
                    view[:] = data
                    offs = np.arange(0, nbytes, unit, dtype=np.uint64)
                    lens = np.minimum(unit, nbytes - offs).astype(np.uint64)
                    algs = int(rng.integers(1, 8))
                    t = e.submit(slot, offs, lens, algs)
                    r = e.result(t, offs.size, algs)
                    bufs = [data[int(o):int(o + l)] for o, l in zip(offs, lens)]
                    for a in (SHA1, MD5, SHA256):
                        if algs & a and not (r[a] == ref_digests(a, bufs)).all():
                            errors.append((tid, it, NAMES[a]))
            except Exception as ex:  # noqa: BLE001
                errors.append((tid, repr(ex)))
        th = [threading.Thread(target=worker, args=(i,)) for i in range(4)]
        for x in th:
            x.start()
        for x in th:
            x.join()
        assert not errors, errors[:4]
        assert e.stats()["submits"] == 48


def test_heavy_tailed_batch_peels_outliers_into_lone_launches(eng):
    """Many small messages plus a few far larger ones (>= 256 KiB, >= 8 x the median): the outliers leave the
    per-lane launch for lone TMA launches (runtime.cu, launch items); digests and launch count say so."""
    rng = np.random.default_rng(31)
    lens = rng.integers(1, 20000, 500).tolist()
    big = [300_000, 1_000_003, 262_144, 777_777, 2_500_000]
    for k, b in enumerate(big):
        lens.insert(37 + 91 * k, b)
    bufs = [rng.integers(0, 256, int(n), dtype=np.uint8) for n in lens]
    s0 = eng.stats()
    out = eng.hash_batch(bufs, SHA1 | MD5 | SHA256)
    s1 = eng.stats()
    for a in (SHA1, MD5, SHA256):
        wrong = np.nonzero((out[a] != ref_digests(a, bufs)).any(axis=1))[0]
        assert wrong.size == 0, (NAMES[a], [lens[i] for i in wrong[:5]])
    assert s1["launches"] - s0["launches"] == 3 + 2 * len(big)  # per-lane remainder x 3 digests; each outlier: SHA-1 + one MD5/SHA-256 launch
    # with a verify table the remainder stays one launch (the bitfield word is written by one ballot)
    exp = ref_digests(SHA1, bufs).copy()
    exp[40, 3] ^= 0x10
    s0 = eng.stats()
    out = eng.hash_batch(bufs, SHA1, expected_sha1=exp)
    assert eng.stats()["launches"] - s0["launches"] == 1
    assert np.nonzero(~out["ok"])[0].tolist() == [40]
