"""Pins the CPU oracle (oracle/) before anything trusts it: golden KATs, libcrypto, hashlib."""
import hashlib

import numpy as np
import pytest

import oracle
from conftest import kat_message

ALGS = [(oracle.SHA1, "sha1"), (oracle.MD5, "md5"), (oracle.SHA256, "sha256")]


def test_kat_vectors_port_and_evp(kat):
    for v in kat["vectors"]:
        m = kat_message(v)
        for alg, name in ALGS:
            assert oracle.digest(alg, m).hex() == v[name], (name, v.get("msg", v.get("count")))
            assert oracle.evp(alg, m).hex() == v[name]
            assert oracle.evp(alg, m, chunk=65536).hex() == v[name]


@pytest.mark.parametrize("alg,name", ALGS)
def test_every_length_0_to_257_matches_libcrypto(alg, name):
    rng = np.random.default_rng(7)
    data = rng.integers(0, 256, 300, dtype=np.uint8)
    for n in range(258):
        m = data[:n].tobytes()
        ref = hashlib.new(name, m).digest()
        assert oracle.digest(alg, m) == ref
        assert oracle.evp(alg, m) == ref


@pytest.mark.parametrize("alg,name", ALGS)
def test_power_of_two_boundaries(alg, name):
    rng = np.random.default_rng(11)
    data = rng.integers(0, 256, (1 << 20) + 1, dtype=np.uint8)
    for k in range(3, 21):
        for n in ((1 << k) - 1, 1 << k, (1 << k) + 1):
            m = data[:n].tobytes()
            assert oracle.digest(alg, m) == hashlib.new(name, m).digest()


@pytest.mark.parametrize("alg,name", ALGS)
def test_streaming_update_in_64k_chunks(alg, name):
    # minio getHashSummer: one update() per 64 KiB stream chunk
    m = oracle.synth(3, 5 * 65536 + 123)
    ref = hashlib.new(name, m.tobytes()).digest()
    assert oracle.digest_chunked(alg, m, 65536) == ref
    assert oracle.digest_chunked(alg, m, 1000) == ref
    assert oracle.digest_chunked(alg, m, 1) == ref if m.size < 4096 else True


def test_batch_threaded_equals_single():
    n, ln = 37, 4099
    buf = oracle.synth_batch(0, n, ln)
    off = np.arange(n, dtype=np.uint64) * ln
    lens = np.full(n, ln, np.uint64)
    for alg, name in ALGS:
        a = oracle.batch(alg, buf, off, lens, threads=1)
        b = oracle.batch(alg, buf, off, lens, threads=4)
        assert (a == b).all()
        _, outs = oracle.evp_batch(alg, buf, off, lens, threads=3, chunk=65536)
        assert (outs[alg] == a).all()
        for i in (0, 17, 36):
            assert a[i].tobytes() == hashlib.new(name, buf[i * ln:(i + 1) * ln].tobytes()).digest()


def test_synth_generator_definition():
    # word j of message i = splitmix64(seed + (i<<40) + j), little-endian bytes
    def sm64(x):
        M = (1 << 64) - 1
        x = (x + 0x9E3779B97F4A7C15) & M
        x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & M
        x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & M
        return x ^ (x >> 31)

    seed = oracle.DEFAULT_SEED
    for i in (0, 1, 65535):
        got = oracle.synth(i, 45, seed)
        exp = b"".join(sm64((seed + (i << 40) + j) & ((1 << 64) - 1)).to_bytes(8, "little") for j in range(6))[:45]
        assert got.tobytes() == exp
    # byte_off addressing and the threaded fill agree with the byte-wise definition
    assert (oracle.synth(5, 20, seed, byte_off=13) == oracle.synth(5, 33, seed)[13:]).all()
    b = oracle.synth_batch(9, 3, 29, seed, threads=2)
    for k in range(3):
        assert (b[k * 29:(k + 1) * 29] == oracle.synth(9 + k, 29, seed)).all()


def test_framing(kat):
    fr = kat["framing"]
    assert oracle.base64_(hashlib.md5(b"abc").digest()) == fr["content_md5_abc"]
    md5s = b"".join(hashlib.md5(bytes([p["byte"]]) * p["count"]).digest() for p in fr["etag_parts"])
    assert oracle.etag_multipart(md5s) == fr["etag"]
    for size, part in fr["calc_part_size"]:
        assert oracle.calc_part_size(size) == part
    assert oracle.hex_(b"\x00\xff\x10") == "00ff10"
    import base64
    for n in range(0, 20):
        b = bytes(range(n))
        assert oracle.base64_(b) == base64.b64encode(b).decode()
