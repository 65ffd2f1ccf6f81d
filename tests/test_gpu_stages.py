"""GPU tests of the callers either side of the hot path: upload stage -> S3 mock, torrent
metainfo -> expected table -> bulk verify -> bitfield, through the stage factories that mirror the reference's
plug-in contract.  Everything hashed on the GPU is checked against hashlib / the oracle."""
import asyncio
import hashlib
import os

import numpy as np
import pytest

import oracle
from oracle.s3mock import S3Mock

pytestmark = pytest.mark.gpu
MiB = 1 << 20


def run(coro):
    return asyncio.new_event_loop().run_until_complete(coro)


class Log:
    def __init__(self):
        self.lines = []

    def _l(self, *a):
        self.lines.append(" ".join(str(x) for x in a))
    info = warn = error = debug = _l


@pytest.fixture(scope="module")
def eng():
    import torch
    assert torch.cuda.is_available()
    from downloader_b200 import Engine
    e = Engine([0], ring_bytes=160 * MiB)
    yield e
    e.close()


def make_file(path, msg_index, size):
    os.makedirs(os.path.dirname(path), exist_ok=True)
    data = oracle.synth(msg_index, size)
    with open(path, "wb") as f:
        f.write(data.tobytes())
    return data


@pytest.mark.parametrize("use_ssl", [True, False])
def test_upload_stage_into_s3_mock(eng, tmp_path, use_ssl):
    from downloader_b200.stages import Job, _NullTelem
    from downloader_b200.stages import upload
    dl = tmp_path / "dl" / "job1"
    sizes = {"Show S1E1.mkv": 40 * MiB + 12345, "small.mp4": 3 * MiB, "empty.webm": 0}
    datas = {n: make_file(str(dl / n), i, s) for i, (n, s) in enumerate(sizes.items())}
    mock, tel, log = S3Mock(), _NullTelem(), Log()
    cfg = {"s3": mock, "engine": eng, "telem": tel, "part_size": 16 * MiB, "override_part_size": True,
           "anonymous": False, "use_ssl": use_ssl}
    stage = run(upload.factory(cfg, {}, log))
    files = [str(dl / n) for n in sizes]
    assert run(stage(Job(media={"id": "job1"}, lastStage={"files": files, "downloadPath": str(dl)}))) is None
    import base64
    for n, s in sizes.items():
        key = os.path.join("job1", "original/", base64.b64encode(n.encode()).decode())
        obj = mock.statObject("triton-staging", key)
        assert obj["size"] == s
        raw = datas[n].tobytes()
        if s > 16 * MiB:
            parts = [raw[o:o + 16 * MiB] for o in range(0, s, 16 * MiB)]
            assert obj["etag"] == oracle.etag_multipart(b"".join(hashlib.md5(p).digest() for p in parts))
        else:
            assert obj["etag"] == hashlib.md5(raw).hexdigest()
    assert mock.statObject("triton-staging", "job1/original/done")["etag"] == hashlib.md5(b"true").hexdigest()
    # the connection mode decides which digest travels (minio enableSHA256 = !anonymous && !useSSL)
    for req in mock.requests:
        h = req[-1]
        if use_ssl:
            assert "Content-MD5" in h and h.get("x-amz-content-sha256", "UNSIGNED-PAYLOAD") == "UNSIGNED-PAYLOAD"
        else:
            assert "Content-MD5" not in h and len(h["x-amz-content-sha256"]) == 64
    assert [p for _, _, p in tel.events] == [66, 83, 100]
    assert not dl.exists()  # lib/upload.js:60-64 removes the download directory


def test_upload_resume_skips_parts_whose_md5_is_already_listed(eng, tmp_path):
    from downloader_b200.stages.upload import f_put_object
    p = tmp_path / "f.mkv"
    data = make_file(str(p), 9, 20 * MiB).tobytes()
    mock = S3Mock()
    mock.makeBucket("b")
    listed = {2: hashlib.md5(data[8 * MiB:16 * MiB]).hexdigest(), 3: "0" * 32}

    class Resuming(S3Mock):
        pass
    # pre-seed the mock's upload with part 2 so completion succeeds without re-sending it
    orig_new = mock.newMultipartUpload

    def new_upload(bucket, name):
        uid = orig_new(bucket, name)
        md5 = hashlib.md5(data[8 * MiB:16 * MiB]).digest()
        mock.uploads[uid]["parts"][2] = {"etag": md5.hex(), "md5": md5, "size": 8 * MiB}
        return uid
    mock.newMultipartUpload = new_upload
    et = run(f_put_object(mock, eng, "b", "o", str(p), 8 * MiB, True, 2, listed_parts=listed))
    sent = sorted(r[3] for r in mock.requests if r[0] == "PUT-PART")
    assert sent == [1, 3]
    assert et == oracle.etag_multipart(b"".join(hashlib.md5(data[o:o + 8 * MiB]).digest() for o in range(0, 20 * MiB, 8 * MiB)))


def test_upload_rejects_bad_job_like_the_reference(eng, tmp_path):
    from downloader_b200.stages import Job, upload
    stage = run(upload.factory({"s3": S3Mock(), "engine": eng}, {}, Log()))
    with pytest.raises(TypeError, match="Invalid files data type, expected array"):
        run(stage(Job(media={"id": "x"}, lastStage={"files": "nope", "downloadPath": str(tmp_path)})))
    with pytest.raises(FileNotFoundError, match="not found"):
        run(stage(Job(media={"id": "x"}, lastStage={"files": [str(tmp_path / "gone.mkv")], "downloadPath": str(tmp_path)})))


def _make_torrent(eng, root, name, sizes, piece_length):
    from downloader_b200 import torrent as tor
    rels, blob = [], b""
    for i, (rel, n) in enumerate(sizes):
        blob += make_file(os.path.join(root, name, rel), 100 + i, n).tobytes()
        rels.append(rel)
    return tor.create_metainfo(eng, root, rels, piece_length, name), blob


def test_metainfo_expected_table_and_info_hash(eng, tmp_path):
    from downloader_b200 import torrent as tor
    sizes = [("S1/e1.mkv", 300000), ("S1/e2.mkv", 0), ("notes.txt", 77), ("S2/e1.mkv", 1 << 20)]
    data, blob = _make_torrent(eng, str(tmp_path), "show", sizes, 65536)
    meta = tor.parse_metainfo(data, eng)
    top = tor.bdecode(data)
    assert meta.info_hash == hashlib.sha1(tor.bencode(top[b"info"])).hexdigest()
    assert meta.length == len(blob) and meta.piece_length == 65536
    assert meta.num_pieces == -(-len(blob) // 65536) and meta.last_piece_length == len(blob) % 65536
    for i in range(meta.num_pieces):  # webtorrent's _hashes: lower-case hex of each 20-byte digest
        assert meta.piece_hex(i) == hashlib.sha1(blob[i * 65536:(i + 1) * 65536]).hexdigest()
    assert [p for p, _ in meta.files] == [os.path.join("show", r) for r, _ in sizes]
    with pytest.raises(ValueError):
        tor.parse_metainfo(tor.bencode({"info": {"name": "x", "piece length": 4, "pieces": b"123"}}), eng)


def test_bulk_verify_bitfield_tracks_corruption_and_missing_files(eng, tmp_path):
    from downloader_b200 import torrent as tor
    from downloader_b200.reader import verify_pieces
    sizes = [("a.mkv", 5 * 262144 + 1000), ("b.mkv", 262144 * 3 - 1000), ("c.mkv", 70000)]
    data, blob = _make_torrent(eng, str(tmp_path), "t", sizes, 262144)
    meta = tor.parse_metainfo(data, eng)
    files = meta.on_disk(str(tmp_path))
    res = verify_pieces(eng, files, meta.piece_length, meta.pieces, batch_bytes=1 << 20)  # several ring slots
    assert res["ok"].all() and res["ok"].size == meta.num_pieces == 9
    # flip one byte inside piece 5 (which spans a.mkv -> b.mkv) and truncate c.mkv
    with open(files[1][0], "r+b") as f:
        f.seek(10)
        b = f.read(1)
        f.seek(10)
        f.write(bytes([b[0] ^ 1]))
    os.remove(files[2][0])
    res = verify_pieces(eng, files, meta.piece_length, meta.pieces)
    assert np.nonzero(~res["ok"])[0].tolist() == [5, 8]


def test_download_stage_verifies_refetches_and_reports(eng, tmp_path, monkeypatch):
    from downloader_b200 import torrent as tor
    from downloader_b200.stages import Job, SourceType, _NullTelem, download
    seed_root = tmp_path / "seed"
    data, blob = _make_torrent(eng, str(seed_root), "movie", [("movie.mkv", 9 * 131072 + 5)], 131072)
    tfile = tmp_path / "movie.torrent"
    tfile.write_bytes(data)
    base = tmp_path / "downloads"
    calls = []

    async def fetcher(meta, missing, download_path):  # stands in for the peer wire: delivers the missing pieces
        calls.append(list(missing))
        dst = os.path.join(download_path, meta.files[0][0])
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        mode = "r+b" if os.path.exists(dst) else "wb"
        with open(dst, mode) as f:
            for i in missing:
                f.seek(i * meta.piece_length)
                f.write(blob[i * meta.piece_length:(i + 1) * meta.piece_length])

    tel = _NullTelem()
    cfg = {"engine": eng, "fetcher": fetcher, "telem": tel, "instance": {"download_path": str(base)}}
    stage = run(download.factory(cfg, {}, Log()))
    job = Job(media={"id": "m1", "source": SourceType.HTTP, "sourceURI": "http://example.invalid" + str(tfile)})
    out = run(stage(job))
    assert out == {"path": str(base / "m1")}
    assert calls == [list(range(10))]                      # nothing on disk -> every piece requested once
    assert tel.events[0][2] == 0 and tel.events[-1][2] == 50
    assert (base / "m1" / "movie" / "movie.mkv").read_bytes() == blob
    # restart with the data on disk: bulk re-verify passes without fetching (webtorrent _verifyPieces)
    calls.clear()
    assert run(stage(job)) == {"path": str(base / "m1")} and calls == []
    # no fetcher and a damaged piece: the stall error the orchestrator acks and drops
    with open(base / "m1" / "movie" / "movie.mkv", "r+b") as f:
        f.seek(131072 * 3 + 7)
        f.write(b"\xff")
    stalled = run(download.factory({"engine": eng, "instance": {"download_path": str(base)}}, {}, Log()))
    with pytest.raises(download.DownloadStalled) as ei:
        run(stalled(job))
    assert ei.value.code == "ERRDLSTALL"
    bad = Job(media={"id": "m1", "source": SourceType.BUCKET, "sourceURI": "bucket://x"})
    with pytest.raises(RuntimeError, match="Protocol not supported."):
        run(stage(bad))
    monkeypatch.delenv("ALLOW_FILE_URLS", raising=False)
    with pytest.raises(PermissionError, match="File URLs are not allowed."):
        run(stage(Job(media={"id": "m1", "source": SourceType.FILE, "sourceURI": "file:///etc/hostname"})))


def test_refetch_rounds_hash_only_the_pieces_that_were_missing(eng, tmp_path):
    """a refetch round must move about len(missing) x pieceLength over PCIe, not the whole torrent
    (webtorrent verifies a piece when it completes — the onChunk shape, call site lib/download.js:64-66)."""
    from downloader_b200 import torrent as tor
    from downloader_b200.reader import verify_pieces
    from downloader_b200.stages import Job, SourceType, _NullTelem, download
    piece = 262144
    n = 160                                                         # 40 MiB torrent
    seed_root = tmp_path / "seed"
    data, blob = _make_torrent(eng, str(seed_root), "show", [("a.mkv", 100 * piece), ("b.mkv", 60 * piece - 777)], piece)
    meta = tor.parse_metainfo(data, eng)
    dl = tmp_path / "dl" / "j"
    os.makedirs(dl / "show")
    damaged = [3, 4, 99, 100, 159]                                  # incl. the piece that spans both files and the short last one
    bad = bytearray(blob)
    for i in damaged:
        bad[i * piece + 11] ^= 0x40
    (dl / "show" / "a.mkv").write_bytes(bytes(bad[:100 * piece]))
    (dl / "show" / "b.mkv").write_bytes(bytes(bad[100 * piece:]))
    files = meta.on_disk(str(dl))
    full = verify_pieces(eng, files, piece, meta.pieces)
    assert np.nonzero(~full["ok"])[0].tolist() == damaged
    # subset verify: only those pieces cross PCIe, rows come back in the order asked for
    s0 = eng.stats()
    sub = verify_pieces(eng, files, piece, meta.pieces, indices=damaged)
    s1 = eng.stats()
    moved = s1["h2d_bytes"] - s0["h2d_bytes"]
    assert not sub["ok"].any() and sub["indices"].tolist() == damaged
    assert (sub[1] == full[1][damaged]).all()                       # same digests as the bulk pass gave these pieces
    assert moved < len(damaged) * piece + (64 << 10), moved          # payload + a few KiB of descriptors
    assert s1["bytes"] - s0["bytes"] == sum(min(piece, len(blob) - i * piece) for i in damaged)
    with pytest.raises(ValueError):
        verify_pieces(eng, files, piece, meta.pieces, indices=[5, 5])
    assert verify_pieces(eng, files, piece, meta.pieces, indices=[])["ok"].size == 0
    # the stage: round 0 reads everything once, the round after the refetch reads only what was refetched
    calls = []

    async def fetcher(m, missing, download_path):
        calls.append(list(missing))
        for i in missing:
            lo, hi = i * piece, min((i + 1) * piece, len(blob))
            for name, start, end in (("a.mkv", 0, 100 * piece), ("b.mkv", 100 * piece, len(blob))):
                a, b = max(lo, start), min(hi, end)
                if b > a:
                    with open(os.path.join(download_path, "show", name), "r+b") as f:
                        f.seek(a - start)
                        f.write(blob[a:b])

    tfile = tmp_path / "show.torrent"
    tfile.write_bytes(data)
    cfg = {"engine": eng, "fetcher": fetcher, "telem": _NullTelem(), "instance": {"download_path": str(tmp_path / "dl")}}
    stage = run(download.factory(cfg, {}, Log()))
    s0 = eng.stats()
    out = run(stage(Job(media={"id": "j", "source": SourceType.TORRENT, "sourceURI": str(tfile)})))
    s1 = eng.stats()
    assert out == {"path": str(dl)} and calls == [damaged]
    hashed = s1["bytes"] - s0["bytes"]  # round 0: the torrent once; round 1: the refetched pieces; + the bencoded info dict
    again = sum(min(piece, len(blob) - i * piece) for i in damaged)
    assert len(blob) + again <= hashed < len(blob) + again + (64 << 10), (hashed, len(blob), again)
    assert (dl / "show" / "a.mkv").read_bytes() + (dl / "show" / "b.mkv").read_bytes() == blob


def test_stage_chain_download_process_upload(eng, tmp_path):
    """lib/main.js:129-140: stages run in order, each receiving the previous result as job.lastStage."""
    from downloader_b200.stages import Job, MediaType, SourceType, _NullTelem, download, process, upload
    seed = tmp_path / "seed"
    data, blob = _make_torrent(eng, str(seed), "Show", [("Season 1/e1.mkv", 700000), ("Extras/x.mkv", 50000), ("Season 1/e1.nfo", 10)], 32768)
    (tmp_path / "s.torrent").write_bytes(data)

    async def fetcher(meta, missing, download_path):
        import shutil
        shutil.copytree(seed / "Show", os.path.join(download_path, "Show"), dirs_exist_ok=True)
    mock, tel, log = S3Mock(), _NullTelem(), Log()
    cfg = {"engine": eng, "fetcher": fetcher, "telem": tel, "s3": mock, "instance": {"download_path": str(tmp_path / "dl")}}
    job = Job(media={"id": "j9", "type": MediaType.TV, "source": SourceType.TORRENT, "sourceURI": str(tmp_path / "s.torrent")})
    last = None
    for mod in (download, process, upload):
        fn = run(mod.factory(cfg, {}, log))
        job["lastStage"] = last
        last = run(fn(job))
    import base64
    keys = sorted(mock.buckets["triton-staging"])
    assert keys == sorted(["j9/original/" + base64.b64encode(b"e1.mkv").decode(), "j9/original/done"])
    e1 = (seed / "Show" / "Season 1" / "e1.mkv").read_bytes()
    assert mock.statObject("triton-staging", keys[0] if keys[0].endswith("done") is False else keys[1])["etag"] == hashlib.md5(e1).hexdigest()


def test_config4_shape_scaled_mixed_pieces_and_parts_from_pinned_ring(eng, tmp_path):
    """BASELINE config 4 scaled down: media files hashed twice from the pinned ring — as torrent pieces
    (1/2/4 MiB, SHA-1 + bitfield) and as S3 parts (8/16/32/64 MiB, MD5 + SHA-256) — with short last piece /
    part; every digest against libcrypto."""
    from downloader_b200 import s3 as s3f
    from downloader_b200.reader import hash_file_parts, piece_bounds, verify_pieces
    rng = np.random.default_rng(40)
    sizes = [int(rng.integers(70, 140)) * MiB + int(rng.integers(1, MiB)) for _ in range(3)]
    for k, size in enumerate(sizes):
        path = str(tmp_path / f"media{k}.mkv")
        data = make_file(path, 500 + k, size)
        plen = [1, 2, 4][k % 3] * MiB
        psize = [8, 16, 32, 64][(k + 1) % 4] * MiB
        pb = piece_bounds(size, plen)
        off = np.array([s for s, _ in pb], np.uint64)
        ln = np.array([e - s for s, e in pb], np.uint64)
        _, exp = oracle.evp_batch(oracle.SHA1, data, off, ln, threads=16)
        bad = [0, len(pb) - 1]
        table = exp[oracle.SHA1].copy()
        for b in bad:
            table[b, 0] ^= 0x55
        res = verify_pieces(eng, [(path, size)], plen, table, batch_bytes=48 * MiB, in_flight=3)
        assert (res[1] == exp[oracle.SHA1]).all()
        assert np.nonzero(~res["ok"])[0].tolist() == bad
        ranges = s3f.part_ranges(size, psize)
        d = hash_file_parts(eng, path, ranges, batch_bytes=150 * MiB)
        off = np.array([s for s, _ in ranges], np.uint64)
        ln = np.array([e - s for s, e in ranges], np.uint64)
        _, exp = oracle.evp_batch(oracle.MD5 | oracle.SHA256, data, off, ln, threads=16, chunk=65536)
        assert (d[2] == exp[oracle.MD5]).all() and (d[4] == exp[oracle.SHA256]).all()
        assert s3f.multipart_etag(eng, d[2])[0] == oracle.etag_multipart(exp[oracle.MD5])


def test_one_pass_feeds_pieces_and_parts_from_the_same_slot(eng, tmp_path):
    """the same bytes hashed as torrent pieces and as S3 parts with one H2D copy per slot."""
    from downloader_b200 import s3 as s3f
    from downloader_b200.reader import hash_media_file, piece_bounds
    size = 150 * MiB + 4321
    path = str(tmp_path / "m.mkv")
    data = make_file(path, 900, size)
    plen, psize = 2 * MiB, 16 * MiB
    pb = piece_bounds(size, plen)
    _, e1 = oracle.evp_batch(oracle.SHA1, data, np.array([s for s, _ in pb], np.uint64), np.array([e - s for s, e in pb], np.uint64), threads=16)
    table = e1[oracle.SHA1].copy()
    table[7, 1] ^= 1
    ranges = s3f.part_ranges(size, psize)
    s0 = eng.stats()
    r = hash_media_file(eng, path, plen, table, ranges, batch_bytes=64 * MiB, in_flight=2)
    s1 = eng.stats()
    assert (r["pieces"][1] == e1[oracle.SHA1]).all()
    assert np.nonzero(~r["pieces"]["ok"])[0].tolist() == [7]
    _, e2 = oracle.evp_batch(oracle.MD5 | oracle.SHA256, data, np.array([s for s, _ in ranges], np.uint64),
                             np.array([e - s for s, e in ranges], np.uint64), threads=16, chunk=65536)
    assert (r["parts"][2] == e2[oracle.MD5]).all() and (r["parts"][4] == e2[oracle.SHA256]).all()
    payload = s1["h2d_bytes"] - s0["h2d_bytes"]
    assert size <= payload < size + 4 * MiB, "the slot must cross PCIe once, not once per submit"
    assert s1["submits"] - s0["submits"] == 2 * 3 and s1["bytes"] - s0["bytes"] == 2 * size


def test_o_direct_reads_and_files_dealt_to_gpus(eng, tmp_path):
    """reader.hash_file_parts(direct=True): the 4 KiB-aligned bodies of the chunks come through O_DIRECT straight into
    the pinned slots (where the file system takes it), the digests are the file's; reader.hash_media_files deals whole
    files to the context's GPUs (here: the one there is) with a feeder thread and reader pool per GPU."""
    from downloader_b200 import ALG_MD5, ALG_SHA1, ALG_SHA256, reader, s3
    sizes = [9 * MiB + 4321, 5 * MiB, 3 * MiB + 1]
    paths, datas = [], []
    for i, n in enumerate(sizes):
        p = str(tmp_path / f"m{i}.bin")
        datas.append(make_file(p, 300 + i, n).tobytes())
        paths.append(p)
    ranges = s3.part_ranges(sizes[0], 4 * MiB)
    for direct in (False, True):
        r = reader.hash_file_parts(eng, paths[0], ranges, ALG_MD5 | ALG_SHA256, batch_bytes=8 * MiB, direct=direct)
        for k, (a, b) in enumerate(ranges):
            assert bytes(r[ALG_MD5][k]) == hashlib.md5(datas[0][a:b]).digest(), (direct, k)
            assert bytes(r[ALG_SHA256][k]) == hashlib.sha256(datas[0][a:b]).digest(), (direct, k)
    jobs = [dict(path=p, piece_length=MiB, expected_sha1=None, part_ranges=s3.part_ranges(n, 4 * MiB), batch_bytes=8 * MiB, direct=True)
            for p, n in zip(paths, sizes)]
    out = reader.hash_media_files(eng, jobs, gpus=[0])
    assert [o["gpu"] for o in out] == [0, 0, 0]
    for o, d, n in zip(out, datas, sizes):
        for k in range(-(-n // MiB)):
            assert bytes(o["pieces"][ALG_SHA1][k]) == hashlib.sha1(d[k * MiB:(k + 1) * MiB]).digest()
        for k, (a, b) in enumerate(s3.part_ranges(n, 4 * MiB)):
            assert bytes(o["parts"][ALG_MD5][k]) == hashlib.md5(d[a:b]).digest()
        assert o["read"]["direct_bytes"] + o["read"]["buffered_bytes"] == n


def test_one_context_drives_two_gpus(tmp_path):
    """th_init with a GPU list: per-GPU rings/streams in one process, pieces sharded by index."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run under gpurun --gpus 2)")
    from downloader_b200 import Engine
    from downloader_b200.sharding import shard_range
    n, ln = 64, 65536
    data = oracle.synth_batch(0, n, ln)
    exp = oracle.batch(oracle.SHA1, data, np.arange(n, dtype=np.uint64) * ln, np.full(n, ln, np.uint64), threads=8)
    with Engine([0, 1], ring_bytes=8 * MiB) as e2:
        tickets = []
        for g in (0, 1):
            lo, hi = shard_range(n, 2, g)
            slot, view = e2.ring_acquire((hi - lo) * ln, gpu=g)
            view[:] = data[lo * ln:hi * ln]
            tickets.append((lo, hi, e2.submit(slot, np.arange(hi - lo, dtype=np.uint64) * ln, np.full(hi - lo, ln, np.uint64), 1, exp[lo:hi])))
        for lo, hi, t in tickets:
            r = e2.result(t, hi - lo, 1, want_ok=True)
            assert (r[1] == exp[lo:hi]).all() and r["ok"].all()
        assert e2.stats(0)["submits"] == 1 and e2.stats(1)["submits"] == 1 and e2.stats()["submits"] == 2
