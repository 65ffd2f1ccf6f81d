"""Host logic around the hot path, on CPU: the process filter (the reference's own three test cases,
test/process/filter_dirs.js:17-82, on the same directory shapes), S3 framing, bencode, the reader's file
concatenation, and the in-process S3 mock."""
import asyncio
import hashlib
import os

import numpy as np
import pytest

import oracle
from downloader_b200 import s3 as s3f
from downloader_b200 import torrent as tor
from downloader_b200.reader import _ConcatFiles, piece_bounds
from downloader_b200.stages import Job, MediaType
from downloader_b200.stages import process as process_stage
from oracle.s3mock import S3Error, S3Mock

HERE = os.path.dirname(os.path.abspath(__file__))
FIX = os.path.join(HERE, "fixtures", "filter_dirs")


class MockLogger:
    def info(self, *a): pass
    def warn(self, *a): pass
    def error(self, *a): pass
    def debug(self, *a): pass


def run(coro):
    return asyncio.new_event_loop().run_until_complete(coro)


@pytest.fixture()
def processor():
    return run(process_stage.factory({}, {}, MockLogger()))


def test_should_filter_non_season_directories(processor):
    base = os.path.join(FIX, "should_filter_non_season_directories")
    res = run(processor(Job(lastStage={"path": base}, media={"id": "<uuid>", "type": MediaType.TV})))
    assert len(res["files"]) == 2
    assert res["files"][0] == os.path.join(base, "S1/KonoSuba S1E1.mkv")
    assert res["downloadPath"] == base


def test_should_read_all_dirs_when_processing_a_movie(processor):
    base = os.path.join(FIX, "should_read_all_dirs_when_processing_a_movie")
    res = run(processor(Job(lastStage={"path": base}, media={"id": "<uuid>", "type": MediaType.MOVIE})))
    assert len(res["files"]) == 1
    assert res["files"][0] == os.path.join(base, "Some Movie Dir/Your Name.mkv")


def test_should_read_with_top_level_dir(processor):
    base = os.path.join(FIX, "should_read_with_top_level_dir")
    res = run(processor(Job(lastStage={"path": base}, media={"id": "<uuid>", "type": MediaType.TV})))
    assert len(res["files"]) == 1
    assert res["files"][0] == os.path.join(base, "Some Movie Dir/Your Name.mkv")


def test_process_rejects_directories_without_media(processor, tmp_path):
    (tmp_path / "a").mkdir()
    (tmp_path / "b").mkdir()
    (tmp_path / "a" / "x.txt").write_text("no media")
    with pytest.raises(RuntimeError, match="Failed to find any suitable media files"):
        run(processor(Job(lastStage={"path": str(tmp_path)}, media={"id": "x", "type": MediaType.TV})))


def test_part_planning_matches_golden_and_oracle(kat):
    for size, part in kat["framing"]["calc_part_size"]:
        assert s3f.calculate_part_size(size) == part == oracle.calc_part_size(size)
    assert s3f.calculate_part_size(1 << 30, 16 << 20, override=True) == 16 << 20
    assert s3f.part_ranges(0, 64) == [(0, 0)]
    r = s3f.part_ranges((1 << 30) + 5, 16 << 20)
    assert len(r) == 65 and r[0] == (0, 16 << 20) and r[-1] == (1 << 30, (1 << 30) + 5)
    p = s3f.plan_object(40 << 20, 16 << 20, override=True)
    assert p.multipart and [e - s for s, e in p.ranges] == [16 << 20, 16 << 20, 8 << 20]
    assert not s3f.plan_object(16 << 20, 16 << 20, override=True).multipart  # size <= partSize -> single PUT
    with pytest.raises(ValueError):
        s3f.calculate_part_size(s3f.MAX_OBJECT_SIZE + 1)


def test_header_encodings_and_mode_selection(kat):
    assert s3f.content_md5(hashlib.md5(b"abc").digest()) == kat["framing"]["content_md5_abc"]
    assert s3f.sha256_hex(hashlib.sha256(b"abc").digest()) == hashlib.sha256(b"abc").hexdigest()
    assert s3f.select_digests(anonymous=False, use_ssl=False) == 4   # plain HTTP + credentials -> SHA-256
    assert s3f.select_digests(anonymous=False, use_ssl=True) == 2    # TLS -> Content-MD5
    assert s3f.select_digests(anonymous=True, use_ssl=False) == 2
    md5s = np.stack([np.frombuffer(hashlib.md5(bytes([i])).digest(), np.uint8) for i in range(3)])
    listed = {1: '"' + hashlib.md5(bytes([0])).hexdigest() + '"', 3: "deadbeef"}
    assert s3f.parts_to_skip(md5s, listed) == [1]


def test_bencode_roundtrip_and_spans():
    obj = {b"announce": b"http://t", b"info": {b"name": b"n", b"piece length": 16384, b"pieces": bytes(40), b"files": [
        {b"length": 5, b"path": [b"a", b"b.mkv"]}, {b"length": 0, b"path": [b"c"]}]}, b"n": -7}
    enc = tor.bencode(obj)
    spans = {}
    assert tor.bdecode(enc, spans) == obj
    s, e = spans[(b"info",)]
    assert enc[s:e] == tor.bencode(obj[b"info"])
    for bad in (b"i12", b"5:abc", b"d3:fooe", b"le "):
        with pytest.raises(ValueError):
            tor.bdecode(bad)


def test_concat_files_reads_across_boundaries(tmp_path):
    sizes = [10, 0, 4097, 33]
    blob = oracle.synth(1, sum(sizes)).tobytes()
    files, pos = [], 0
    for i, n in enumerate(sizes):
        p = tmp_path / f"f{i}"
        p.write_bytes(blob[pos:pos + n])
        files.append((str(p), n))
        pos += n
    files.append((str(tmp_path / "missing"), 20))  # a file that is not there reads as zeros
    src = _ConcatFiles(files)
    try:
        for start, ln in [(0, 10), (5, 100), (9, 4200), (0, sum(sizes) + 20), (4100, 60)]:
            buf = np.full(ln, 0xEE, np.uint8)
            src.readinto(buf, start)
            exp = (blob + bytes(20))[start:start + ln]
            assert buf.tobytes() == exp.ljust(ln, b"\0")
    finally:
        src.close()
    assert piece_bounds(10, 4) == [(0, 4), (4, 8), (8, 10)]


def test_concat_files_chunked_thread_pool_reads(tmp_path, monkeypatch):
    """Large ranges are cut into chunks read by a thread pool; same bytes as the serial read, short file included."""
    from downloader_b200 import reader
    monkeypatch.setattr(reader, "READ_CHUNK", 1000)
    monkeypatch.setattr(reader, "READ_THREADS", 4)
    sizes = [12345, 7, 50000]
    blob = oracle.synth(2, sum(sizes)).tobytes()
    files, pos = [], 0
    for i, n in enumerate(sizes):
        p = tmp_path / f"g{i}"
        p.write_bytes(blob[pos:pos + n] if i != 2 else blob[pos:pos + n - 3000])  # last file is 3000 bytes short on disk
        files.append((str(p), n))
        pos += n
    exp_all = blob[:sum(sizes) - 3000] + bytes(3000)
    src = _ConcatFiles(files)
    try:
        for start, ln in [(0, sum(sizes)), (999, 30001), (12340, 20), (60000, 2352 + 100)]:
            buf = np.full(ln, 0xEE, np.uint8)
            src.readinto(buf, start)
            assert buf.tobytes() == exp_all[start:start + ln].ljust(ln, b"\0"), (start, ln)
    finally:
        src.close()


def test_s3_mock_plays_the_server():
    m = S3Mock()
    assert not m.bucketExists("b")
    m.makeBucket("b")
    body = b"true"
    assert m.putObject("b", "k", body, {"Content-MD5": s3f.content_md5(hashlib.md5(body).digest())}) == hashlib.md5(body).hexdigest()
    with pytest.raises(S3Error, match="BadDigest"):
        m.putObject("b", "k", body, {"Content-MD5": s3f.content_md5(hashlib.md5(b"x").digest())})
    with pytest.raises(S3Error, match="XAmzContentSHA256Mismatch"):
        m.putObject("b", "k", body, {"x-amz-content-sha256": hashlib.sha256(b"x").hexdigest()})
    with pytest.raises(S3Error, match="InvalidArgument"):
        m.putObject("b", "k", body, {"x-amz-content-sha256": "abcd"})
    uid = m.newMultipartUpload("b", "o")
    parts = [b"a" * 100, b"b" * 50, b"c"]
    etags = [m.putPart("b", "o", uid, i + 1, p, {"x-amz-content-sha256": hashlib.sha256(p).hexdigest()}) for i, p in enumerate(parts)]
    et = m.completeMultipartUpload("b", "o", uid, etags)
    assert et == oracle.etag_multipart(b"".join(hashlib.md5(p).digest() for p in parts))


# ---- js/lib/process.js is tied to the mirror that runs ---------------------------------------------
ROOT = os.path.dirname(HERE)
RULES_JSON = os.path.join(ROOT, "js", "lib", "process_rules.json")
REF_PROCESS = "/root/reference/lib/process.js"


def test_process_rules_are_one_table_for_js_and_python():
    """Node is absent, so js/lib/process.js cannot run its fixtures.  Instead both it and the Python mirror (which
    does run them, above) take every literal from js/lib/process_rules.json: the JS file must hold no extension
    list or directory regex of its own, and must apply the rules in the reference's order."""
    import json
    import re
    with open(RULES_JSON) as f:
        rules = json.load(f)
    assert process_stage.RULES == rules
    assert list(process_stage.MEDIA_EXTS) == rules["mediaExts"] == [".mp4", ".mkv", ".mov", ".webm"]
    assert process_stage._RE_SKIP.pattern == "/extras|/commentary" and process_stage._RE_SKIP.flags & re.I
    assert process_stage._RE_SEASON.pattern == r"s\d+|season" and process_stage._RE_SEASON.flags & re.I
    with open(os.path.join(ROOT, "js", "lib", "process.js")) as f:
        js = f.read()
    code = re.sub(r"/\*.*?\*/", "", js, flags=re.S)
    code = "\n".join(line.split("//")[0] for line in code.splitlines())
    assert "require('./process_rules.json')" in code
    for literal in ("'.mkv'", "'.mp4'", "'.mov'", "'.webm'", "|season", "extras", "commentary", "suitable media files"):
        assert literal not in code, f"js/lib/process.js carries its own copy of {literal!r}"
    assert not re.search(r"[=(,]\s*/[^/\n]+/[a-z]*\.test\(", code), "regex literal in js/lib/process.js"
    # decision order of wantDirectory = lib/process.js:36-67: only top-level dir, movie mode, skip, season
    body = code[code.index("function wantDirectory"):code.index("function collect")]
    order = [body.index(k) for k in ("onlyTopLevelDirectoryIsAlwaysEntered", "movieModeEntersEveryDirectory", "SKIP_DIR.test(full)", "SEASON_DIR.test(base)")]
    assert order == sorted(order)
    assert "RULES.noMediaError" in code and "new RegExp(RULES.skipDirectory.source, RULES.skipDirectory.flags)" in code


@pytest.mark.skipif(not os.path.exists(REF_PROCESS), reason="reference tree not present (GPU box)")
def test_process_rules_equal_the_reference_literals():
    """The table's values are the reference's own literals (lib/process.js:15-20,59,64 and the error text, :111)."""
    import json
    import re
    with open(RULES_JSON) as f:
        rules = json.load(f)
    with open(REF_PROCESS) as f:
        ref = f.read()
    exts = re.findall(r"'(\.[a-z0-9]+)'", ref[ref.index("const mediaExts"):ref.index("]", ref.index("const mediaExts"))])
    assert exts == rules["mediaExts"]
    regexes = re.findall(r"if \(/(.+?)/([a-z]*)\.test\((\w+)\)\)", ref)
    assert regexes == [(rules["skipDirectory"]["source"], rules["skipDirectory"]["flags"], "item"),
                       (rules["seasonDirectory"]["source"], rules["seasonDirectory"]["flags"], "name")]
    assert f"throw new Error('{rules['noMediaError']}')" in ref


def test_top_level_rule_uses_the_basename_like_the_reference(tmp_path):
    """lib/process.js:40-45 tests `path.join(absPath, name)`: ANY directory whose basename equals the only top-level
    entry's name is entered, at any depth — kept as is, in both mirrors."""
    base = tmp_path / "only"
    (base / "only").mkdir(parents=True)          # nested directory with the top-level directory's name
    (base / "only" / "deep.mkv").write_bytes(b"")
    (base / "other").mkdir()
    (base / "other" / "no.mkv").write_bytes(b"")
    p = run(process_stage.factory({}, {}, MockLogger()))
    res = run(p(Job(lastStage={"path": str(tmp_path)}, media={"id": "x", "type": MediaType.TV})))
    assert res["files"] == [str(base / "only" / "deep.mkv")]


def _aligned_u8(n, align=4096):
    raw = np.zeros(n + align, np.uint8)
    off = (-raw.ctypes.data) % align
    return raw[off:off + n]


def test_o_direct_reads_give_the_same_bytes_as_buffered_reads(tmp_path):
    """reader._ConcatFiles(direct=True): the 4 KiB-aligned body of a chunk goes through an O_DIRECT descriptor, the
    unaligned head / tail and file systems without O_DIRECT through the page cache.  Whatever the mix, the bytes in
    the slot are the file's bytes, missing tails are zeros, and the two byte counters add up to what was read."""
    rng = np.random.default_rng(4096)
    sizes = [3 * 4096 * 700 + 1234, 0, 4096 * 5, 777]
    files, blob = [], b""
    for i, n in enumerate(sizes):
        p = tmp_path / f"f{i}.bin"
        data = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        p.write_bytes(data)
        blob += data
        files.append((str(p), n))
    total = len(blob)
    for direct in (False, True):
        src = _ConcatFiles(files, direct=direct)
        try:
            for pos, n in [(0, total), (4096, 4096 * 600), (4096 * 3, 4096 * 10 + 5), (17, 50000), (total - 100, 300)]:
                view = _aligned_u8(n)
                view[:] = 0xAA
                src.readinto(view, pos)
                want = blob[pos:pos + n] + bytes(max(0, pos + n - total))
                assert view.tobytes() == want, (direct, pos, n)
            assert src.direct_bytes + src.buffered_bytes == sum(min(n, total - pos) for pos, n in
                                                                 [(0, total), (4096, 4096 * 600), (4096 * 3, 4096 * 10 + 5), (17, 50000), (total - 100, 300)])
            if not direct:
                assert src.direct_bytes == 0
            elif src._dfds and any(fd >= 0 for fd in src._dfds.values()):
                assert src.direct_bytes > 0  # this file system takes O_DIRECT: the aligned bodies used it
        finally:
            src.close()


# ---- job runner: the Python mirror of lib/main.js's stage loop runs; js/lib/main.js is checked to have the same shape ----
class _Msg:
    def __init__(self, content=b"job"):
        self.content, self.acked, self.nacked = content, 0, 0

    def ack(self):
        self.acked += 1

    def nack(self):
        self.nacked += 1


class _Codec:
    def decode_download(self, content):
        return {"media": {"id": "job-1", "type": 0}}

    def encode_convert(self, payload):
        return repr(sorted(payload)).encode()


class _Queue:
    def __init__(self, fail=False):
        self.sent, self.fail = [], fail

    def publish(self, topic, body):
        if self.fail:
            raise ConnectionError("broker gone")
        self.sent.append((topic, body))


class _S3:
    def __init__(self, done=False):
        self.done, self.asked = done, []

    def getObject(self, bucket, key):
        self.asked.append((bucket, key))
        if not self.done:
            raise KeyError("NoSuchKey")
        return b"true"


def _runner(s3, queue, behaviour):
    from downloader_b200.stages import _NullTelem
    from downloader_b200.stages import main as main_stage
    calls = []

    def make(name):
        async def fac(config, emitter, logger):
            async def stage(job):
                calls.append((name, dict(job.get("lastStage") or {})))
                if behaviour.get(name) is not None:
                    raise behaviour[name]
                return {"from": name}
            return stage
        return fac
    tel = _NullTelem()
    cfg = {"telem": tel, "s3": s3, "queue": queue, "codec": _Codec()}
    handle = run(main_stage.factory(cfg, MockLogger(), factories={n: make(n) for n in main_stage.STAGES}))
    return handle, calls, tel


def test_job_runner_runs_stages_in_order_and_publishes():
    s3, q = _S3(done=False), _Queue()
    handle, calls, tel = _runner(s3, q, {})
    m = _Msg()
    run(handle(m))
    assert [c[0] for c in calls] == ["download", "process", "upload"]
    assert calls[0][1] == {} and calls[1][1] == {"from": "download"} and calls[2][1] == {"from": "process"}  # lastStage threading
    assert s3.asked == [("triton-staging", "job-1/original/done")]
    assert [t for t, _ in q.sent] == ["v1.convert"] and (m.acked, m.nacked) == (1, 0)
    assert tel.events[0] == ("job-1", "status", 2)


def test_job_runner_skips_a_job_that_is_already_staged():
    """lib/main.js:117-124,153-155: the `done` marker exists -> no stage runs, the convert job is published again, ack."""
    s3, q = _S3(done=True), _Queue()
    handle, calls, _ = _runner(s3, q, {})
    m = _Msg()
    run(handle(m))
    assert calls == [] and len(q.sent) == 1 and (m.acked, m.nacked) == (1, 0)


def test_job_runner_error_policy_matches_the_reference():
    """lib/main.js:141-151,157-166: a stalled download is acked and dropped; any other stage error -> status 6 + nack,
    nothing published; a failed publish is only logged — neither ack nor nack, no ERRORED status."""
    from downloader_b200.stages.download import DownloadStalled
    handle, calls, tel = _runner(_S3(), _Queue(), {"download": DownloadStalled("Download stalled.")})
    m = _Msg()
    run(handle(m))
    assert (m.acked, m.nacked) == (1, 0) and [c[0] for c in calls] == ["download"] and ("job-1", "status", 6) not in tel.events
    q = _Queue()
    handle, calls, tel = _runner(_S3(), q, {"process": RuntimeError("Failed to find any suitable media files")})
    m = _Msg()
    run(handle(m))
    assert (m.acked, m.nacked) == (0, 1) and q.sent == [] and tel.events[-1] == ("job-1", "status", 6)
    handle, calls, tel = _runner(_S3(), _Queue(fail=True), {})
    m = _Msg()
    run(handle(m))
    assert (m.acked, m.nacked) == (0, 0) and len(calls) == 3 and ("job-1", "status", 6) not in tel.events


def test_js_main_has_the_shape_of_the_mirror_that_runs():
    """js/lib/main.js cannot be executed here; check the order of its steps against the Python mirror's: done-marker
    lookup before the stage loop, ERRDLSTALL -> ack and otherwise ERRORED + nack inside the stage try/catch, and the
    publish in a try/catch of its own that neither acks nor nacks."""
    import re
    with open(os.path.join(ROOT, "js", "lib", "main.js")) as f:
        js = f.read()
    code = re.sub(r"/\*.*?\*/", "", js, flags=re.S)
    code = "\n".join(line.split("//")[0] for line in code.splitlines())
    assert "['download', 'process', 'upload']" in code
    order = [code.index(k) for k in ("getObject('triton-staging', path.join(jobId, 'original/', 'done'))",
                                     "for (const name of STAGES) lastStage = await table[name]",
                                     "err.code === 'ERRDLSTALL'", "emitStatus(jobId, STATUS_ERRORED)",
                                     "config.queue.publish('v1.convert'", "failed to create job:")]
    assert order == sorted(order)
    publish_catch = code[code.index("config.queue.publish('v1.convert'"):code.index("message.ack()\n    active--")]
    assert "nack" not in publish_catch and "STATUS_ERRORED" not in publish_catch


CONTRACT = {  # strings of the stage contract: (reference file, JS module, Python mirror)
    "Download stalled.": ("download.js", "download.js", "download.py"),
    "ERRDLSTALL": ("download.js", "download.js", "download.py"),
    "File URLs are not allowed.": ("download.js", "download.js", "download.py"),
    "Protocol not supported.": ("download.js", "download.js", "download.py"),
    "Invalid files data type, expected array, got '": ("upload.js", "upload.js", "upload.py"),
    "triton-staging": ("upload.js", "upload.js", "upload.py"),
    "'original/'": ("upload.js", "upload.js", None),
    " not found.": ("upload.js", "upload.js", "upload.py"),
}


def test_stage_contract_strings_are_the_same_in_reference_js_and_python():
    """Error texts, the error code, the bucket and the object-name pieces are contract: lib/main.js and operators
    match on them.  Each must appear in the JS module, in the Python mirror that runs, and (when the reference tree
    is present) in the reference file it was taken from."""
    for text, (ref, js, py) in CONTRACT.items():
        with open(os.path.join(ROOT, "js", "lib", js)) as f:
            assert text in f.read(), (text, js)
        if py:
            with open(os.path.join(ROOT, "downloader_b200", "stages", py)) as f:
                assert text.strip("'") in f.read(), (text, py)
        refp = os.path.join("/root/reference/lib", ref)
        if os.path.exists(refp):
            with open(refp) as f:
                assert text in f.read(), (text, ref)
