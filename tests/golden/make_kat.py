"""Generates tests/golden/kat.json — known-answer vectors for the byte-integrity path.

The reference (tritonmedia/downloader) pins no digest anywhere, so the golden
set is: the published RFC 1321 A.5 / FIPS 180-4 example vectors (hard-coded below, as published)
plus padding-boundary and framing cases whose expected values are produced by this image's OpenSSL
through Python hashlib — the same libcrypto Node's crypto binds.  The hard-coded published values
are asserted against hashlib before anything is written.

Run:  python tests/golden/make_kat.py      (rewrites kat.json deterministically)
"""
import base64
import hashlib
import json
import os

PUBLISHED = [  # (message spec, md5, sha1, sha256) exactly as published / listed in SURVEY.md §8c
    ("", "d41d8cd98f00b204e9800998ecf8427e", "da39a3ee5e6b4b0d3255bfef95601890afd80709",
     "e3b0c44298fc1c149afbf4c8996fb92427ae41e4649b934ca495991b7852b855"),
    ("a", "0cc175b9c0f1b6a831c399e269772661", "86f7e437faa5a7fce15d1ddcb9eaeaea377667b8",
     "ca978112ca1bbdcafac231b39a23dc4da786eff8147c4e72b9807785afee48bb"),
    ("abc", "900150983cd24fb0d6963f7d28e17f72", "a9993e364706816aba3e25717850c26c9cd0d89d",
     "ba7816bf8f01cfea414140de5dae2223b00361a396177a9cb410ff61f20015ad"),
    ("message digest", "f96b697d7cb7938d525a2f31aaf161d0", "c12252ceda8be8994d5fa0290a47231c1d16aae3",
     "f7846f55cf23e14eebeab5b4e1550cad5b509e3348fbc4efa3a1413d393cb650"),
    ("abcdefghijklmnopqrstuvwxyz", "c3fcd3d76192e4007dfb496cca67e13b",
     "32d10c7b8cf96570ca04ce37f2a19d84240d3a89",
     "71c480df93d6ae2f1efad1447c66c9525e316218cf51fc8d9ed832f2daf18b73"),
    ("ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789", "d174ab98d277d9f5a5611c2c9f419d9f",
     "761c457bf73b14d27e9e9265c46f4b4dda11f940",
     "db4bfcbd4da0cd85a60c3c37d3fbd8805c77f15fc6b1fdfe614ee0a7c8fdb4c0"),
    ("1234567890" * 8, "57edf4a22be3c955ac49da2e2107b67a", "50abf5706a150990a08b2c5ea40fa0e585554732",
     "f371bc4a311f2b009eef952dd83ca80e2b60026c8e935592d0f9c308453c813e"),
    ("abcdbcdecdefdefgefghfghighijhijkijkljklmklmnlmnomnopnopq", "8215ef0796a20bcaaae116d3876c664a",
     "84983e441c3bd26ebaae4aa1f95129e5e54670f1",
     "248d6a61d20638b8e5c026930c3e6039a33ce45964ff2167f6ecedd419db06c1"),
    ("abcdefghbcdefghicdefghijdefghijkefghijklfghijklmghijklmnhijklmnoijklmnopjklmnopqklmnopqrlmnopqrsmnopqrstnopqrstu",
     "03dd8807a93175fb062dfb55dc7d359c", "a49b2446a02c645bf419f995b67091253a04a259",
     "cf5b16a778af8380036ce59e7b0492370b249b11e8f07a51afac45037afee9d1"),
]
MILLION_A = ("7707d6ae4e027c70eea2a935c2296f21", "34aa973cd4c4daa4f61eeb2bdbad27316534016f",
             "cdc76e5c9914fb9281a1c7e284d73e67f1809a48a497200e046d39ccc7112cd0")


def dig(b):
    return {"md5": hashlib.md5(b).hexdigest(), "sha1": hashlib.sha1(b).hexdigest(),
            "sha256": hashlib.sha256(b).hexdigest()}


def main():
    vectors = []
    for msg, m5, s1, s256 in PUBLISHED:
        b = msg.encode()
        d = dig(b)
        assert (d["md5"], d["sha1"], d["sha256"]) == (m5, s1, s256), msg
        vectors.append({"kind": "ascii", "msg": msg, **d})
    d = dig(b"a" * 1_000_000)
    assert (d["md5"], d["sha1"], d["sha256"]) == MILLION_A
    vectors.append({"kind": "repeat", "byte": 0x61, "count": 1_000_000, **d})
    # padding boundaries: 55/56/63/64/65/119/120 zero bytes, + a few more
    for n in (55, 56, 63, 64, 65, 119, 120, 127, 128, 129):
        vectors.append({"kind": "repeat", "byte": 0, "count": n, **dig(bytes(n))})
    # lib/upload.js:55 writes the 4-byte body 'true' (small-object digest)
    vectors.append({"kind": "ascii", "msg": "true", **dig(b"true")})
    framing = {
        "content_md5_abc": base64.b64encode(hashlib.md5(b"abc").digest()).decode(),
        "etag_parts": [{"byte": 0x61, "count": 5 << 20}, {"byte": 0x62, "count": 5 << 20}, {"byte": 0x63, "count": 1}],
    }
    assert framing["content_md5_abc"] == "kAFQmDzST7DWlj99KOF/cg=="
    cat = b"".join(hashlib.md5(bytes([p["byte"]]) * p["count"]).digest() for p in framing["etag_parts"])
    framing["etag"] = hashlib.md5(cat).hexdigest() + "-3"
    assert framing["etag"] == "0a97f1336a2298a6c3e9adaa562a9eec-3"
    # calculatePartSize (minio-js 7.0.12): 64 MiB, +16 MiB until partSize*10000 > size
    MiB = 1 << 20
    framing["calc_part_size"] = [[0, 64 * MiB], [1 << 30, 64 * MiB], [64 * MiB * 10000 - 1, 64 * MiB],
                                 [64 * MiB * 10000, 80 * MiB], [5 << 40, 528 * MiB]]
    out = {"source": "RFC 1321 A.5, FIPS 180-4 examples, OpenSSL 3.0 via hashlib (SURVEY.md §8c)",
           "vectors": vectors, "framing": framing}
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "kat.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
        f.write("\n")
    print("wrote", path, len(vectors), "vectors")


if __name__ == "__main__":
    main()
