"""In-process S3 mock (BASELINE config 1) — TEST INFRASTRUCTURE, lives with the oracle on purpose.

It plays the *server*: it re-hashes every body it receives on the CPU (hashlib = OpenSSL) and rejects a
part whose `Content-MD5` / `x-amz-content-sha256` header does not match, returns the part ETag (hex MD5)
and, on completion, `md5(md5_1 || ... || md5_N)-N`.  Method names follow minio-js.  Bodies are not kept.
"""
from __future__ import annotations

import base64
import hashlib


class S3Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"{code}: {msg}")
        self.code = code


class S3Mock:
    def __init__(self):
        self.buckets = {}
        self.uploads = {}
        self.requests = []
        self._next = 1

    def bucketExists(self, b):
        return b in self.buckets

    def makeBucket(self, b):
        self.buckets.setdefault(b, {})

    def _check(self, body: bytes, headers: dict):
        md5 = hashlib.md5(body).digest()
        if "Content-MD5" in headers and base64.b64decode(headers["Content-MD5"]) != md5:
            raise S3Error("BadDigest", "The Content-MD5 you specified did not match what we received.")
        sha = headers.get("x-amz-content-sha256")
        if sha and sha != "UNSIGNED-PAYLOAD":
            if len(sha) != 64:  # minio makeRequestStream rejects other lengths before sending
                raise S3Error("InvalidArgument", "x-amz-content-sha256 must be 64 hex characters")
            if sha != hashlib.sha256(body).hexdigest():
                raise S3Error("XAmzContentSHA256Mismatch", "The provided 'x-amz-content-sha256' header does not match what was computed.")
        return md5

    def putObject(self, bucket, name, body, headers=None):
        if isinstance(body, str):
            body = body.encode()
        headers = headers or {}
        md5 = self._check(bytes(body), headers)
        self.requests.append(("PUT", bucket, name, len(body), dict(headers)))
        etag = md5.hex()
        self.buckets.setdefault(bucket, {})[name] = {"etag": etag, "size": len(body)}
        return etag

    def newMultipartUpload(self, bucket, name):
        uid = f"upload-{self._next}"
        self._next += 1
        self.uploads[uid] = {"bucket": bucket, "name": name, "parts": {}}
        return uid

    def putPart(self, bucket, name, upload_id, part_number, body, headers=None):
        headers = headers or {}
        md5 = self._check(bytes(body), headers)
        self.requests.append(("PUT-PART", bucket, name, part_number, len(body), dict(headers)))
        self.uploads[upload_id]["parts"][part_number] = {"etag": md5.hex(), "md5": md5, "size": len(body)}
        return '"' + md5.hex() + '"'

    def listParts(self, bucket, name, upload_id):
        return {n: p["etag"] for n, p in self.uploads[upload_id]["parts"].items()}

    def completeMultipartUpload(self, bucket, name, upload_id, etags):
        up = self.uploads.pop(upload_id)
        parts = [up["parts"].get(i + 1) for i in range(len(etags))]
        for i, (p, e) in enumerate(zip(parts, etags)):
            if p is None or p["etag"] != e.strip('"'):
                raise S3Error("InvalidPart", f"part {i + 1} missing or ETag mismatch")
        etag = hashlib.md5(b"".join(p["md5"] for p in parts)).hexdigest() + f"-{len(parts)}"
        self.buckets.setdefault(bucket, {})[name] = {"etag": etag, "size": sum(p["size"] for p in parts)}
        return etag

    def statObject(self, bucket, name):
        return self.buckets[bucket][name]
