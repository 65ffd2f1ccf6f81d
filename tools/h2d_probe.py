#!/usr/bin/env python
"""h2d_probe.py — what the host side of a box can deliver to its GPUs, with no hashing at all.

The end-to-end leg of bench.py is one cudaMemcpyAsync per ring slot from pinned host memory (runtime.cu, enqueue)
with the kernels hidden under it, so its ceiling is whatever plain pinned->device copies reach when the same set of
GPUs copy at once.  This tool measures exactly that: one worker process per GPU (as bench.py's ranks are), each with
its own pinned buffers (cudaHostAllocDefault and cudaHostAllocWriteCombined) and one or two copy streams; for every
subset of GPUs asked for, the participating workers start together and each times its copies with CUDA events on the
copy stream.  A row reports every GPU's rate and the aggregate (total bytes / slowest GPU's time).

  python tools/h2d_probe.py                      # singles, pairs/quads (contiguous and spread), all GPUs
  python tools/h2d_probe.py --subsets 0,1 0,4    # just these

Output: one JSON line per (subset, allocation flag, streams) + a topology record; keep it under profiles/.
"""
from __future__ import annotations

import argparse
import ctypes as C
import glob
import json
import multiprocessing as mp
import os
import subprocess
import sys
import time

GB = 1e9
H2D = 1
FLAGS = {"default": 0, "write_combined": 4}


def cudart() -> C.CDLL:
    """The CUDA runtime torch ships (no torch import needed: the probe is plain cudart)."""
    try:
        return _bind(C.CDLL("libcudart.so.12"))
    except OSError:
        pass
    for base in sys.path:
        for p in glob.glob(os.path.join(base, "nvidia", "cuda_runtime", "lib", "libcudart.so.*")):
            return _bind(C.CDLL(p))
    return _bind(C.CDLL("/usr/local/cuda/lib64/libcudart.so"))


def _bind(L: C.CDLL) -> C.CDLL:
    vp = C.c_void_p
    L.cudaSetDevice.argtypes = [C.c_int]
    L.cudaHostAlloc.argtypes = [C.POINTER(vp), C.c_size_t, C.c_uint]
    L.cudaFreeHost.argtypes = [vp]
    L.cudaMalloc.argtypes = [C.POINTER(vp), C.c_size_t]
    L.cudaFree.argtypes = [vp]
    L.cudaStreamCreateWithFlags.argtypes = [C.POINTER(vp), C.c_uint]
    L.cudaMemcpyAsync.argtypes = [vp, vp, C.c_size_t, C.c_int, vp]
    L.cudaEventCreate.argtypes = [C.POINTER(vp)]
    L.cudaEventRecord.argtypes = [vp, vp]
    L.cudaEventSynchronize.argtypes = [vp]
    L.cudaEventElapsedTime.argtypes = [C.POINTER(C.c_float), vp, vp]
    L.cudaStreamWaitEvent.argtypes = [vp, vp, C.c_uint]
    L.cudaDeviceSynchronize.argtypes = []
    L.cudaGetDeviceCount.argtypes = [C.POINTER(C.c_int)]
    L.cudaGetErrorString.restype = C.c_char_p
    L.cudaGetErrorString.argtypes = [C.c_int]
    return L


def ck(L, rc, what):
    if rc != 0:
        raise RuntimeError(f"{what}: {L.cudaGetErrorString(rc).decode()}")


def timed_copies(L, dst: int, src: int, chunk: int, reps: int, streams: list, ev) -> float:
    """reps copies of `chunk` bytes src->dst, dealt round-robin to `streams`; milliseconds by CUDA events: start is
    recorded on streams[0] before the first copy, end after streams[0] has waited for every other stream's tail."""
    e0, e1, tails = ev
    ck(L, L.cudaDeviceSynchronize(), "sync")
    ck(L, L.cudaEventRecord(e0, streams[0]), "record")
    for s in streams[1:]:
        ck(L, L.cudaStreamWaitEvent(s, e0, 0), "wait")
    ns = len(streams)
    part = chunk // ns // 256 * 256
    for r in range(reps):
        for k, s in enumerate(streams):  # each repetition is split across the streams (same bytes in total)
            ck(L, L.cudaMemcpyAsync(dst + k * part, src + k * part, part, H2D, s), "cudaMemcpyAsync")
    for k, s in enumerate(streams[1:]):
        ck(L, L.cudaEventRecord(tails[k], s), "record")
        ck(L, L.cudaStreamWaitEvent(streams[0], tails[k], 0), "wait")
    ck(L, L.cudaEventRecord(e1, streams[0]), "record")
    ck(L, L.cudaEventSynchronize(e1), "sync")
    ms = C.c_float()
    ck(L, L.cudaEventElapsedTime(C.byref(ms), e0, e1), "elapsed")
    return float(ms.value), part * ns * reps


def worker(dev: int, chunk: int, cmd_q, res_q):
    L = cudart()
    ck(L, L.cudaSetDevice(dev), "cudaSetDevice")
    host = {}
    for name, fl in FLAGS.items():
        p = C.c_void_p()
        ck(L, L.cudaHostAlloc(C.byref(p), chunk, fl), f"cudaHostAlloc({name})")
        C.memset(p, 0x5A, chunk)  # touch every page
        host[name] = p.value
    d = C.c_void_p()
    ck(L, L.cudaMalloc(C.byref(d), chunk), "cudaMalloc")
    streams = []
    for _ in range(2):
        s = C.c_void_p()
        ck(L, L.cudaStreamCreateWithFlags(C.byref(s), 1), "stream")
        streams.append(s)
    evs = []
    for _ in range(3):
        e = C.c_void_p()
        ck(L, L.cudaEventCreate(C.byref(e)), "event")
        evs.append(e)
    ev = (evs[0], evs[1], evs[2:])
    timed_copies(L, d.value, host["default"], chunk, 2, streams[:1], ev)  # warm-up
    res_q.put(("ready", dev))
    while True:
        cmd = cmd_q.get()
        if cmd is None:
            break
        flag, nstreams, reps, start_at = cmd
        while time.time() < start_at:
            pass
        ms, nbytes = timed_copies(L, d.value, host[flag], chunk, reps, streams[:nstreams], ev)
        res_q.put((dev, ms, nbytes))
    L.cudaFree(d)
    for p in host.values():
        L.cudaFreeHost(p)


def topology() -> dict:
    out = {}
    try:
        out["nvidia_smi_topo"] = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=30).stdout
    except Exception as e:  # noqa: BLE001
        out["nvidia_smi_topo"] = repr(e)
    try:
        q = subprocess.run(["nvidia-smi", "--query-gpu=index,pci.bus_id,pcie.link.gen.current,pcie.link.width.current",
                            "--format=csv,noheader"], capture_output=True, text=True, timeout=30).stdout
        out["pcie"] = [l.strip() for l in q.splitlines() if l.strip()]
        numa = {}
        for l in out["pcie"]:
            f = [x.strip() for x in l.split(",")]
            bus = f[1].lower()
            bus = bus[4:] if bus.startswith("0000") and len(bus) > 12 else bus  # 00000000:1B:00.0 -> 0000:1b:00.0
            p = f"/sys/bus/pci/devices/{bus}/numa_node"
            numa[f[0]] = open(p).read().strip() if os.path.exists(p) else "n/a"
        out["numa_node"] = numa
    except Exception as e:  # noqa: BLE001
        out["pcie"] = repr(e)
    out["cpus"] = os.cpu_count()
    try:
        out["lscpu"] = [l for l in subprocess.run(["lscpu"], capture_output=True, text=True).stdout.splitlines()
                        if any(k in l for k in ("Model name", "Socket", "NUMA", "Thread", "Core"))]
    except Exception:  # noqa: BLE001
        pass
    return out


def default_subsets(n: int) -> list:
    subs = [[g] for g in range(n)]
    if n >= 2:
        subs.append([0, 1])
    if n >= 8:
        subs += [[0, 4], [0, 1, 2, 3], [0, 1, 4, 5], [0, 2, 4, 6], list(range(8))]
    elif n >= 4:
        subs += [[0, 2], [0, 1, 2, 3]]
    return subs


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--chunk-mib", type=int, default=2048, help="bytes per copy (bench.py's slots are 2 GiB)")
    ap.add_argument("--reps", type=int, default=8)
    ap.add_argument("--subsets", nargs="*", default=None, help="e.g. 0,1 0,4 0,1,2,3")
    ap.add_argument("--flags", nargs="*", default=list(FLAGS))
    ap.add_argument("--streams", nargs="*", type=int, default=[1, 2])
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    L = cudart()
    n = C.c_int()
    ck(L, L.cudaGetDeviceCount(C.byref(n)), "cudaGetDeviceCount")
    n = n.value
    subsets = [[int(x) for x in s.split(",")] for s in args.subsets] if args.subsets else default_subsets(n)
    chunk = args.chunk_mib << 20
    ctx = mp.get_context("spawn")
    res_q = ctx.Queue()
    cmd_qs, procs = {}, {}
    for g in sorted({g for s in subsets for g in s}):
        cmd_qs[g] = ctx.Queue()
        procs[g] = ctx.Process(target=worker, args=(g, chunk, cmd_qs[g], res_q), daemon=True)
        procs[g].start()
    for _ in procs:
        assert res_q.get(timeout=300)[0] == "ready"
    sink = open(args.out, "w") if args.out else None

    def emit(rec):
        line = json.dumps(rec)
        print(line, flush=True)
        if sink:
            sink.write(line + "\n")
            sink.flush()

    emit({"topology": topology(), "gpus": n, "chunk_bytes": chunk, "reps": args.reps})
    for sub in subsets:
        for flag in args.flags:
            for ns in args.streams:
                start_at = time.time() + 0.25
                for g in sub:
                    cmd_qs[g].put((flag, ns, args.reps, start_at))
                got = [res_q.get(timeout=600) for _ in sub]
                per = {str(g): round(nb / (ms * 1e-3) / GB, 2) for g, ms, nb in got}
                worst = max(ms for _, ms, _ in got)
                total = sum(nb for _, _, nb in got)
                emit({"gpus": sub, "host_alloc": flag, "streams_per_gpu": ns, "per_gpu_gbs": per,
                      "aggregate_gbs": round(total / (worst * 1e-3) / GB, 2), "slowest_ms": round(worst, 2),
                      "timing": "CUDA events on each GPU's copy stream; aggregate = total bytes / slowest GPU"})
    for q in cmd_qs.values():
        q.put(None)
    for p in procs.values():
        p.join(timeout=30)
    if sink:
        sink.close()


if __name__ == "__main__":
    main()
