#!/usr/bin/env python
"""bench.py — the byte-integrity pass of tritonmedia/downloader on B200, one JSON line.

Headline workload = BASELINE.json configs[1]: SHA-1 verification of 65,536 synthetic 256 KiB torrent
pieces (16 GiB) per GPU, fused compare -> verified bitfield (the reference's path:
lib/download.js:64-66 -> webtorrent -> simple-sha1 -> Node crypto).  A "step" is one pass of the hot
path over that batch.

  value     device-resident: inputs already in HBM, one th_sha1_batch launch per step, CUDA events
            on the launching stream, max over ranks; inputs (16 GiB) >> L2 (126 MB), no flush needed.
  e2e       the same pieces through the public API with HOST buffers: pinned ring -> H2D -> kernel ->
            D2H digests + bitfield, copies inside the timed region; beside it the plain pinned->device
            copy rate of the same ranks measured in the same run (e2e.h2d_probe_gbs) and their ratio.
  roofline  integer (alu-pipe) bound of SURVEY §8d next to the HBM figure; see DESIGN.md.
  cpu_baseline / --impl reference   libcrypto EVP (the code Node crypto runs) on the host cores.
  workloads the other two digests and north_star's target shape under the same clock: BASELINE configs[2]
            (MD5 + SHA-256 over 4,096 x 16 MiB parts; at N > 1 both the weak variant and the strong split) and
            the largest 16 MiB-part batch one GPU holds (the chip's fill point), each digest alone and together,
            with roofline fraction, parallelism ceiling, e2e and what was verified against libcrypto.

Every digest of every timed batch is compared with libcrypto (--verify all, the default); the expected table
of the fused piece verify is libcrypto's, not the kernel's.

Multi-GPU: pieces shard by index, one process per GPU (torchrun), no data-path collective; only
digests are gathered on the host (weak scaling: every GPU gets the full per-GPU workload).
`--single-process` instead drives all N GPUs from ONE process and one th_ctx (the N-API add-on's shape).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# kernels of different submits run on different streams; give them distinct hardware queues (read at context creation)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

GB = 1e9
MiB = 1 << 20
SM_COUNT = 148
ALU_LANES_PER_CLK_SM = 64  # B300_MICROARCH.md: alu pipe = 16 lanes/clk/SMSP x 4
# algorithmic 32-bit ops per 64-byte block
OPS_PER_BLOCK = {"sha1": 613, "md5": 324, "sha256": 1400, "md5+sha256": 1724}
METRIC = "hashed GB/s (SHA-1/MD5/SHA-256) at 1/2/4/8 B200 vs Node crypto on host cores"
DLEN = {"sha1": 20, "md5": 16, "sha256": 32}
ALG_BIT = {"sha1": 1, "md5": 2, "sha256": 4}
SEED = 0x747269746F6E

WORKLOADS = {
    # name: (algs, n messages per GPU, message bytes, description)
    "sha1_pieces": ("sha1", 65536, 262144, "SHA-1 verification of 65,536 synthetic 256 KiB torrent pieces (16 GiB) per GPU"),
    "md5_sha256_parts": ("md5+sha256", 4096, 16 << 20, "MD5+SHA-256 over 4,096 x 16 MiB multipart parts (64 GiB) per GPU"),
}
FILL_POINT_PARTS = 9472  # 296 groups = one chain + one helper warp on every pair of SM sub-partitions


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d.get("hbm_gbs", 6650.0)), float(d.get("sm_max_mhz", 1965.0)), "measured"
    return 6650.0, 1965.0, "fallback"


def load_counters():
    """ncu counters of committed captures (profiles/ncu_counters.json): constants read from a file, labelled so."""
    p = os.path.join(ROOT, "profiles", "ncu_counters.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f)
    return {}


class ClockSampler:
    """nvidia-smi -lms during the timed regions (B200_PROFILING.md clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def mark(self):
        return time.time()

    def stop(self):
        if self.proc:
            time.sleep(0.25)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
            self.proc = None

    def summary(self, t0, t1):
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        near = []
        for t, line in list(self.rows):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                c, m = float(f[0]), float(f[1])
            except ValueError:
                continue
            mx = max(mx, m)
            near.append((abs(t - (t0 + t1) / 2), c))
            if t0 - 0.05 <= t <= t1 + 0.15:
                sm.append(c)
                for k, nm in enumerate(names):
                    if f[3 + k].lower().startswith("active"):
                        reasons.add(nm)
        if not sm and near:  # region shorter than the sampling period: use the nearest samples
            sm = [c for _, c in sorted(near)[:3]]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def dist_setup():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist  # plumbing only: barrier, max-reduce of times, digest gather
        dist.init_process_group("gloo")
    return rank, world, local, dist


# ------------------------------------------------------------------------------------------ CPU reference arm
def cpu_reference(algs: str, msg_len: int, n_sample: int, steps: int, warmup: int, threads: int):
    """libcrypto EVP over a bounded sample of the workload; returns (GB/s, s per step, detail)."""
    import oracle
    mask = sum(ALG_BIT[a] for a in algs.split("+"))
    buf = oracle.synth_batch(0, n_sample, msg_len, threads=threads)
    off = (np.arange(n_sample, dtype=np.uint64) * msg_len)
    ln = np.full(n_sample, msg_len, np.uint64)
    chunk = 0 if algs == "sha1" else 65536  # simple-sha1: one update; minio: one update per 64 KiB
    expected = None
    times = []
    for it in range(warmup + steps):
        dt, outs = oracle.evp_batch(mask, buf, off, ln, threads=threads, chunk=chunk)
        if algs == "sha1":  # the piece verify: compare with the expected digest table
            t0 = time.perf_counter()
            if expected is None:
                expected = outs[oracle.SHA1].copy()
            okv = (outs[oracle.SHA1] == expected).all(axis=1)
            dt += time.perf_counter() - t0
            assert okv.all()
        if it >= warmup:
            times.append(dt)
    per = float(np.mean(times))
    return n_sample * msg_len / per / GB, per, {"openssl": oracle.openssl_version()}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    algs, n, ln, desc = WORKLOADS[args.workload]
    threads = os.cpu_count() or 1
    n_sample = max(threads * 8, min(n, (2 << 30) // ln))  # ~2 GiB per step
    v, per, det = cpu_reference(algs, ln, n_sample, args.steps, args.warmup, threads)
    v1, _, _ = cpu_reference(algs, ln, max(8, n_sample // 16), 1, 1, 1)
    line = {
        "impl": "reference", "metric": METRIC, "value": round(v, 3), "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(per * 1e3, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": desc, "algs": algs, "messages_per_step": n_sample, "message_bytes": ln,
                   "note": "reference CPU path = libcrypto EVP (what Node crypto binds), all host threads, messages dealt by index"},
        "cpu_baseline": {"value": round(v, 3), "unit": "GB/s", "cores": threads, "kind": "port",
                         "sample": f"{n_sample} x {ln} B per step, libcrypto {det['openssl']}; 1 thread (Node's own event-loop semantics): {v1:.3f} GB/s",
                         "value_1thread": round(v1, 3)},
        "e2e": {"value": round(v, 3), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    if args.workload == "sha1_pieces" and not args.headline_only:
        # the other digests of the metric on their own shape (16 MiB parts, 64 KiB updates as minio feeds them)
        extra = []
        pl = 16 << 20
        ns = max(threads * 2, (2 << 30) // pl)
        for a in ("md5", "sha256", "md5+sha256"):
            va, pera, _ = cpu_reference(a, pl, ns, 2, 1, threads)
            extra.append({"name": "md5_sha256_parts", "algs": a, "value": round(va, 3), "unit": "GB/s", "cores": threads,
                          "sample": f"{ns} x {pl} B per step x 2 steps", "ms_per_step": round(pera * 1e3, 2)})
        line["workloads"] = extra
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------ GPU arm helpers
class Env:
    pass


def barrier(E):
    E.torch.cuda.synchronize()
    if E.dist is not None:
        E.dist.barrier()


def max_over_ranks(E, x: float) -> float:
    if E.dist is None:
        return x
    tm = E.torch.tensor([x], dtype=E.torch.float64)
    E.dist.all_reduce(tm, op=E.dist.ReduceOp.MAX)
    return float(tm.item())


def make_step(E, data, n, ln, names, outs, expected=None, ok=None):
    """One pass of the hot path over the resident batch: one launch per digest; several digests of the same bytes run
    on separate streams (separate warps), forked from and joined to the timing stream."""
    torch, L = E.torch, E.L
    side = [torch.cuda.Stream() for _ in names[1:]]
    ev_fork = torch.cuda.Event()
    ev_join = [torch.cuda.Event() for _ in side]
    base = data.data_ptr()

    def launch(name, sp):
        if name == "sha1":
            rc = L.th_sha1_batch(base, None, None, None, ln, ln, n, outs["sha1"].data_ptr(),
                                 expected.data_ptr() if expected is not None else None,
                                 ok.data_ptr() if ok is not None else None, sp)
        elif name == "md5":
            rc = L.th_md5_batch(base, None, None, None, ln, ln, n, outs["md5"].data_ptr(), sp)
        else:
            rc = L.th_sha256_batch(base, None, None, None, ln, ln, n, outs["sha256"].data_ptr(), sp)
        assert rc == 0, (name, rc)

    if set(names) == {"md5", "sha256"} and L.th_md5_sha256_fused(ln, ln, n) == 1:
        # few large parts: ONE launch — SHA-256 chain, its schedule helper and an MD5 chain per group, each warp on a
        # sub-partition of its own (digest_parts_kernel; what th_submit launches for MD5 | SHA-256)
        def fused_step():
            rc = L.th_md5_sha256_batch(base, ln, ln, n, outs["md5"].data_ptr(), outs["sha256"].data_ptr(), E.sp)
            assert rc == 0, rc
        return fused_step, 1

    def step():  # launch order = the runtime's (SHA-1, MD5, SHA-256): the shorter MD5 chains are placed first
        if side:
            ev_fork.record(E.stream)
        launch(names[0], E.sp)
        for nm, s2, ej in zip(names[1:], side, ev_join):
            s2.wait_event(ev_fork)
            launch(nm, C.c_void_p(s2.cuda_stream))
            ej.record(s2)
        for ej in ev_join:
            E.stream.wait_event(ej)
    return step, len(names)


def time_steps(E, step, steps, warmup):
    """W untimed steps, then exactly K steps between CUDA events on the launching stream, barrier + synchronize on
    both sides, max over ranks.  Returns (ms per step, host marks of the timed region for the clock sampler)."""
    torch = E.torch
    for _ in range(warmup):
        step()
    barrier(E)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = E.sampler.mark()
    e0.record(E.stream)
    for _ in range(steps):
        step()
    e1.record(E.stream)
    barrier(E)
    t1 = E.sampler.mark()
    return max_over_ranks(E, e0.elapsed_time(e1)) / steps, (t0, t1)


def cpu_digests(E, data, first, count, ln, names, threads):
    """libcrypto digests of messages [first, first+count) of the resident batch: device -> pinned staging -> EVP."""
    import oracle
    torch = E.torch
    mask = sum(ALG_BIT[a] for a in names)
    chunk_msgs = max(1, min(count, E.stage.numel() // ln))
    outs = {a: np.empty((count, DLEN[a]), np.uint8) for a in names}
    upd = 0 if names == ("sha1",) else 65536  # simple-sha1: one update per piece; minio: one per 64 KiB
    for i in range(0, count, chunk_msgs):
        k = min(chunk_msgs, count - i)
        E.stage[:k * ln].copy_(data[(first + i) * ln:(first + i + k) * ln])
        torch.cuda.synchronize()
        _, o = oracle.evp_batch(mask, E.stage.numpy()[:k * ln], np.arange(k, dtype=np.uint64) * ln,
                                np.full(k, ln, np.uint64), threads=threads, chunk=upd)
        for a in names:
            outs[a][i:i + k] = o[ALG_BIT[a]]
    return outs


def verify(E, data, n, ln, outs, mode, threads, known=None):
    """Compares the GPU digests of the timed batch with libcrypto.  mode 'all': every message; 'sample': 64
    stratified.  `known`: {alg: (m, len) array} of digests already verified against libcrypto for the first m
    messages of the same synthetic stream — those are compared directly and, where every digest has them, only
    the rest goes to the CPU.  Returns ({alg: what was checked}, {alg: (n, len) host array})."""
    names = tuple(outs)
    got = {a: outs[a].cpu().numpy().reshape(n, DLEN[a]) for a in names}
    report = {}
    start, via = 0, {}
    if known:
        for a in names:
            if a in known:
                m = min(n, known[a].shape[0])
                assert (got[a][:m] == known[a][:m]).all(), f"{a}: digests differ from the verified ones of the same messages"
                via[a] = m
        if all(a in via for a in names):
            start = min(via.values())
    tail = {a: (f"; first {via[a]} equal the libcrypto-verified digests of the same messages" if a in via else "") for a in names}
    if mode == "all":
        if start < n:
            cpu = cpu_digests(E, data, start, n - start, ln, names, threads)
            for a in names:
                assert (got[a][start:] == cpu[a]).all(), f"{a}: GPU digests differ from libcrypto"
        for a in names:
            if start >= n:
                report[a] = f"all {n}: equal to the libcrypto-verified digests of the same messages"
            else:
                report[a] = (f"all {n} vs libcrypto" if start == 0 else f"all {n}: {n - start} vs libcrypto") + tail[a]
        return report, got
    pick = sorted(set(np.linspace(start, n - 1, 64).astype(int).tolist())) if start < n else []
    for i in pick:
        cpu = cpu_digests(E, data, i, 1, ln, names, threads)
        for a in names:
            assert (got[a][i] == cpu[a][0]).all(), f"{a}: GPU digest {i} differs from libcrypto"
    for a in names:
        report[a] = f"{len(pick)} of {n} sampled vs libcrypto" + tail[a]
    return report, got


def roofline_record(E, algs, per_gpu, n, total_bytes, clocks, counters_key=None):
    """SURVEY §8d roofline of one digest set on one batch shape; counters_key = "<workload>:<algs>" looks up the
    committed ncu capture of that kernel on that shape (profiles/ncu_counters.json)."""
    ops_per_byte = OPS_PER_BLOCK[algs] / 64.0
    clk = E.sm_max_mhz
    int_peak = SM_COUNT * ALU_LANES_PER_CLK_SM * clk * 1e6 / ops_per_byte / GB
    obs = (clocks or {}).get("sm_mhz") or clk
    warps = (n + 31) // 32
    w_sat = 2 if algs == "md5" else 1
    peak = min(int_peak, E.hbm_peak)
    r = {
        "bound": "alu", "achieved": round(per_gpu, 2), "peak": round(peak, 2), "unit": "GB/s",
        "frac": round(per_gpu / peak, 4), "traffic": None,
        "definition": f"min(HBM, {SM_COUNT} SMs x {ALU_LANES_PER_CLK_SM} int lane-ops/clk x {clk:.0f} MHz / {ops_per_byte:.2f} ops/B); per GPU; of {E.peak_src} clock/HBM",
        "frac_at_observed_clock": round(per_gpu / (int_peak * obs / clk), 4),
        "hbm": {"achieved": round(per_gpu, 2), "peak": E.hbm_peak, "frac": round(per_gpu / E.hbm_peak, 4), "source": E.peak_src},
        "algorithmic_bytes_per_launch": total_bytes,
        "parallelism_ceiling": round(min(1.0, warps / (592.0 * w_sat)), 4),
    }
    c = E.counters.get(counters_key or "", None)
    if c:
        # constants from a committed ncu capture of this kernel on this shape — not measured in this run
        r["traffic"] = c.get("dram_bytes_per_launch")
        r["traffic_source"] = c.get("source")
        r["alu_pipe_busy"] = c.get("alu_pipe_busy")
        r["fma_pipe_busy"] = c.get("fma_pipe_busy")
        r["pipe_busy_source"] = c.get("source")
        r["pipe_busy_note"] = ("fraction of cycles the alu / fma pipe was busy in that capture: the tight bound — what is left of "
                               "the alu pipe is the kernel's real headroom, whatever `frac` reads")
    else:
        r["traffic_source"] = None
    return r


_CUDART = None


def cudart():
    global _CUDART
    if _CUDART is None:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import h2d_probe
        _CUDART = h2d_probe.cudart()
    return _CUDART


def h2d_probe(E, host_ptrs, slot_bytes, data, reps=2):
    """Plain pinned->device cudaMemcpyAsync of the e2e leg's own ring slots, every rank at once, no hashing: the
    ceiling the e2e figure is to be read against.  Destination = the slot's own source range of the
    resident batch (same bytes, so nothing changes).  CUDA events on the copy stream, max over ranks."""
    torch = E.torch
    R = cudart()
    barrier(E)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for timed in (False, True):
        if timed:
            barrier(E)
            e0.record(E.stream)
        for _ in range(reps if timed else 1):
            for k, hp in enumerate(host_ptrs):
                rc = R.cudaMemcpyAsync(data.data_ptr() + k * slot_bytes, hp, slot_bytes, 1, E.sp)
                assert rc == 0, rc
        if timed:
            e1.record(E.stream)
        torch.cuda.synchronize()
    ms = max_over_ranks(E, e0.elapsed_time(e1))
    return E.world * reps * len(host_ptrs) * slot_bytes / (ms * 1e-3) / GB


def e2e_leg(E, args, data, n, ln, names, dev_outs, exp_host, corrupt, steps, warmup):
    """The same messages through the public API with HOST buffers (pinned ring -> H2D -> kernels -> D2H digests)."""
    from downloader_b200 import Engine
    torch = E.torch
    slots = args.e2e_slots
    per_slot = n // slots
    # pinned-memory budget: never ask for more than ~45 % of the host's available RAM over all ranks
    avail = _mem_available()
    budget = int(0.45 * avail / E.world) if avail else n * ln
    budget = min(budget, args.e2e_gib << 30)
    while per_slot > 32 and slots * per_slot * ln > budget:
        per_slot //= 2
    slot_bytes = per_slot * ln
    dev_mult = args.e2e_device_mult or (1 if names == ("sha1",) else 3)
    eng = Engine([E.local], ring_bytes=slots * slot_bytes, device_bytes=dev_mult * slots * slot_bytes)
    mask = sum(ALG_BIT[a] for a in names)
    offs_s = np.arange(per_slot, dtype=np.uint64) * ln
    lens_s = np.full(per_slot, ln, np.uint64)
    # fill the pinned ring once (the stand-in for fs.read into pinned memory); contents persist
    held, host_ptrs = [], []
    for k in range(slots):
        sid, view = eng.ring_acquire(slot_bytes)
        torch.from_numpy(view).copy_(data[k * slot_bytes:(k + 1) * slot_bytes])
        held.append(sid)
        host_ptrs.append(view.ctypes.data)
    torch.cuda.synchronize()
    probe = h2d_probe(E, host_ptrs, slot_bytes, data)
    for sid in held:
        eng.ring_release(sid)
    key = ALG_BIT[names[0]]

    # A stream of batches, as the files loop of lib/upload.js is: step s submits its slots, then collects
    # the digests of step s-1 (whose chains ran under this step's copies); the last step's are collected
    # before the clock stops, so every step's H2D and D2H lie inside the timed region.
    def collect(tickets):
        res = [eng.result(t, per_slot, mask, want_ok=exp_host is not None) for t in tickets]
        if E.dist is not None:  # digests are gathered on the host; no GPU collective
            mine = torch.from_numpy(np.concatenate([r[key] for r in res]))
            gl = [torch.empty_like(mine) for _ in range(E.world)] if E.rank == 0 else None
            E.dist.gather(mine, gl, dst=0)
        return res

    def e2e_run(nsteps):
        prev = None
        for _ in range(nsteps):
            tickets = []
            for k in range(slots):
                sid, _ = eng.ring_acquire(slot_bytes)
                exp_k = exp_host[k * per_slot:(k + 1) * per_slot] if exp_host is not None else None
                tickets.append(eng.submit(sid, offs_s, lens_s, mask, exp_k))
            if prev is not None:
                collect(prev)
            prev = tickets
        return collect(prev)

    e2e_run(max(1, min(warmup, 2)))
    s0 = eng.stats()
    barrier(E)
    w0 = time.perf_counter()
    res = e2e_run(steps)
    barrier(E)
    w1 = time.perf_counter()
    s1 = eng.stats()
    dt = max_over_ranks(E, (w1 - w0) / steps)
    # parity of the e2e path with the (verified) device-resident digests
    m = per_slot * slots
    for a in names:
        got = np.concatenate([r[ALG_BIT[a]] for r in res])
        assert (got == dev_outs[a][:m]).all(), f"e2e {a} digests differ"
    if exp_host is not None:
        okv = np.concatenate([r["ok"] for r in res])
        assert sorted(np.nonzero(~okv)[0].tolist()) == [c for c in corrupt if c < m]
    value = E.world * slots * slot_bytes / dt / GB
    out = {"value": round(value, 3), "unit": "GB/s",
           "h2d_bytes_per_step": int((s1["h2d_bytes"] - s0["h2d_bytes"]) // steps),
           "d2h_bytes_per_step": int((s1["d2h_bytes"] - s0["d2h_bytes"]) // steps),
           "messages_per_gpu": m, "steps": steps, "ms_per_step": round(dt * 1e3, 3),
           "timing": "host wall clock around %d streamed steps (submit all slots, collect the previous step's digests; last step drained inside), max over ranks" % steps,
           "pinned_ring_bytes": slots * slot_bytes, "device_ring_bytes": dev_mult * slots * slot_bytes,
           "launches_per_step": int((s1["launches"] - s0["launches"]) // steps),
           "overflow_allocs": int(s1["overflow_allocs"]),
           "h2d_probe_gbs": round(probe, 3),
           "h2d_probe": "plain cudaMemcpyAsync of the same pinned slots by the same ranks at once, no hashing, CUDA events, max over ranks",
           "frac_of_h2d_ceiling": round(value / probe, 4),
           "verified": "equal to the device-resident digests of the same messages (verified vs libcrypto)"}
    eng.close()
    return out


def run_workload(E, args, name, algs, n, ln, desc, first_msg, steps, warmup, do_e2e, verify_mode, data=None,
                 variants=None, known=None, counters_key=None):
    """Times one batch shape: `algs` together (the entry's value) and, if asked, each variant alone."""
    torch, L = E.torch, E.L
    total_bytes = n * ln
    names = tuple(algs.split("+"))
    own = data is None
    if own:
        data = torch.empty(total_bytes, dtype=torch.uint8, device="cuda")
        assert L.th_synth_fill(data.data_ptr(), first_msg, n, ln, ln, SEED, E.sp) == 0
        torch.cuda.synchronize()
    threads = max(1, (os.cpu_count() or 1) // E.world)
    all_names = tuple(dict.fromkeys(names + tuple(a for v in (variants or []) for a in v.split("+"))))
    outs = {a: torch.zeros(n * DLEN[a], dtype=torch.uint8, device="cuda") for a in all_names}
    expected = ok = exp_host = None
    corrupt = []
    cpu_sha1 = None
    if names == ("sha1",):
        # expected table of the fused verify = libcrypto's digests of every piece (not the kernel's), K corrupted
        ok = torch.zeros((n + 31) // 32, dtype=torch.int32, device="cuda")
        if verify_mode == "all":
            cpu_sha1 = cpu_digests(E, data, 0, n, ln, ("sha1",), threads)["sha1"]
            exp_host = cpu_sha1.copy()
        else:  # tuning runs only: the table comes from the kernel, a sample is checked below
            tmp = torch.zeros(n * 20, dtype=torch.uint8, device="cuda")
            assert L.th_sha1_batch(data.data_ptr(), None, None, None, ln, ln, n, tmp.data_ptr(), None, None, E.sp) == 0
            exp_host = tmp.cpu().numpy().reshape(n, 20).copy()
        corrupt = sorted({0, n // 3, n - 1})
        for b in corrupt:
            exp_host[b, 5] ^= 0x10
        expected = torch.from_numpy(exp_host).cuda()

    entry = {"name": name, "config": desc, "algs": algs, "messages_per_gpu": n, "message_bytes": ln,
             "l2": "inputs (%.1f GiB per GPU) exceed the 126 MB L2; no flush" % (total_bytes / 2**30)}
    step, per_step = make_step(E, data, n, ln, names, outs, expected, ok)
    ms, marks = time_steps(E, step, steps, warmup)
    clocks = E.sampler.summary(*marks)
    per_gpu = total_bytes / (ms * 1e-3) / GB
    entry.update({"value": round(E.world * per_gpu, 3), "unit": "GB/s", "ms_per_step": round(ms, 4), "steps": steps,
                  "warmup": warmup, "gpu_launches": per_step * steps,
                  "roofline": roofline_record(E, algs, per_gpu, n, total_bytes, clocks, f"{counters_key}:{algs}" if counters_key else None),
                  "clocks": clocks, "kernel": _plan(algs, ln, n)})
    if variants:
        entry["digests"] = {}
        for v in variants:
            vn = tuple(v.split("+"))
            if vn == names:
                continue
            vstep, vper_step = make_step(E, data, n, ln, vn, outs)
            vms, vmarks = time_steps(E, vstep, steps, warmup)
            vper = total_bytes / (vms * 1e-3) / GB
            entry["digests"][v] = {"value": round(E.world * vper, 3), "ms_per_step": round(vms, 4), "gpu_launches": vper_step * steps,
                                   "roofline": roofline_record(E, v, vper, n, total_bytes, E.sampler.summary(*vmarks),
                                                               f"{counters_key}:{v}" if counters_key else None),
                                   "kernel": _plan(v, ln, n)}
    # ---- correctness of what was timed -------------------------------------------------------------------
    if cpu_sha1 is not None:  # every piece's digest was computed by libcrypto above (it made the expected table)
        got = {"sha1": outs["sha1"].cpu().numpy().reshape(n, 20)}
        assert (got["sha1"] == cpu_sha1).all(), "sha1: GPU digests differ from libcrypto"
        report = {"sha1": f"all {n} vs libcrypto"}
    else:
        report, got = verify(E, data, n, ln, outs, verify_mode, threads, known=known)
    if ok is not None:
        words = ok.cpu().numpy().view(np.uint32)
        bits = ((words[np.arange(n) >> 5] >> (np.arange(n) & 31)) & 1).astype(bool)
        assert sorted(np.nonzero(~bits)[0].tolist()) == corrupt, "verify bitfield wrong"
        report["bitfield"] = f"flags exactly the {len(corrupt)} corrupted of {n} (expected table: libcrypto)" if verify_mode == "all" \
            else f"flags exactly the {len(corrupt)} corrupted of {n}"
    entry["verified_vs_oracle"] = report
    if do_e2e:
        # 5 streamed steps whatever `steps` is: the last step's chains drain after the last copy, and with 16 MiB parts
        # that tail (one chain latency, 0.24 s) is most of a step — over 3 steps it alone costs 20 % of the figure
        entry["e2e"] = e2e_leg(E, args, data, n, ln, names, got, exp_host, corrupt, max(1, min(args.steps, 5)), warmup)
    if own:
        del data
    del outs, expected, ok
    torch.cuda.empty_cache()
    return entry, got


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="sha1_pieces", choices=sorted(WORKLOADS))
    ap.add_argument("--headline-only", action="store_true", help="skip the `workloads` array (configs[2] and the fill point)")
    ap.add_argument("--verify", default="all", choices=["all", "sample"], help="compare every digest of every timed batch with libcrypto (default) or a 64-message sample")
    ap.add_argument("--messages", type=int, default=0, help="override messages per GPU (not the BASELINE config)")
    ap.add_argument("--strong", action="store_true", help="divide the workload's messages over the ranks (BASELINE config 3 as stated) instead of giving each GPU all of them")
    ap.add_argument("--skip-e2e", action="store_true")
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--e2e-slots", type=int, default=8)
    ap.add_argument("--e2e-gib", type=int, default=16, help="cap on the pinned bytes the e2e leg cycles through per GPU")
    ap.add_argument("--e2e-device-mult", type=int, default=0,
                    help="device staging ring = this many pinned rings (0: 1 for pieces, 3 for 16 MiB parts, whose chains outlast the copy)")
    ap.add_argument("--staging", default="", help="override bulk staging geometry, e.g. 512x2 (tuning only)")
    ap.add_argument("--path", type=int, default=0, help="0 auto, 1 bulk, 2 tile, 3 sliced (tuning only)")
    ap.add_argument("--device-order", default="auto", help="comma-separated CUDA devices, rank r uses entry r; 'auto' interleaves the two halves of the device list, 'identity' is LOCAL_RANK")
    ap.add_argument("--single-process", action="store_true", help="drive --gpus N devices from ONE process and one th_ctx (no torchrun)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        return run_reference(args)
    if args.single_process:
        return run_single_process(args)

    import torch

    from downloader_b200 import _capi

    E = Env()
    E.torch = torch
    E.rank, E.world, local, E.dist = dist_setup()
    order, order_why = device_order(args.device_order, torch.cuda.device_count(), E.world)
    E.local = order[local]
    torch.cuda.set_device(E.local)
    E.L = L = _capi.lib()
    if args.staging:
        ch_, st_ = map(int, args.staging.split("x"))
        assert L.th_set_staging(7, ch_, st_) == 0, "no such staging variant"
    assert L.th_set_option(b"path", args.path) == 0
    E.hbm_peak, E.sm_max_mhz, E.peak_src = load_peaks()
    E.counters = load_counters()
    E.stream = torch.cuda.current_stream()
    E.sp = C.c_void_p(E.stream.cuda_stream)
    E.stage = torch.empty(1 << 30, dtype=torch.uint8).pin_memory()  # verification staging: device -> host -> libcrypto
    E.sampler = ClockSampler(E.local)
    E.sampler.start()

    algs, n, ln, desc = WORKLOADS[args.workload]
    if args.messages:
        n = args.messages
    if args.strong:
        n = max(32, n // E.world)
        desc += " / %d ranks (strong scaling)" % E.world
    first_msg = E.rank * n  # shard by index: rank r owns messages [r*n, (r+1)*n)
    head, head_digs = run_workload(E, args, args.workload, algs, n, ln, desc, first_msg, args.steps, args.warmup,
                                   not args.skip_e2e, args.verify,
                                   counters_key=args.workload if not args.messages and not args.strong else None)

    line = {
        "metric": METRIC, "value": head["value"], "unit": "GB/s", "n_gpus": E.world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True,
        "scaling": "strong" if args.strong else "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": desc, "algs": algs, "messages_per_gpu": n, "message_bytes": ln, "l2": head["l2"],
                   "kernel": head["kernel"], "verified_vs_oracle": head["verified_vs_oracle"],
                   "fused_compare": algs == "sha1", "devices": {"rank_to_cuda": order[:E.world], "why": order_why}},
        "roofline": head["roofline"], "clocks": head["clocks"], "gpu_launches": head["gpu_launches"],
    }
    if "e2e" in head:
        line["e2e"] = head["e2e"]
    if "digests" in head:
        line["digests"] = head["digests"]
    del head_digs

    # ---- the other digests and the 16 MiB-part target shape, under the same clock --------------------------
    if args.workload == "sha1_pieces" and not args.headline_only and not args.messages:
        line["workloads"] = extra_workloads(E, args)

    E.sampler.stop()
    if E.rank == 0 and E.world == 1 and not args.skip_cpu:
        threads = os.cpu_count() or 1
        n_sample = max(threads * 8, min(n, (2 << 30) // ln))
        v, per, det = cpu_reference(algs, ln, n_sample, 3, 1, threads)
        v1, _, _ = cpu_reference(algs, ln, max(8, n_sample // 16), 1, 1, 1)
        line["cpu_baseline"] = {"value": round(v, 3), "unit": "GB/s", "cores": threads, "kind": "port",
                                "sample": f"{n_sample} x {ln} B x 3 passes, libcrypto EVP ({det['openssl']}) = the code Node crypto runs; "
                                          f"1 thread (reference semantics): {v1:.3f} GB/s",
                                "value_1thread": round(v1, 3)}
    if E.rank == 0:
        print(json.dumps(line))
    if E.dist is not None:
        E.dist.barrier()
        E.dist.destroy_process_group()
    return 0


def extra_workloads(E, args):
    """BASELINE configs[2] (weak and, at N > 1, the strong split) and the fill-point batch of 16 MiB parts."""
    torch, L = E.torch, E.L
    steps, warmup = max(1, min(args.steps, 3)), min(args.warmup, 3)
    algs, n, ln, desc = WORKLOADS["md5_sha256_parts"]
    out = []
    free, _ = torch.cuda.mem_get_info()
    if free < n * ln + (52 << 30):
        return [{"name": "md5_sha256_parts", "skipped": "not enough free device memory (%.0f GiB)" % (free / 2**30)}]
    first = E.rank * n
    data = torch.empty(n * ln, dtype=torch.uint8, device="cuda")
    assert L.th_synth_fill(data.data_ptr(), first, n, ln, ln, SEED, E.sp) == 0
    torch.cuda.synchronize()
    weak, digs = run_workload(E, args, "md5_sha256_parts", algs, n, ln,
                              "BASELINE configs[2], weak variant: " + desc + "; MD5 and SHA-256 kernels run concurrently; `digests` = each alone",
                              first, steps, warmup, not args.skip_e2e, args.verify, data=data, variants=["md5", "sha256"],
                              counters_key="md5_sha256_parts")
    weak["scaling"] = "weak"
    out.append(weak)
    if E.world > 1:
        ns = max(32, n // E.world)
        strong, _ = run_workload(E, args, "md5_sha256_parts_strong", algs, ns, ln,
                                 "BASELINE configs[2] as stated: 4,096 x 16 MiB parts split over %d GPUs (%d per GPU; each GPU's share is the head of its weak batch)" % (E.world, ns),
                                 first, steps, warmup, False, args.verify, data=data[:ns * ln], variants=["md5", "sha256"], known=digs)
        strong["scaling"] = "strong"
        strong["note"] = "flat by construction: the step is one SHA-256 chain latency however many GPUs share the parts (SURVEY §7.3-1)"
        out.append(strong)
    del data
    torch.cuda.empty_cache()
    # ---- fill point: the largest 16 MiB-part batch one GPU holds -------------------------------------------
    free, _ = torch.cuda.mem_get_info()
    nf = min(FILL_POINT_PARTS, int((free - (3 << 30)) // ln) // 32 * 32)
    if nf < n:
        out.append({"name": "parts_fill_point", "skipped": "not enough free device memory (%.0f GiB)" % (free / 2**30)})
        return out
    data = torch.empty(nf * ln, dtype=torch.uint8, device="cuda")
    assert L.th_synth_fill(data.data_ptr(), first, nf, ln, ln, SEED, E.sp) == 0
    torch.cuda.synchronize()
    fill, _ = run_workload(E, args, "parts_fill_point", algs, nf, ln,
                           "fill point: %d x 16 MiB parts (%.0f GiB) per GPU = %d groups, a chain warp + a helper warp on every pair of SM "
                           "sub-partitions; north_star's >= 70 %% bar is stated on 16 MiB-part batches; messages [0, 4096) are configs[2]'s"
                           % (nf, nf * ln / 2**30, (nf + 31) // 32),
                           first, steps, warmup, False, "sample", data=data, variants=["sha1", "md5", "sha256"], known=digs,
                           counters_key="parts_fill_point" if nf == FILL_POINT_PARTS else None)
    fill["scaling"] = "weak"
    out.append(fill)
    del data
    torch.cuda.empty_cache()
    return out


def run_single_process(args):
    """One process, one th_ctx over N GPUs (the N-API add-on's real shape; lib/upload.js:34-52 is one event loop
    feeding everything): the headline pieces through the ring of every GPU, fed (a) by one thread that deals slots
    round-robin and (b) by one feeder thread per GPU.  Prints one JSON line (mode: single_process)."""
    import torch

    from downloader_b200 import ALG_SHA1, Engine, _capi
    L = _capi.lib()
    N = args.gpus
    algs, n, ln, desc = WORKLOADS["sha1_pieces"]
    slots = args.e2e_slots
    avail = _mem_available()
    budget = min(int(0.45 * avail / N) if avail else n * ln, args.e2e_gib << 30)
    per_slot = n // slots
    while per_slot > 32 and slots * per_slot * ln > budget:
        per_slot //= 2
    slot_bytes = per_slot * ln
    eng = Engine(list(range(N)), ring_bytes=slots * slot_bytes)
    offs = np.arange(per_slot, dtype=np.uint64) * ln
    lens = np.full(per_slot, ln, np.uint64)
    ref = []
    for g in range(N):  # fill every ring once with that GPU's shard of the synthetic stream; keep its device digests
        torch.cuda.set_device(g)
        d = torch.empty(slots * slot_bytes, dtype=torch.uint8, device="cuda")
        assert L.th_synth_fill(d.data_ptr(), g * n, slots * per_slot, ln, ln, SEED, None) == 0
        o = torch.zeros(slots * per_slot * 20, dtype=torch.uint8, device="cuda")
        assert L.th_sha1_batch(d.data_ptr(), None, None, None, ln, ln, slots * per_slot, o.data_ptr(), None, None, None) == 0
        held = []
        for k in range(slots):
            sid, view = eng.ring_acquire(slot_bytes, gpu=g)
            torch.from_numpy(view).copy_(d[k * slot_bytes:(k + 1) * slot_bytes])
            held.append(sid)
        torch.cuda.synchronize()
        for sid in held:
            eng.ring_release(sid)
        ref.append(o.cpu().numpy().reshape(-1, 20))
        del d, o
        torch.cuda.empty_cache()
    steps = max(1, min(args.steps, 5))

    def feed(gpus, nsteps, check):
        prev = []
        for s in range(nsteps + 1):
            cur = []
            if s < nsteps:
                for k in range(slots):
                    for g in gpus:  # round-robin over the GPUs this thread feeds
                        sid, _ = eng.ring_acquire(slot_bytes, gpu=g)
                        cur.append((g, k, eng.submit(sid, offs, lens, ALG_SHA1)))
            for g, k, t in prev:
                r = eng.result(t, per_slot, ALG_SHA1)
                if check:
                    assert (r[ALG_SHA1] == ref[g][k * per_slot:(k + 1) * per_slot]).all(), "digest mismatch"
            prev = cur

    results = {}
    for mode in ("one_thread", "thread_per_gpu"):
        groups = [list(range(N))] if mode == "one_thread" else [[g] for g in range(N)]
        for timed in (False, True):
            ths = [threading.Thread(target=feed, args=(gr, steps if timed else 1, not timed)) for gr in groups]
            w0 = time.perf_counter()
            for t in ths:
                t.start()
            for t in ths:
                t.join()
            dt = time.perf_counter() - w0
        results[mode] = {"value": round(N * steps * slots * slot_bytes / dt / GB, 3), "unit": "GB/s",
                         "ms_per_step": round(dt / steps * 1e3, 2), "feeder_threads": len(groups)}
    st = eng.stats()
    eng.close()
    print(json.dumps({"mode": "single_process", "metric": METRIC, "n_gpus": N, "steps": steps, "workload": desc,
                      "pinned_ring_bytes_per_gpu": slots * slot_bytes, "e2e_single_process": results,
                      "overflow_allocs": st["overflow_allocs"],
                      "timing": "host wall clock around the streamed steps, last step drained inside; digests checked in the warm-up pass"}))
    return 0


def device_order(spec, count, world):
    """Which CUDA device each local rank uses.  'auto': the two halves of the device list interleaved (0, n/2, 1, n/2+1,
    ...).  On this pool's 8-GPU boxes GPUs 0-3 and 4-7 sit behind two host bridges that each deliver ~115 GB/s of
    pinned->device copies however many of their GPUs copy (profiles/r2_h2d_probe_8gpu.jsonl: {0,1,2,3} 115 GB/s,
    {0,1,4,5} 212 GB/s), so a 2- or 4-rank run that takes its GPUs from both halves has twice the end-to-end ceiling of
    one that takes GPUs 0..N-1.  With every GPU in use the order makes no difference."""
    if spec not in ("auto", "identity", ""):
        return [int(x) for x in spec.split(",")], "given on the command line"
    if spec == "auto" and count >= 4 and count % 2 == 0 and world <= count:
        half = count // 2
        order = [x for pair in zip(range(half), range(half, count)) for x in pair]
        return order, "halves of the device list interleaved: GPUs 0..n/2-1 and n/2..n-1 share one host bridge each (profiles/r2_h2d_probe_8gpu.jsonl)"
    return list(range(max(count, world))), "LOCAL_RANK"


def _mem_available():
    try:
        with open("/proc/meminfo") as f:
            for line in f:
                if line.startswith("MemAvailable:"):
                    return int(line.split()[1]) * 1024
    except OSError:
        pass
    return 0


def _plan(algs, ln, n):
    from downloader_b200 import _capi
    names = algs.split("+")
    if set(names) == {"md5", "sha256"} and _capi.lib().th_md5_sha256_fused(ln, ln, n) == 1:
        return "parts: one launch, per group a SHA-256 chain warp + schedule-helper warp + MD5 chain warp on separate sub-partitions; SHA-256 half: " + _capi.plan(4, ln, ln, n)
    if len(names) == 1:
        return _capi.plan(ALG_BIT[names[0]], ln, ln, n)
    return {a: _capi.plan(ALG_BIT[a], ln, ln, n) for a in names}


if __name__ == "__main__":
    sys.exit(main())
