#!/usr/bin/env python
"""bench.py — the byte-integrity pass of tritonmedia/downloader on B200, one JSON line.

Default workload = BASELINE.json configs[1]: SHA-1 verification of 65,536 synthetic 256 KiB torrent
pieces (16 GiB) per GPU, fused compare -> verified bitfield (the reference's path:
lib/download.js:64-66 -> webtorrent -> simple-sha1 -> Node crypto).  A "step" is one pass of the hot
path over that batch.

  value     device-resident: inputs already in HBM, one th_sha1_batch launch per step, CUDA events
            on the launching stream, max over ranks; inputs (16 GiB) >> L2 (126 MB), no flush needed.
  e2e       the same pieces through the public API with HOST buffers: pinned ring -> H2D -> kernel ->
            D2H digests + bitfield, copies inside the timed region.
  roofline  integer (alu-pipe) bound of SURVEY §8d next to the HBM figure; see DESIGN.md.
  cpu_baseline / --impl reference   libcrypto EVP (the code Node crypto runs) on the host cores.

Multi-GPU: pieces shard by index, one process per GPU (torchrun), no data-path collective; only
digests are gathered on the host (weak scaling: every GPU gets the full per-GPU workload).
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
SM_COUNT = 148
ALU_LANES_PER_CLK_SM = 64  # B300_MICROARCH.md: alu pipe = 16 lanes/clk/SMSP x 4
# algorithmic 32-bit ops per 64-byte block
OPS_PER_BLOCK = {"sha1": 613, "md5": 324, "sha256": 1400, "md5+sha256": 1724}
METRIC = "hashed GB/s (SHA-1/MD5/SHA-256) at 1/2/4/8 B200 vs Node crypto on host cores"

WORKLOADS = {
    # name: (algs, n messages per GPU, message bytes, description)
    "sha1_pieces": ("sha1", 65536, 262144, "SHA-1 verification of 65,536 synthetic 256 KiB torrent pieces (16 GiB) per GPU"),
    "md5_sha256_parts": ("md5+sha256", 4096, 16 << 20, "MD5+SHA-256 over 4,096 x 16 MiB multipart parts (64 GiB) per GPU"),
}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d.get("hbm_gbs", 6650.0)), float(d.get("sm_max_mhz", 1965.0)), "measured"
    return 6650.0, 1965.0, "fallback"


class ClockSampler:
    """nvidia-smi -lms 200 during the timed region (B200_PROFILING.md clocks line)."""
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

    def summary(self, t0, t1):
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in self.rows:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                c, m = float(f[0]), float(f[1])
            except ValueError:
                continue
            mx = max(mx, m)
            if t0 - 0.05 <= t <= t1 + 0.15:
                sm.append(c)
                for k, nm in enumerate(names):
                    if f[3 + k].lower().startswith("active"):
                        reasons.add(nm)
        if not sm:  # region shorter than the sampling period: use the nearest samples
            sm = [float(l.split(",")[0]) for _, l in self.rows[-3:] if l.split(",")[0].strip().replace(".", "").isdigit()]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def dist_setup(n_gpus):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist  # plumbing only: barrier, max-reduce of times, digest gather
        dist.init_process_group("gloo")
    return rank, world, local, dist


def cpu_reference(algs: str, msg_len: int, n_sample: int, steps: int, warmup: int, threads: int):
    """libcrypto EVP over a bounded sample of the workload; returns (GB/s, s per step, detail)."""
    import oracle
    mask = {"sha1": oracle.SHA1, "md5+sha256": oracle.MD5 | oracle.SHA256}[algs]
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
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="sha1_pieces", choices=sorted(WORKLOADS))
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
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        return run_reference(args)

    import torch

    from downloader_b200 import ALG_MD5, ALG_SHA1, ALG_SHA256, Engine, _capi

    rank, world, local, dist = dist_setup(args.gpus)
    torch.cuda.set_device(local)
    L = _capi.lib()
    if args.staging:
        ch_, st_ = map(int, args.staging.split("x"))
        assert L.th_set_staging(7, ch_, st_) == 0, "no such staging variant"
    assert L.th_set_option(b"path", args.path) == 0
    algs, n, ln, desc = WORKLOADS[args.workload]
    if args.messages:
        n = args.messages
    if args.strong:
        n = max(32, n // world)
        desc += " / %d ranks (strong scaling)" % world
    hbm_peak, sm_max_mhz, peak_src = load_peaks()
    total_bytes = n * ln
    first_msg = rank * n  # shard by index: rank r owns messages [r*n, (r+1)*n)
    stream = torch.cuda.current_stream()
    sp = C.c_void_p(stream.cuda_stream)

    # ---- device-resident set-up (untimed): synthetic pieces + expected digests ---------------------
    data = torch.empty(total_bytes, dtype=torch.uint8, device="cuda")
    assert L.th_synth_fill(data.data_ptr(), first_msg, n, ln, ln, 0x747269746F6E, sp) == 0
    outs, ok = {}, None
    if algs == "sha1":
        outs["sha1"] = torch.zeros(n * 20, dtype=torch.uint8, device="cuda")
        expected = torch.zeros(n * 20, dtype=torch.uint8, device="cuda")
        ok = torch.zeros((n + 31) // 32, dtype=torch.int32, device="cuda")
        assert L.th_sha1_batch(data.data_ptr(), None, None, None, ln, ln, n, expected.data_ptr(), None, None, sp) == 0
        corrupt = sorted({0, n // 3, n - 1})
        for b in corrupt:
            expected.view(n, 20)[b, 5] ^= 0x10

        def step():
            rc = L.th_sha1_batch(data.data_ptr(), None, None, None, ln, ln, n, outs["sha1"].data_ptr(),
                                 expected.data_ptr(), ok.data_ptr(), sp)
            assert rc == 0, rc
        launches_per_step = 1
    else:
        outs["md5"] = torch.zeros(n * 16, dtype=torch.uint8, device="cuda")
        outs["sha256"] = torch.zeros(n * 32, dtype=torch.uint8, device="cuda")
        s2 = torch.cuda.Stream()
        ev_fork, ev_join = torch.cuda.Event(), torch.cuda.Event()

        def step():  # MD5 and SHA-256 of the same parts on two streams (separate warps)
            ev_fork.record(stream)
            s2.wait_event(ev_fork)
            assert L.th_md5_batch(data.data_ptr(), None, None, None, ln, ln, n, outs["md5"].data_ptr(), C.c_void_p(s2.cuda_stream)) == 0
            assert L.th_sha256_batch(data.data_ptr(), None, None, None, ln, ln, n, outs["sha256"].data_ptr(), sp) == 0
            ev_join.record(s2)
            stream.wait_event(ev_join)
        launches_per_step = 2

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()

    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(args.warmup):
        step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = sampler.mark()
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    barrier()
    t1 = sampler.mark()
    ms = e0.elapsed_time(e1)
    if dist is not None:
        tm = torch.tensor([ms], dtype=torch.float64)
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        ms = float(tm.item())
    ms_per_step = ms / args.steps
    value = world * total_bytes / (ms_per_step * 1e-3) / GB
    per_gpu = total_bytes / (ms_per_step * 1e-3) / GB

    # ---- correctness of what was timed (rank-local, bounded) ------------------------------------------
    import oracle
    checked = {}
    pick = sorted(set(np.linspace(0, n - 1, 24).astype(int).tolist()))
    host = np.concatenate([data[i * ln:(i + 1) * ln].cpu().numpy() for i in pick])
    offs = (np.arange(len(pick)) * ln).astype(np.uint64)
    lens = np.full(len(pick), ln, np.uint64)
    for name, alg in (("sha1", oracle.SHA1), ("md5", oracle.MD5), ("sha256", oracle.SHA256)):
        if name in outs:
            got = outs[name].cpu().numpy().reshape(n, -1)[pick]
            exp = oracle.batch(alg, host, offs, lens, threads=min(16, os.cpu_count() or 1))
            assert (got == exp).all(), f"{name}: GPU digests differ from the oracle"
            checked[name] = len(pick)
    if ok is not None:
        words = ok.cpu().numpy().view(np.uint32)
        bits = ((words[np.arange(n) >> 5] >> (np.arange(n) & 31)) & 1).astype(bool)
        assert sorted(np.nonzero(~bits)[0].tolist()) == corrupt, "verify bitfield wrong"

    # ---- end to end: pinned host ring -> H2D -> kernel -> D2H digests --------------------------------
    e2e = None
    if not args.skip_e2e:
        slots = args.e2e_slots
        per_slot = n // slots
        # pinned-memory budget: never ask for more than ~45 % of the host's available RAM over all ranks
        avail = _mem_available()
        budget = int(0.45 * avail / world) if avail else total_bytes
        budget = min(budget, args.e2e_gib << 30)
        while per_slot > 32 and slots * per_slot * ln > budget:
            per_slot //= 2
        slot_bytes = per_slot * ln
        dev_mult = args.e2e_device_mult or (1 if algs == "sha1" else 3)
        eng = Engine([local], ring_bytes=slots * slot_bytes, device_bytes=dev_mult * slots * slot_bytes)
        mask = ALG_SHA1 if algs == "sha1" else (ALG_MD5 | ALG_SHA256)
        offs_s = np.arange(per_slot, dtype=np.uint64) * ln
        lens_s = np.full(per_slot, ln, np.uint64)
        exp_host = expected.cpu().numpy().reshape(n, 20) if algs == "sha1" else None
        # fill the pinned ring once (the stand-in for fs.read into pinned memory); contents persist
        held = []
        for k in range(slots):
            sid, view = eng.ring_acquire(slot_bytes)
            torch.from_numpy(view).copy_(data[k * slot_bytes:(k + 1) * slot_bytes])
            held.append(sid)
        torch.cuda.synchronize()
        for sid in held:
            eng.ring_release(sid)

        # A stream of batches, as the files loop of lib/upload.js is: step s submits its slots, then collects
        # the digests of step s-1 (whose chains ran under this step's copies); the last step's are collected
        # before the clock stops, so every step's H2D and D2H lie inside the timed region.
        def collect(tickets):
            res = [eng.result(t, per_slot, mask, want_ok=exp_host is not None) for t in tickets]
            if dist is not None:  # digests are gathered on the host; no GPU collective
                mine = torch.from_numpy(np.concatenate([r[ALG_SHA1 if algs == "sha1" else ALG_MD5] for r in res]))
                gl = [torch.empty_like(mine) for _ in range(world)] if rank == 0 else None
                dist.gather(mine, gl, dst=0)
            return res

        def e2e_run(nsteps):
            prev, res = None, None
            for _ in range(nsteps):
                tickets = []
                for k in range(slots):
                    sid, _ = eng.ring_acquire(slot_bytes)
                    exp_k = exp_host[k * per_slot:(k + 1) * per_slot] if exp_host is not None else None
                    tickets.append(eng.submit(sid, offs_s, lens_s, mask, exp_k))
                if prev is not None:
                    res = collect(prev)
                prev = tickets
            return collect(prev)

        res = e2e_run(max(1, min(args.warmup, 2)))
        s0 = eng.stats()
        barrier()
        w0 = time.perf_counter()
        e2e_steps = max(1, min(args.steps, 5))
        res = e2e_run(e2e_steps)
        barrier()
        w1 = time.perf_counter()
        s1 = eng.stats()
        dt = (w1 - w0) / e2e_steps
        if dist is not None:
            tm = torch.tensor([dt], dtype=torch.float64)
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            dt = float(tm.item())
        # parity of the e2e path with the device-resident path
        key = ALG_SHA1 if algs == "sha1" else ALG_MD5
        name = "sha1" if algs == "sha1" else "md5"
        got = np.concatenate([r[key] for r in res])
        assert (got == outs[name].cpu().numpy().reshape(n, -1)[:per_slot * slots]).all(), "e2e digests differ"
        if exp_host is not None:
            okv = np.concatenate([r["ok"] for r in res])
            assert sorted(np.nonzero(~okv)[0].tolist()) == [c for c in corrupt if c < per_slot * slots]
        e2e = {"value": round(world * slots * slot_bytes / dt / GB, 3), "unit": "GB/s",
               "h2d_bytes_per_step": int((s1["h2d_bytes"] - s0["h2d_bytes"]) // e2e_steps),
               "d2h_bytes_per_step": int((s1["d2h_bytes"] - s0["d2h_bytes"]) // e2e_steps),
               "messages_per_gpu": slots * per_slot, "steps": e2e_steps, "ms_per_step": round(dt * 1e3, 3), "timing": "host wall clock around %d streamed steps (submit all slots, collect the previous step's digests; last step drained inside), max over ranks" % e2e_steps,
               "pinned_ring_bytes": slots * slot_bytes, "device_ring_bytes": dev_mult * slots * slot_bytes,
               "launches_per_step": int((s1["launches"] - s0["launches"]) // e2e_steps)}
        eng.close()
    sampler.stop()
    clocks = sampler.summary(t0, t1)

    # ---- roofline: integer alu pipe; HBM beside it ---------------------------------------
    ops_per_byte = OPS_PER_BLOCK[algs] / 64.0
    clk = sm_max_mhz
    int_peak = SM_COUNT * ALU_LANES_PER_CLK_SM * clk * 1e6 / ops_per_byte / GB
    obs = clocks.get("sm_mhz") or clk
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        with open(tp) as f:
            traffic = json.load(f).get(args.workload)
    warps = (n + 31) // 32
    w_sat = 2 if algs == "md5" else 1
    roofline = {
        "bound": "alu", "achieved": round(per_gpu, 2), "peak": round(min(int_peak, hbm_peak), 2), "unit": "GB/s",
        "frac": round(per_gpu / min(int_peak, hbm_peak), 4), "traffic": traffic,
        "definition": f"min(HBM, {SM_COUNT} SMs x {ALU_LANES_PER_CLK_SM} int lane-ops/clk x {clk:.0f} MHz / {ops_per_byte:.2f} ops/B); per GPU; of {peak_src} clock/HBM",
        "frac_at_observed_clock": round(per_gpu / (int_peak * obs / clk), 4),
        "hbm": {"achieved": round(per_gpu, 2), "peak": hbm_peak, "frac": round(per_gpu / hbm_peak, 4), "source": peak_src},
        "algorithmic_bytes_per_launch": total_bytes,
        "parallelism_ceiling": round(min(1.0, warps / (592.0 * w_sat)), 4),
        "note": "frac can exceed 1: SURVEY 8d's bound counts every op on the 64-lane/clk/SM alu pipe, and the kernels issue "
                "the plain adds as IMAD on the fma pipe (ncu: alu pipe ~90 % busy at this rate, profiles/r1_*_ncu_summary.txt); "
                "below parallelism_ceiling < 1 the batch is chain-latency-bound, not pipe-bound",
    }

    line = {
        "metric": METRIC, "value": round(value, 3), "unit": "GB/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4), "higher_is_better": True, "scaling": "strong" if args.strong else "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": desc, "algs": algs, "messages_per_gpu": n, "message_bytes": ln,
                   "l2": "inputs (%.1f GiB per GPU) exceed the 126 MB L2; no flush" % (total_bytes / 2**30),
                   "kernel": _plan(algs, ln, n),
                   "verified_vs_oracle": checked, "fused_compare": algs == "sha1"},
        "roofline": roofline, "clocks": clocks, "gpu_launches": launches_per_step * args.steps,
    }
    if e2e is not None:
        line["e2e"] = e2e
    if rank == 0 and world == 1 and not args.skip_cpu:
        threads = os.cpu_count() or 1
        n_sample = max(threads * 8, min(n, (2 << 30) // ln))
        v, per, det = cpu_reference(algs, ln, n_sample, 3, 1, threads)
        v1, _, _ = cpu_reference(algs, ln, max(8, n_sample // 16), 1, 1, 1)
        line["cpu_baseline"] = {"value": round(v, 3), "unit": "GB/s", "cores": threads, "kind": "port",
                                "sample": f"{n_sample} x {ln} B x 3 passes, libcrypto EVP ({det['openssl']}) = the code Node crypto runs; "
                                          f"1 thread (reference semantics): {v1:.3f} GB/s",
                                "value_1thread": round(v1, 3)}
    if rank == 0:
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


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
    if algs == "sha1":
        return _capi.plan(1, ln, ln, n)
    return {"md5": _capi.plan(2, ln, ln, n), "sha256": _capi.plan(4, ln, ln, n)}


if __name__ == "__main__":
    sys.exit(main())
