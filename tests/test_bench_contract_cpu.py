"""bench.py contract checks that need no GPU: the reference arm prints one JSON line with the agreed keys,
and under a 2-rank launch only rank 0 prints."""
import json
import os
import socket
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = {"impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
        "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"}


def test_reference_arm_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert KEYS <= set(d), KEYS - set(d)
    assert d["impl"] == "reference" and d["unit"] == "GB/s" and d["value"] > 0 and d["vs_baseline"] is None
    assert d["cpu_baseline"]["kind"] in ("port", "reference") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"] == {"value": d["value"], "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"] and "model" not in d["config"]


def test_reference_arm_under_torchrun_prints_once():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0",
           "--workload", "md5_sha256_parts"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1 and json.loads(lines[0])["n_gpus"] == 2


def test_device_order_takes_gpus_from_both_halves_of_the_box():
    """bench.py --device-order auto: rank r's GPU alternates between the two halves of the device list (the two host
    bridges of this pool's 8-GPU boxes, profiles/r2_h2d_probe_8gpu.jsonl); every rank still gets its own device."""
    sys.path.insert(0, ROOT)
    import bench
    order, why = bench.device_order("auto", 8, 4)
    assert order[:4] == [0, 4, 1, 5] and "halves" in why
    assert sorted(bench.device_order("auto", 8, 8)[0]) == list(range(8))
    assert bench.device_order("auto", 8, 2)[0][:2] == [0, 4]
    assert bench.device_order("auto", 1, 1)[0] == [0] and bench.device_order("auto", 2, 2)[0] == [0, 1]
    assert bench.device_order("identity", 8, 4)[0][:4] == [0, 1, 2, 3]
    assert bench.device_order("3,2,1,0", 8, 4) == ([3, 2, 1, 0], "given on the command line")
