"""bench.py contract (driver-facing): one JSON line with the required keys, on CPU with the torch engine and a tiny model; the
reference arm reports `unavailable` and exits 0."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="2")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], cwd=ROOT, env=env, capture_output=True,
                       text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-3000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, r.stdout
    return json.loads(lines[0])


def test_bench_json_line_has_the_contract_keys():
    d = _run("--steps", "3", "--warmup", "3", "--backend", "torch", "--batch", "2", "--model-config",
             "configs/model_config_tiny.yaml")
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches", "clocks"):
        assert k in d, k
    assert d["n_gpus"] == 1 and d["steps"] == 3 and d["warmup"] >= 3 and d["higher_is_better"] is True
    assert d["unit"] == "tokens/s" and d["value"] > 0 and d["scaling"] == "strong"   # the headline arm keeps the reference's GLOBAL batch 8
    assert set(d["e2e"]) >= {"value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"}
    assert d["e2e"]["h2d_bytes_per_step"] == 2 * 2 * 128 * 4 and d["e2e"]["d2h_bytes_per_step"] == 4
    assert d["config"]["global_batch"] == 2 and d["config"]["seq_len"] == 128 and d["config"]["parallelism"] == "dp1"
    assert abs(d["vs_baseline"] - d["value"] / 27887.0) < 1e-6
    assert set(d["clocks"]) >= {"sm_mhz", "sm_max_mhz", "reasons"}


def test_reference_arm_reports_unavailable_and_exits_zero():
    d = _run("--impl", "reference", "--gpus", "1", "--steps", "2", "--warmup", "1")
    assert d["impl"] == "reference" and "unavailable" in d and len(d["unavailable"]) > 20


def test_a_failing_arm_does_not_lose_the_headline():
    """Two gloo ranks, arms dp + tp: rank 1 raises inside the TP arm.  Its peer is stuck in a collective; the per-arm
    watchdog sees the failure marker, rank 0 still prints ONE line with the DP headline and `extra.tp.error`, and the
    launcher exits without waiting for a collective timeout."""
    import socket
    import time
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="2", DTC_BENCH_FAULT="tp:1", DTC_DIST_TIMEOUT="300")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "bench.py"), "--gpus", "2", "--steps", "2", "--warmup", "1",
           "--backend", "torch", "--batch", "2", "--model-config", "configs/model_config_tiny.yaml", "--arms", "dp,tp"]
    t0 = time.time()
    r = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, r.stdout[-2000:] + r.stderr[-3000:]
    d = json.loads(lines[0])
    assert d["value"] and d["value"] > 0 and d["config"]["parallelism"] == "dp2"
    assert "error" in d["extra"]["tp"]
    assert time.time() - t0 < 250
