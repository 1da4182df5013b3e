"""bench.py contract pieces that can be checked without a GPU: the reference arm prints the contracted JSON line (the CPU
port timed on the host cores), ranks > 0 of the reference arm exit without work, and the product arm FAILS LOUDLY without a
CUDA device (no CPU fallback)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SMALL = ["--batch", "256", "--horizon", "10", "--iters", "2", "--steps", "1", "--warmup", "1", "--cpu-sample", "64"]


def _run(extra, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + extra, capture_output=True, text=True, env=e, timeout=600)


def test_reference_arm_prints_the_contract_line(built):
    r = _run(["--impl", "reference"] + SMALL)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert line["impl"] == "reference" and line["metric"] == "fddp_mpc_solves_per_sec" and line["unit"] == "solves/s"
    assert line["higher_is_better"] is True and line["dtype"] == "f64" and line["data"] == "synthetic" and line["vs_baseline"] is None
    assert line["value"] > 0 and line["steps"] == 1 and line["warmup"] == 1
    assert line["e2e"] == {"value": line["value"], "unit": "solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = line["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == line["value"] and "sample" in cb
    assert "workload" in line["config"] and "model" not in line["config"]


def test_reference_arm_other_ranks_exit_without_work(built):
    r = _run(["--impl", "reference", "--gpus", "2"] + SMALL, env={"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert r.returncode == 0 and not [l for l in r.stdout.splitlines() if l.startswith("{")]


def test_product_arm_fails_loudly_without_a_gpu(built):
    import torch

    if torch.cuda.is_available():
        return   # on a GPU box the product arm runs (covered by the driver's bench run)
    r = _run(["--steps", "1", "--warmup", "1"])
    assert r.returncode != 0 and "no CUDA device" in (r.stderr + r.stdout)
