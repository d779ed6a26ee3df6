"""The reference arm of bench.py on the host cores (no GPU): one JSON line in the contract's shape, for the lifter and for
the transformer lifter, and the GPU arm's refusal to run without a B200."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args, timeout=600):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, timeout=timeout,
                       cwd=ROOT)
    return r


@pytest.mark.parametrize("workload,metric", [("train", "lifted poses/sec (fwd+bwd+Adam)"), ("vit_eval", "lifted poses/sec (eval inference)")])
def test_reference_arm_prints_one_contract_line(workload, metric):
    r = _run("--impl", "reference", "--workload", workload, "--steps", "1", "--warmup", "1", "--gpus", "1")
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1, r.stdout[-2000:]
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == metric and d["unit"] == "poses/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["n_gpus"] == 1 and d["vs_baseline"] is None
    assert d["e2e"] == {"value": d["value"], "unit": "poses/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "baselineModel.py")):
        assert cb["kind"] == "reference" and d["reference_files"]["matches_recorded"] is True
    if workload == "train":
        assert d["c1"]["value"] > 0 and "batch 64" in d["c1"]["config"]


def test_gpu_arm_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = _run("--steps", "1", "--warmup", "1", timeout=300)
    assert r.returncode != 0 and not [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
