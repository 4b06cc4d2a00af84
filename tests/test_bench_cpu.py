"""bench.py contract on CPU: the reference arm's JSON line, its behaviour under a multi-rank
launch, and that the product arm fails loudly without a GPU (no CPU fallback)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BENCH = os.path.join(ROOT, "bench.py")


def _run(args, env=None, timeout=600):
    e = dict(os.environ)
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
        e.pop(k, None)
    e.update(env or {})
    return subprocess.run([sys.executable, BENCH] + args, cwd=ROOT, env=e, timeout=timeout,
                          stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)


@pytest.fixture(scope="module")
def reference_line():
    r = _run(["--impl", "reference", "--steps", "1", "--warmup", "0"])
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, "stdout must carry exactly one JSON line"
    return json.loads(lines[0])


def test_reference_arm_line_has_the_contract_keys(reference_line):
    d = reference_line
    assert d["impl"] == "reference"
    assert d["metric"] == "nbody_fp64_interactions_per_s" and d["unit"] == "interactions/s"
    assert d["higher_is_better"] is True and d["dtype"] == "f64" and d["vs_baseline"] is None
    assert d["steps"] == 1 and d["warmup"] == 0 and d["n_gpus"] == 1
    assert d["value"] > 0 and d["ms_per_step"] > 0
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == d["value"]
    assert "bodies" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["gpu_launches"] == 0


def test_reference_arm_names_the_same_workload_as_the_product_arm(reference_line):
    sys.path.insert(0, ROOT)
    import bench
    # same workload string, followed by the CPU arm's bounded-sample caveat
    assert reference_line["config"]["workload"].startswith(bench.workload(bench.N_BODIES) + " [CPU arm:")
    assert "N=1048576" in reference_line["config"]["workload"]
    assert "bounded sample" in reference_line["config"]["workload"]


def test_reference_arm_value_is_the_sample_over_the_step_time(reference_line):
    d = reference_line
    n = int(d["cpu_baseline"]["sample"].split()[1])
    assert n % 1024 == 0 and 8192 <= n <= 1 << 20
    assert d["value"] == pytest.approx(float(n) * n / (d["ms_per_step"] * 1e-3), rel=1e-9)


def test_reference_arm_other_ranks_exit_quietly():
    r = _run(["--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
             env={"RANK": "1", "LOCAL_RANK": "1", "WORLD_SIZE": "2"}, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_product_arm_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = _run(["--steps", "1", "--warmup", "0", "--skip-conv", "--skip-cpu"], timeout=300)
    assert r.returncode != 0
    assert r.stdout.strip() == "", "no JSON line may be printed when the CUDA path cannot run"
