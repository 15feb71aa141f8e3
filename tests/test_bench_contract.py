"""bench.py's contract on a box without a GPU: the reference arm (the CPU port of the path, the one leg that may execute
oracle/) prints the agreed JSON line, and the product arm refuses to run -- there is no CPU fallback."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, timeout):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, cwd=ROOT, capture_output=True, text=True, timeout=timeout)


@pytest.mark.timeout(600)
def test_reference_arm_prints_the_contract_line():
    r = _run(["--impl", "reference", "--steps", "1", "--warmup", "0"], 500)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference"
    assert line["metric"].startswith("tracked frames/s") and line["unit"] == "frames/s" and line["higher_is_better"] is True
    assert line["n_gpus"] == 1 and line["steps"] == 1 and line["warmup"] == 0 and line["ms_per_step"] > 0
    assert line["value"] > 1.0  # frames/s of the CPU port on this box's cores
    assert "configs[1]" in line["config"]["workload"]
    cb = line["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["sample"] and cb["value"] == line["value"]
    e = line["e2e"]
    assert e["value"] == line["value"] and e["unit"] == line["unit"] and e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0


def test_product_arm_has_no_cpu_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present: the product arm runs")
    r = _run(["--steps", "1", "--warmup", "1"], 300)
    assert r.returncode != 0
    assert "no CPU fallback" in r.stderr
    assert not any(l.startswith("{") for l in r.stdout.splitlines())  # no result line of any kind
