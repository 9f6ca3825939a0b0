"""bench.py prints the driver's JSON contract (own arm), on a reduced shape so the test stays short."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_bench_line_carries_the_contract_keys():
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--P", "20000", "--H", "256", "--W", "256",
           "--steps", "4", "--warmup", "3", "--no-cpu-baseline"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches", "clocks", "roofline"):
        assert k in line, k
    assert line["n_gpus"] == 1 and line["steps"] == 4 and line["warmup"] == 3
    assert line["value"] > 0 and line["e2e"]["value"] > 0
    # end to end means host inputs: it cannot beat the resident number, and it moved real bytes
    assert line["e2e"]["value"] <= line["value"] * 1.02
    assert line["e2e"]["h2d_bytes_per_step"] == 2 * 20000 * 43 * 4 + 2 * 32 * 4
    assert line["e2e"]["d2h_bytes_per_step"] > 0
    # the eight library kernels of a step, every step: preprocess, tile scan, scatter, tile sort, the two compositors,
    # the loss reduction and the backward preprocess (the packing kernel went when the backward began to sum the views)
    assert line["gpu_launches"] >= 4 * 8
    r = line["roofline"]
    assert r["bound"] == "hbm" and 0 < r["frac"] < 1 and r["achieved"] > 0 and r["peak"] > 0
    assert "workload" in line["config"] and "model" not in line["config"]
