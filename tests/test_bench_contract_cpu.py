"""bench.py contract on CPU: the reference arm (the reference's own code from oracle/_ref on the host cores, else the
oracle port) prints ONE JSON line with the keys the driver reads; the product arm refuses to run without a GPU (no
CPU fallback)."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_json_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "0", "--size", "8", "--T", "4"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "frame-steps/s" and d["unit"] == "frame-steps/s"
    assert d["higher_is_better"] is True and d["value"] > 0
    staged = os.path.exists(os.path.join(ROOT, "oracle", "_ref", "seva", "sampling.py")) or os.path.isdir("/root/reference/seva")
    assert d["cpu_baseline"]["kind"] == ("reference" if staged else "port")
    assert d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["sample"] and d["steps"] == 1 and d["warmup"] == 0
    assert "T=4 " in d["config"]["workload"] and "CFG batch 8" in d["config"]["workload"]
    assert d["e2e"] == {"value": d["value"], "unit": "frame-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


@pytest.mark.skipif(torch.cuda.is_available(), reason="CPU-only behaviour")
def test_product_arm_fails_loudly_without_gpu():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode != 0 and "no CPU fallback" in (out.stderr + out.stdout)
