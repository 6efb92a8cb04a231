"""bench.py's output contract: stdout is exactly ONE JSON line with the keys the driver reads.  The reference arm runs on
the host cores (CPU test); the olbank arm needs a B200."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype",
             "data", "config", "e2e", "cpu_baseline"}


def run(args):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True, cwd=ROOT, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines  # build chatter belongs on stderr
    return json.loads(lines[0])


def test_reference_arm_prints_one_json_line():
    d = run(["--impl", "reference", "--steps", "1", "--warmup", "1"])
    assert BASE_KEYS <= set(d)
    assert d["impl"] == "reference" and d["n_gpus"] == 1 and d["steps"] == 1
    assert d["metric"].startswith("channel-samples/s") and d["unit"] == "channel-samples/s" and d["higher_is_better"] is True
    assert d["cpu_baseline"]["kind"] in ("port", "reference") and d["cpu_baseline"]["cores"] >= 1
    assert d["cpu_baseline"]["value"] == d["value"] == d["e2e"]["value"] > 0
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert "workload" in d["config"] and "model" not in d["config"]


@pytest.mark.gpu
def test_olbank_arm_prints_one_json_line():
    d = run(["--steps", "3", "--warmup", "3"])
    assert BASE_KEYS | {"roofline", "clocks", "gpu_launches", "kernel", "configs1"} <= set(d)
    assert d["gpu_launches"] == 3 and d["kernel"] == "k_bq_tma" and d["dtype"] == "f32" and d["scaling"] == "weak"
    rf = d["roofline"]
    assert rf["bound"] == "hbm" and rf["unit"] == "GB/s" and abs(rf["frac"] - rf["achieved"] / rf["peak"]) < 1e-9
    assert rf["algorithmic_bytes_per_launch"] == 8 * 65536 * 24000
    assert d["e2e"]["h2d_bytes_per_step"] == d["e2e"]["d2h_bytes_per_step"] > 0 and 0 < d["e2e"]["value"] < d["value"]
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["value"] > 0
    assert {"sm_mhz", "sm_max_mhz", "reasons"} <= set(d["clocks"])
    assert d["configs1"]["kernel"] == "k_bq_skew" and d["configs1"]["value"] > 0
