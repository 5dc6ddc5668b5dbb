"""The parts of bench.py's contract that can be checked without a GPU: the reference arm prints ONE JSON line with the
required keys, runs the unmodified reference functions on the requested config, and marks reduced smoke runs."""
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
REQUIRED = {"impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
            "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches"}


def test_reference_arm_prints_one_json_line(oracle):
    env = dict(os.environ, KD_BENCH_REF_LOG2N="16", KD_BENCH_REF_WORLD="2")
    r = subprocess.run([sys.executable, "bench.py", "--impl", "reference", "--steps", "2", "--warmup", "0"], cwd=ROOT,
                       env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert REQUIRED <= set(d)
    assert d["impl"] == "reference" and d["reduced"] is True          # 2^16 is not the headline size: said so
    assert d["config"]["n_points"] == 1 << 16 and d["config"]["dim"] == 3 and d["config"]["world"] == 2
    assert d["steps"] == 2 and d["value"] > 0 and d["gpu_launches"] == 0
    assert abs(d["value"] - (1 << 16) / (d["ms_per_step"] * 1e-3)) / d["value"] < 1e-6


def test_reference_arm_defaults_to_the_headline_config():
    src = (ROOT / "bench.py").read_text()
    assert 'ap.add_argument("--log2n", type=int, default=28)' in src and 'ap.add_argument("--dim", type=int, default=3)' in src
    assert 'os.environ.get("KD_BENCH_REF_LOG2N", str(args.log2n))' in src      # no reduced default
