"""bench.py end to end on a GPU: the JSON line carries every key of the contract, the two arms describe
the same workload, and the counters make sense.  Short runs (a handful of steps).  Needs a B200: -m gpu."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

pytestmark = pytest.mark.gpu


def run_bench(*args):
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                         env=env, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, f"expected exactly one line on stdout, got {len(lines)}"
    return json.loads(lines[0])


def test_b200_arm_line(cuda_device):
    d = run_bench("--steps", "6", "--warmup", "3", "--no-cpu-baseline", "--e2e-steps", "5")
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "roofline", "e2e", "gpu_launches", "clocks"):
        assert key in d, key
    assert d["metric"] == "preprocessed_clips_per_sec" and d["unit"] == "clips/s" and d["higher_is_better"] is True
    assert d["steps"] == 6 and d["warmup"] == 3 and d["n_gpus"] == 1 and d["vs_baseline"] is None
    assert d["gpu_launches"] == 6 and d["config"]["workload"] == "slowfast_train_b8" and "model" not in d["config"]
    assert abs(d["value"] - 8 * 6 / (d["ms_per_step"] * 6 / 1000.0)) < 1e-3 * d["value"]
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    assert r["kernel"] == "pva_tiled_kernel" and r["traffic"] is None or r["traffic"] >= r["algorithmic_bytes_per_launch"]
    e = d["e2e"]
    assert e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] == 8 * 3 * 40 * 256 * 256 * 2
    assert 0 < e["value"] < d["value"] and e["ceiling"]["value"] > 0 and 0 < e["frac_of_ceiling"] < 1.2
    assert set(d["clocks"]) >= {"sm_mhz", "sm_max_mhz", "reasons"}
    assert d["launch_us_isolated"] > 0 and d["host_us_per_call"] > 0


def test_reference_arm_prints_the_same_config(cuda_device):
    ref = run_bench("--impl", "reference", "--steps", "2", "--warmup", "3")
    assert ref["impl"] == "reference" and ref["gpu_launches"] == 0 and ref["value"] > 0
    assert ref["cpu_baseline"]["kind"] == "port" and ref["cpu_baseline"]["cores"] >= 1
    assert ref["e2e"]["h2d_bytes_per_step"] == 0 and ref["e2e"]["d2h_bytes_per_step"] == 0
    import bench

    assert ref["config"] == bench.workload_config("slowfast_train_b8", bench.WORKLOADS["slowfast_train_b8"], 1)
