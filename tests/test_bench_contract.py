"""bench.py prints exactly one JSON line with the contract's keys (reference arm here on CPU, our arm on the GPU)."""
import json
import os
import subprocess
import sys

import pytest

from conftest import ROOT

BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data",
             "config", "e2e"}


def run(args, timeout):
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True, timeout=timeout, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, f"stdout must hold one JSON line, got {len(lines)}"
    return json.loads(lines[0])


def test_reference_arm_line():
    j = run(["--impl", "reference", "--steps", "2", "--warmup", "1"], 600)
    assert BASE_KEYS <= set(j) and j["impl"] == "reference"
    assert j["metric"] == "odometry_frames_per_s" and j["unit"] == "frames/s" and j["higher_is_better"] is True
    assert j["cpu_baseline"]["kind"] == "port" and j["cpu_baseline"]["cores"] >= 1 and j["cpu_baseline"]["value"] == j["value"] > 0
    assert j["e2e"] == {"value": j["value"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in j["config"]


@pytest.mark.gpu
def test_our_arm_line():
    j = run(["--steps", "16", "--warmup", "3", "--no-cpu-baseline"], 900)
    assert BASE_KEYS | {"roofline", "gpu_launches", "clocks"} <= set(j)
    assert j["n_gpus"] == 1 and j["steps"] == 16 and j["warmup"] == 3 and j["vs_baseline"] is None and j["dtype"] == "f32"
    r = j["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    assert r["traffic"] is None or r["traffic"] > 0          # from the committed ncu summary of this workload, null without one
    assert 0 < r["frac_step_anchored"] < 1
    assert set(j["config"]) == {"workload", "width", "height", "icpWeight", "iterations", "so3", "parallelism"}   # same keys in both arms
    lat = j["latency"]
    assert lat["value"] > 0 and lat["host_latency_ms"]["p50"] > 0 and lat["roofline"]["concurrent_launches"] == 1
    e = j["e2e"]
    assert e["value"] > 0 and e["h2d_bytes_per_step"] > 15e6 and e["d2h_bytes_per_step"] > 0 and e["value"] != j["value"]
    assert j["gpu_launches"] == 6 * 16 + 1 and j["result_ring"]["records_received"] == [16]
    assert True and j["value"] > 0 and "workload" in j["config"]
    assert {"sm_mhz", "sm_max_mhz", "reasons"} <= set(j["clocks"])
    assert j["last_frame_abs_translation_error_m"] < 0.05 and e["last_frame_abs_translation_error_m"] < 0.05   # the tracker pulls the 3 cm prior onto the ground truth
    fe = j["e2e_frontend"]   # sensor frame uploaded, model view predicted on the device: every frame tracked, pose on the ground truth
    assert fe["value"] > 0 and fe["frames_tracked"] == fe["steps"] and 2e6 < fe["h2d_bytes_per_step"] < 4e6 and fe["last_frame_abs_translation_error_m"] < 0.03
