"""bench.py's one-line JSON contract, checked on the committed lines under profiles/ (CPU only: no run is started here).
The driver reads exactly these keys; the roofline / e2e / clocks objects must be internally consistent."""
import glob
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LINES = sorted(glob.glob(os.path.join(ROOT, "profiles", "r02_bench_sd15_b8_ddim50_v*.json")))


def test_bench_help_runs_without_gpu():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--help"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0
    for flag in ("--gpus", "--steps", "--warmup", "--impl"):
        assert flag in out.stdout


@pytest.mark.parametrize("path", LINES, ids=[os.path.basename(p) for p in LINES])
def test_committed_bench_line_keeps_the_contract(path):
    d = json.load(open(path))
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype",
              "data", "config", "e2e", "gpu_launches", "clocks", "roofline"):
        assert k in d, k
    assert d["metric"] == "images/s" and d["unit"] == "images/s" and d["higher_is_better"] is True
    assert d["dtype"] == "bf16" and d["scaling"] == "weak" and d["vs_baseline"] is None
    cfg = d["config"]
    assert cfg["workload"] == "sd15_512_b8_ddim50_cfg" and "model" in cfg
    # value = images of all ranks / device time of the timed jobs
    images_per_job = cfg["batch_per_gpu"] * d["n_gpus"]
    assert d["value"] == pytest.approx(images_per_job / (d["ms_per_step"] * 1e-3), rel=1e-6)
    e = d["e2e"]
    assert e["unit"] == "images/s" and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0
    assert 0.5 * d["value"] < e["value"] <= d["value"] * 1.02     # host copies inside the timed region: never faster than the device-timed rate
    assert d["gpu_launches"] > 0 and d["launches_per_sampler_step"] > 100
    r = d["roofline"]
    assert r["bound"] == "tensor" and r["unit"] == "TFLOP/s"
    assert r["frac"] == pytest.approx(r["achieved"] / r["peak"], rel=1e-9) and 0.0 < r["frac"] < 1.0
    assert r["achieved"] == pytest.approx(r["algorithmic_flop_per_launch"] / (r["avg_launch_ms"] * 1e-3) / 1e12, rel=1e-6)
    # the dominant kernel's time per step cannot exceed the step it is part of
    assert r["launches"] * r["avg_launch_ms"] < r["step_ms_in_graph"]
    c = d["clocks"]
    assert c["sm_mhz"] <= c["sm_max_mhz"]
    assert not set(c["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    hb = d.get("roofline_hbm")
    if hb:
        for ent in hb.values():
            if ent["frac"] is not None:
                assert 0.0 < ent["frac"] < 1.0 and ent["achieved_gbs"] == pytest.approx(ent["frac"] * ent["peak_gbs"], rel=1e-6)
