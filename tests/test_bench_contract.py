"""bench.py contract (driver side): the reference arm prints one JSON line with the agreed keys on CPU, non-zero ranks stay
silent, and the GPU arm's line carries roofline / cpu_baseline / e2e / clocks / gpu_launches."""
import json
import os
import subprocess
import sys

import pytest

from conftest import ROOT


def _run(args, env=None, timeout=600):
    e = dict(os.environ)
    e.update(env or {})
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True, env=e, timeout=timeout, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    return [l for l in p.stdout.splitlines() if l.startswith("{")]


def test_reference_arm_line():
    lines = _run(["--impl", "reference", "--steps", "1", "--warmup", "0"])
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "env-steps/s" and d["unit"] == "env-steps/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["steps"] == 1 and d["dtype"] == "f64" and d["vs_baseline"] is None
    assert d["cpu_baseline"]["kind"] in ("port", "mujoco") and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    # the workload string names the CPU arm's REAL sample (worlds, steps, threads, which implementation), not the GPU batch
    assert "CPU arm" in d["config"]["workload"] and "continuing rollout" in d["config"]["workload"] and d["cpu_baseline"]["sample"] in d["config"]["workload"]
    assert d["parity"].startswith("unpinned")


def test_reference_arm_other_ranks_are_silent():
    assert _run(["--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"], env={"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}) == []


@pytest.mark.gpu
def test_gpu_arm_line():
    lines = _run(["--steps", "20", "--warmup", "3"])
    assert len(lines) == 1
    d = json.loads(lines[0])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype",
              "data", "config", "clocks", "gpu_launches", "e2e", "roofline", "cpu_baseline"):
        assert k in d, k
    assert d["n_gpus"] == 1 and d["steps"] == 20 and d["scaling"] == "weak" and d["dtype"] == "f32" and d["data"] == "synthetic"
    # per step: kin, collide, collideMpr, collideCompact, makecon, tendon, tenout, inertia, velact, posout, conout, newton + 2 empty ordering launches
    assert d["gpu_launches"] == 280 and d["config"]["nworld_per_gpu"] == 8192 and d["config"]["fields"] == "all"
    assert d["parity"].startswith("unpinned")
    ss = d["steady_state"]
    assert ss["after_steps"] >= 300 and 0 < ss["value"] <= 1.05 * d["value"] and ss["ncon_mean"] > 5, "steady state = the clenched hand"
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9 and r["traffic"] > 0
    assert abs(r["algorithmic_bytes_per_launch"] - d["config"]["bytes_per_world_step"] * 8192) < 1
    e = d["e2e"]
    assert 0 < e["value"] < d["value"] and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0
    assert d["cpu_baseline"]["kind"] in ("port", "mujoco") and d["cpu_baseline"]["value"] > 0
