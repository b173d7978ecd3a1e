"""bench.py's one-line JSON contract, checked on the GPU (the CPU side is covered by tests/test_host_logic_cpu.py)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def test_bench_line_has_every_contract_key():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "2", "--warmup", "3", "--no-cpu-baseline"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.strip().splitlines() if l.startswith("{")]
    assert len(lines) == 1                                   # exactly ONE JSON line
    j = json.loads(lines[0])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
              "dtype", "data", "config", "e2e", "gpu_launches", "clocks", "roofline"):
        assert k in j, k
    assert j["metric"] == "event_windows_per_s_to_visual_tokens" and j["unit"] == "windows/s" and j["n_gpus"] == 1
    assert j["steps"] == 2 and j["warmup"] == 3 and j["higher_is_better"] is True and j["scaling"] == "weak" and j["vs_baseline"] is None
    assert "workload" in j["config"] and "model" not in j["config"]
    assert abs(j["value"] - 32 / (j["ms_per_step"] * 1e-3)) < 1e-6 * j["value"]
    assert j["gpu_launches"] > 100
    e = j["e2e"]
    assert e["value"] > 0 and e["h2d_bytes_per_step"] == 32 * 1_000_000 * 5 + 33 * 8 and e["d2h_bytes_per_step"] == 32 * 582 * 4096 * 2 + 4
    r = j["roofline"]
    for k in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert k in r, k
    assert r["bound"] == "tensor" and r["unit"] == "TFLOP/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9 and 0.3 < r["frac"] < 1.2
    assert abs(r["share_of_step"] - r["avg_launch_ms"] * r["launches"] / j["steps"] / j["ms_per_step"]) < 1e-9
    assert "traffic_source" in r and j["raster_roofline"]["bound"] == "hbm"
    c = j["clocks"]
    assert c["sm_mhz"] and c["sm_max_mhz"] and isinstance(c["reasons"], list)


def test_reference_arm_runs_the_configured_window():
    """`--impl reference`: one window of the configured size per step on the host cores, nothing scaled; the keys that make
    the two arms comparable are in `config`."""
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    j = json.loads([l for l in out.stdout.strip().splitlines() if l.startswith("{")][-1])
    assert j["impl"] == "reference" and j["metric"] == "event_windows_per_s_to_visual_tokens" and j["unit"] == "windows/s"
    assert j["config"]["events_per_window"] == 1_000_000 and j["config"]["sensor"] == [480, 640] and j["config"]["windows_per_step"] == 1
    cb = j["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and abs(cb["value"] - j["value"]) < 1e-9
    assert cb["raster_all_cores"] is None or cb["raster_all_cores"]["processes"] == cb["cores"]
    assert j["e2e"]["h2d_bytes_per_step"] == 0 and j["e2e"]["d2h_bytes_per_step"] == 0
