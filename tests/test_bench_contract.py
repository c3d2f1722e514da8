"""bench.py's reference arm runs on host cores only: its JSON line (the keys the driver reads) is checked here on a
tiny workload; the GPU arm's line has the same contract keys plus `roofline`, `train`, `clocks`, `gpu_launches`."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_contract_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                          "--num-gaussians", "20000", "--res", "64", "--cpu-pixels", "64"],
                         capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert k in d, k
    assert d["impl"] == "reference" and d["unit"] == "Mrays/s" and d["higher_is_better"] is True and d["value"] > 0
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0 and d["e2e"]["value"] == d["value"]
    assert "workload" in d["config"] and d["vs_baseline"] is None


def test_config_table_names_baseline_shapes():
    sys.path.insert(0, ROOT)
    import bench

    assert set(bench.CONFIGS) == {"c2", "c3", "c4"}
    assert bench.CONFIGS["c2"]["metric"] == bench.METRIC and bench.CONFIGS["c2"]["n"] == 1_000_000
    assert bench.CONFIGS["c3"]["bounces"] == 3 and not bench.CONFIGS["c3"]["grads"] and (bench.CONFIGS["c3"]["w"], bench.CONFIGS["c3"]["h"]) == (1920, 1080)
    assert bench.CONFIGS["c4"]["n"] == 3_000_000 and len(bench.kernel_sources_sha256()) == 64
