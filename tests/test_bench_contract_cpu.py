"""The reference arm of bench.py (`--impl reference`: the reference's OWN model code on the host cores through the seam of
oracle/ref_cpu.py, staged under oracle/_ref; the oracle port only where that staging is absent) on a tiny geometry: exits 0,
prints exactly one JSON line, and that line carries the keys the driver reads — with the number of steps ACTUALLY timed and
nothing extrapolated.  The GPU arm needs a B200 and
is exercised by the driver itself; its line is built by the same code from the same `base` dictionary."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_contract_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--geom", "tiny", "--bsz", "4",
                        "--steps", "1", "--warmup", "1"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines
    d = json.loads(lines[0])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
              "dtype", "data", "config", "impl", "cpu_baseline", "e2e", "gpu_launches"):
        assert k in d, k
    assert d["impl"] == "reference" and d["n_gpus"] == 1 and d["steps"] == 1 and d["higher_is_better"] is True
    assert d["unit"] == "tok/s" and d["value"] > 0 and d["vs_baseline"] is None and d["gpu_launches"] == 0
    assert "workload" in d["config"] and "model" not in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert cb["extrapolated"] is False and abs(d["ms_per_step"] - 1e3 * 4 / d["value"]) < 1e-6 * d["ms_per_step"] + 1e-9
    if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "Albatross", "rwkv7.py")):
        assert cb["kind"] == "reference"
    c0 = d["cpu_baseline_configs0"]                      # 0.1B, bsz 1, 16-token prompt, 3 warm-up + 128 greedy tokens
    assert c0["steps_timed"] == 128 and c0["value"] > 0 and "0.1B" in c0["sample"]
    e = d["e2e"]
    assert e["value"] == d["value"] and e["unit"] == d["unit"] and e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0
    base = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    metric = base.get("metric")
    if isinstance(metric, str):
        assert d["metric"] == metric
