"""CPU tests of bench.py's contract: the reference arm (`--impl reference`: the CPU restatement timed on the host cores, the
one place besides the parity check where bench.py may run the oracle) on a small domain, and the keys of the last committed
GPU bench line (profiles/), which the round's driver parses."""
import glob
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
             "dtype", "data", "config"}


def test_reference_arm_prints_one_line_with_the_contract_keys():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--n", "32", "--steps", "2", "--warmup", "1"],
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert BASE_KEYS <= set(d) and d["impl"] == "reference"
    assert d["unit"] == "GDOF/s" and d["higher_is_better"] is True and d["dtype"] == "f64" and d["value"] > 0
    assert d["steps"] == 2 and d["warmup"] == 1 and "32x32x32" in d["config"]["workload"]        # what was run is what is printed
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "32x32x32" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_last_committed_gpu_line_carries_the_contract_keys():
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r02z*_bench_512_default.json")))
    assert files, "no default bench line of the round's last passes under profiles/"
    d = json.loads(open(files[-1]).read().strip().splitlines()[0])
    assert BASE_KEYS <= set(d)
    assert "FAS V-cycle" in d["metric"] and d["unit"] == "GDOF/s"            # BASELINE.json: "3D Poisson FAS V-cycle GDOF/s ..."
    assert d["n_gpus"] == 1 and d["dtype"] == "f64" and "512x512x512" in d["config"]["workload"] and "model" not in d["config"]
    rf = d["roofline"]
    assert {"bound", "achieved", "peak", "unit", "frac", "traffic"} <= set(rf) and rf["bound"] == "hbm" and rf["unit"] == "GB/s"
    assert abs(rf["frac"] - rf["achieved"] / rf["peak"]) < 1e-9 and 0 < rf["frac"] < 1
    assert {"value", "unit", "cores", "kind", "sample"} <= set(d["cpu_baseline"])
    assert {"value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"} <= set(d["e2e"]) and d["e2e"]["h2d_bytes_per_step"] > 0
    assert d["gpu_launches"] > 0 and {"sm_mhz", "sm_max_mhz", "reasons"} <= set(d["clocks"])
    assert d["parity_check"]["bitwise"] is True and d["general_launches_per_step"] == 0
