"""bench.py's contract on the CPU side: the reference arm prints one JSON line with the keys the driver reads,
and rank != 0 of a multi-rank launch exits quietly."""
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def run_reference(extra_env):
    env = dict(os.environ, OMNIA_BENCH_ROWS="4000", OMNIA_BENCH_QUERIES="32", OMNIA_BENCH_DIM="128", **extra_env)
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1"],
                         capture_output=True, text=True, env=env, cwd=str(ROOT), timeout=240)
    assert out.returncode == 0, out.stderr[-2000:]
    return out.stdout.strip().splitlines()


def test_reference_arm_prints_the_contract_line():
    lines = run_reference({})
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["higher_is_better"] is True and d["unit"] == "queries/s"
    assert d["metric"].startswith("recall queries/sec") and d["value"] > 0 and d["steps"] == 2 and d["warmup"] == 1
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"] and "model" not in d["config"]
    assert d["vs_baseline"] is None and d["data"] == "synthetic"


def test_reference_arm_runs_on_rank_zero_only():
    assert run_reference({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}) == []
