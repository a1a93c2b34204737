"""bench.py's reference arm on a small n (CPU only): the JSON line carries the keys the driver
reads, both CPU strategies are reported, and ranks other than 0 print nothing."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(extra_env=None):
    env = dict(os.environ)
    env.update(extra_env or {})
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--n", "2048",
                        "--steps", "2", "--warmup", "1"], capture_output=True, text=True, env=env, timeout=300)
    assert p.returncode == 0, p.stderr
    return p.stdout.strip()


def test_reference_arm_line():
    out = run()
    line = json.loads(out.splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "Tbit-op/s" and line["higher_is_better"] is True
    assert line["steps"] == 2 and line["warmup"] == 1 and line["n_gpus"] == 1 and line["gpu_launches"] == 0
    assert line["config"]["workload"].startswith("square mzd_mul n=2048")
    cb = line["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == line["value"] > 0
    assert set(cb["strategies"]) == {"strassen", "m4rm"}
    assert line["value"] == max(s["value"] for s in cb["strategies"].values())
    e2e = line["e2e"]
    assert e2e["value"] == line["value"] and e2e["h2d_bytes_per_step"] == 0 and e2e["d2h_bytes_per_step"] == 0
    assert line["vs_baseline"] is None  # BASELINE.md holds no published number for this metric


def test_reference_arm_other_ranks_are_silent():
    assert run({"RANK": "1", "LOCAL_RANK": "1", "WORLD_SIZE": "2"}) == ""
