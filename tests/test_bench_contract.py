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


def test_reference_arm_ignores_torchrun_thread_cap():
    """torchrun exports OMP_NUM_THREADS=1: the arm must still use every core it may run on and
    the same sample (round 1's N > 1 reference lines were one-thread runs on a smaller cube)."""
    import ast
    free = json.loads(run().splitlines()[-1])
    capped = json.loads(run({"OMP_NUM_THREADS": "1", "RANK": "0", "LOCAL_RANK": "0", "WORLD_SIZE": "2"}).splitlines()[-1])
    cores = len(os.sched_getaffinity(0))
    assert free["cpu_baseline"]["cores"] == capped["cpu_baseline"]["cores"] == cores
    # `config` carries what the product arm's carries and nothing else (same_config)
    assert set(capped["config"]) == set(free["config"]) == {"workload", "m", "k", "n", "algo", "l2"}
    for k in ("strassen", "m4rm"):
        assert free["cpu_baseline"]["strategies"][k]["sample"] == capped["cpu_baseline"]["strategies"][k]["sample"]
    if cores > 1:  # a one-thread run would be several times slower
        assert capped["value"] > 0.4 * free["value"]
