"""bench.py's reference arm (the oracle on the host cores) prints the contract's JSON line; runs without a GPU.

The GPU arm's line is checked by the driver on a B200; here the keys both arms share are pinned so that the two lines stay
comparable (same metric, unit, config.workload prefix, higher_is_better)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(extra_env=None, args=()):
    env = dict(os.environ)
    env.update(extra_env or {})
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0", *args],
                       capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, p.stdout
    return json.loads(lines[0])


def test_reference_arm_line_has_the_contract_keys():
    j = _run()
    assert j["impl"] == "reference" and j["metric"] == "nrc_queries_per_s" and j["unit"] == "queries/s"
    assert j["higher_is_better"] is True and j["n_gpus"] == 1 and j["steps"] == 1 and j["gpu_launches"] == 0
    assert j["value"] > 0 and abs(j["ms_per_step"] * 1e-3 * j["value"] - j["cpu_baseline"]["queries"]) < 1e-3 * j["cpu_baseline"]["queries"]
    cb = j["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == j["value"] and "rows" in cb["sample"]
    assert j["e2e"] == {"value": j["value"], "unit": j["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert j["config"]["workload"].startswith("cloud_sixteenth") and "1920x1080" in j["config"]["workload"]
    assert j["vs_baseline"] is None and j["data"] == "synthetic"


def test_reference_arm_other_ranks_exit_without_work():
    """Under torchrun only rank 0 runs the CPU arm; the other ranks print nothing and exit 0."""
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT="29655")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=120, env=env, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    assert not [l for l in p.stdout.splitlines() if l.startswith("{")]
