"""bench.py's reference arm (`--impl reference`) runs on host cores only, so its contract can be checked here: one JSON line with the
keys the driver reads, the GPU arm's metric / unit / workload, and under torchrun only rank 0 prints."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(env_extra=None):
    env = dict(os.environ)
    env.update(env_extra or {})
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0", "--workload", "tum"],
                       capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    return [l for l in r.stdout.splitlines() if l.strip()]


def test_reference_arm_prints_one_json_line():
    lines = _run()
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "ORB frames/sec (extract+describe)" and d["unit"] == "images/s"
    assert d["higher_is_better"] is True and d["steps"] == 1 and d["value"] > 0 and d["gpu_launches"] == 0
    assert "TUM" in d["config"]["workload"]
    # both arms print the SAME config object (the driver compares them); arm-specific facts live in `run`
    sys.path.insert(0, ROOT)
    import bench
    assert d["config"] == bench.shared_config("tum")
    assert json.loads(json.dumps(bench.shared_config("kitti"))).keys() == d["config"].keys()
    assert "images_per_step" in d["run"] and "rank 0" in d["run"]["ranks"]
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_exit_quietly():
    assert _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}) == []
