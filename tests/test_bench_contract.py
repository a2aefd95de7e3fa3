"""The reference arm of bench.py runs on CPU (libsodium on the host cores), so its side of the driver's contract can be
checked without a GPU: one JSON line, the keys the driver reads, the same `config` the GPU arm prints, and under a
multi-rank launch only rank 0 works."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(extra_env=None, args=()):
    env = dict(os.environ)
    env.update(extra_env or {})
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "3", "--warmup", "1",
                          *args], capture_output=True, text=True, env=env, cwd=ROOT, timeout=300)
    assert res.returncode == 0, res.stderr[-2000:]
    return [l for l in res.stdout.splitlines() if l.strip()]


def test_reference_arm_prints_the_contract_line():
    lines = _run()
    assert len(lines) == 1, lines
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "frames/s" and d["higher_is_better"] is True
    assert d["n_gpus"] == 1 and d["warmup"] >= 1 and 1 <= d["steps"] <= 3 and d["value"] > 0 and d["ms_per_step"] > 0
    assert d["scaling"] == "weak" and d["vs_baseline"] is None and d["data"] == "synthetic" and d["gpu_launches"] == 0
    assert d["metric"] == "sealed_frames_per_sec" or "frames" in d["metric"]
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    e = d["e2e"]
    assert e["value"] == d["value"] and e["unit"] == d["unit"] and e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0
    # the workload description both arms print comes from one function
    sys.path.insert(0, ROOT)
    import bench

    assert d["config"] == bench.make_config("c2", 1, 3, False, bench.plan_mode_for("c2"))
    assert "workload" in d["config"] and "model" not in d["config"]


def test_reference_arm_other_ranks_exit_without_work():
    assert _run({"RANK": "1", "LOCAL_RANK": "1", "WORLD_SIZE": "2"}, ("--gpus", "2")) == []
    (line,) = _run({"RANK": "0", "LOCAL_RANK": "0", "WORLD_SIZE": "2"}, ("--gpus", "2"))
    d = json.loads(line)
    assert d["n_gpus"] == 2 and d["config"]["n_gpus"] == 2 and d["impl"] == "reference"
