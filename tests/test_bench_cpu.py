"""CPU-side checks of bench.py's contract: the reference arm (the oracle port on the host cores) runs without a GPU and
prints ONE JSON line with the keys the driver reads; the workload table names BASELINE.json's configurations."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_json_line():
    env = dict(os.environ, OMP_NUM_THREADS="4")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "config2", "--steps", "1",
                        "--warmup", "0"], capture_output=True, text=True, cwd=ROOT, env=env, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["higher_is_better"] is True and d["unit"] == "pixel-pairs/s"
    assert d["value"] > 0 and d["steps"] == 1 and d["n_gpus"] == 1
    assert d["cpu_baseline"]["kind"] in ("port", "reference") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert "config2" in d["config"]["workload"]


def test_workloads_cover_the_baseline_configs():
    sys.path.insert(0, ROOT)
    import bench
    base = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    assert len(base["configs"]) == 5
    for key in ("config2", "config3", "config4", "config5"):
        assert key in bench.WORKLOADS
    c2, c3, c4 = (bench.WORKLOADS[k] for k in ("config2", "config3", "config4"))
    assert (c2["global_batch"], c2["h"], c2["w"], c2["dim"]) == (8, 56, 56, 128)       # batch 8, 448 / 8, proj dim 128
    assert (c3["global_batch"], c3["h"], c3["w"]) == (16, 64, 64)                      # batch 16, 512 / 8
    assert (c4["global_batch"], c4["h"], c4["w"]) == (64, 64, 64)                      # global batch 64, 512 / 8
