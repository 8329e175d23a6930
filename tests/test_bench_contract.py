"""bench.py's stdout contract: exactly ONE JSON line, whatever libraries and child processes print (CPU arm, no GPU needed)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_json_line(tmp_path):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                        "--scale", "0.005"], capture_output=True, text=True, timeout=600, cwd=ROOT,
                       env=dict(os.environ, SVTK_BENCH_CACHE=str(tmp_path)))
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout[:2000]
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "SV records audited/s" and d["unit"] == "records/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 1
    assert d["value"] > 0 and d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    # both arms print the same `config` for the same arguments (the driver compares them)
    sys.path.insert(0, ROOT)
    import bench
    args = type("A", (), dict(workload="c2", scale=0.005))
    assert d["config"] == bench.job_config(args, 1)
