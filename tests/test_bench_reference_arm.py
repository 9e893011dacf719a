"""bench.py --impl reference (the reference's CPU path timed on the host cores) prints ONE JSON line with the keys the
driver reads, takes its config from the same function as our arm, and loads no library of this repo into the process
that is timed (the scene comes from a child process)."""
import json
import os
import subprocess
import sys

from parity import ROOT


def test_reference_arm_line(built_libs):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "c1",
                          "--steps", "2", "--warmup", "1"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["metric"] == "Mrays/s" and j["unit"] == "Mrays/s" and j["higher_is_better"] is True
    assert j["value"] > 0 and j["steps"] == 2 and j["warmup"] == 1 and j["gpu_launches"] == 0
    assert j["e2e"] == {"value": j["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = j["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == j["value"] and "sample" in cb
    sys.path.insert(0, ROOT)
    import bench
    assert j["config"] == bench.bench_config("c1", 1)        # the same dict as our arm's line
    # the timed workers hold the checker (oracle/), never the product libraries
    libs = j["libraries_loaded_by_the_workers"]
    assert all("opencl_ray_tracer_b200" not in p for p in libs)
    assert any("oracle" in p for p in libs)
