"""bench.py --impl reference: the CPU arm prints one JSON line and times the reference's own shader code
(oracle/_ref, kind "reference") where that library is available, the oracle port (kind "port") otherwise."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(*extra):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "C1", "--steps", "1",
                          "--warmup", "1", "--ref-frames", "1", *extra], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    return json.loads(lines[0])


def test_reference_arm_line():
    from oracle.ref import binding as R
    d = run()
    assert d["impl"] == "reference" and d["unit"] == "Mrays/s" and d["value"] > 0 and d["higher_is_better"] is True
    assert d["cpu_baseline"]["kind"] == ("reference" if R.available() else "port")
    assert d["cpu_baseline"]["cores"] == os.cpu_count() and d["e2e"]["value"] == d["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    p = run("--ref-port")
    assert p["cpu_baseline"]["kind"] == "port"
    # both arms trace the same rays of the same sample (640x480, every tile, one frame)
    assert "307200 rays/frame" in d["cpu_baseline"]["sample"] and "307200 rays/frame" in p["cpu_baseline"]["sample"]
