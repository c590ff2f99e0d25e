"""bench.py contract pieces that need no GPU."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_reports_unavailable():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference"],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120, cwd=ROOT)
    assert p.returncode == 0
    r = json.loads([l for l in p.stdout.splitlines() if l.startswith("{")][-1])
    assert r["impl"] == "reference" and "unavailable" in r
