"""examples/quickstart.py runs end to end on the CPU oracle."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(backend, extra=()):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "quickstart.py"), "--backend",
                        backend, "--workers", "6", "--dim", "12", "--iters", "300", *extra],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout
    lines = [l for l in r.stdout.splitlines() if "iterations" in l]
    assert len(lines) == 3, r.stdout
    return lines


def test_quickstart_oracle():
    _run("oracle")
