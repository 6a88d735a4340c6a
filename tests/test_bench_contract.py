"""bench.py contract checks that do not need a GPU: the reference arm answers with one JSON line
and exit code 0, and the own arm's emitted keys (parsed from the source) cover the driver's list."""
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_reports_unavailable_and_exits_zero():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "1",
                        "--steps", "2", "--warmup", "1"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-500:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    rec = json.loads(lines[0])
    assert rec["impl"] == "reference" and isinstance(rec["unavailable"], str) and rec["unavailable"]


def test_own_arm_emits_every_contract_key():
    src = open(os.path.join(ROOT, "bench.py")).read()
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
                "scaling", "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches",
                "h2d_bytes_per_step", "d2h_bytes_per_step", "global_batch", "seq_len", "parallelism"):
        assert re.search(r"[\"']%s[\"']" % key, src), key
    # warm-up floor of the timing rules
    assert re.search(r"max\(\s*3\s*,|warmup.*3", src)


def test_graft_entry_has_build_and_smoke():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as ge
    assert callable(ge.build) and callable(ge.smoke)
