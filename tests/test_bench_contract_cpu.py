"""The parts of the bench.py contract that can be checked without a GPU: the reference arm answers with one JSON
line and exit code 0 (the reference is Python-2 / TensorFlow-0.x source without packaging: DESIGN.md section 4),
also when launched as several ranks; the argument surface the driver uses is accepted."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, env=None):
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True,
                       env=dict(os.environ, **(env or {})), timeout=300)
    return p.returncode, [l for l in p.stdout.splitlines() if l.startswith("{")]


def test_reference_arm_reports_unavailable():
    rc, lines = _run(["--impl", "reference", "--gpus", "1", "--steps", "5", "--warmup", "3"])
    assert rc == 0 and len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and isinstance(d["unavailable"], str) and "\n" not in d["unavailable"]


def test_reference_arm_prints_once_per_job():
    # under torchrun every rank runs bench.py; only rank 0 answers
    rc0, l0 = _run(["--impl", "reference", "--gpus", "2"], {"RANK": "0", "WORLD_SIZE": "2", "LOCAL_RANK": "0"})
    rc1, l1 = _run(["--impl", "reference", "--gpus", "2"], {"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert rc0 == 0 and rc1 == 0 and len(l0) == 1 and len(l1) == 0


def test_help_lists_the_driver_flags():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--help"], capture_output=True, text=True,
                       timeout=120)
    assert p.returncode == 0
    for flag in ("--gpus", "--steps", "--warmup", "--impl"):
        assert flag in p.stdout
