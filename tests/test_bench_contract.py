"""bench.py's stdout contract on a box without a GPU: exactly one JSON line, everything else on stderr."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_reports_unavailable_strategy_as_one_json_line():
    r = subprocess.run([sys.executable, "bench.py", "--impl", "reference", "--strategy", "zero3"], cwd=ROOT,
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0
    lines = r.stdout.splitlines()
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and "DeepSpeed" in d["unavailable"]


def test_only_the_result_line_reaches_stdout():
    code = ("import bench, os; bench.quiet_stdout(); print('python noise'); os.system('echo child noise'); "
            "bench.emit({'metric': 'tokens_per_sec', 'value': 1.0})")
    r = subprocess.run([sys.executable, "-c", code], cwd=ROOT, capture_output=True, text=True, timeout=60)
    assert r.returncode == 0
    assert r.stdout == '{"metric": "tokens_per_sec", "value": 1.0}\n'
    assert "python noise" in r.stderr and "child noise" in r.stderr
