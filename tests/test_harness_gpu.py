"""The public CLI on a GPU: native engine, result schema, device timer, loss going down."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _cli(tmp_path, *extra):
    cmd = [sys.executable, os.path.join(ROOT, "benchmarking", "train_harness.py"), "--strategy", "ddp",
           "--world-size", "1", "--rank", "0", "--tier", "T", "--seq-len", "128", "--steps", "24",
           "--warmup-steps", "4", "--per-device-batch", "2", "--grad-accum", "1", "--results-dir", str(tmp_path),
           "--log-interval", "0"] + list(extra)
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return r.stdout, json.load(open(tmp_path / "result_ddp_ws1_seq128_tierT.json"))


def test_cli_native_path(tmp_path):
    out, res = _cli(tmp_path, "--profile")
    assert res["impl"] == "NativeEngine" and res["timer"] == "device" and res["dtype"] == "bfloat16"
    assert res["tokens_per_sec"] > 0 and 0 < res["mean_loss"] < 7.0
    assert res["gpu_launches_per_step"] > 20
    assert res["profile"]["forward_ms"] > 0 and res["profile"]["backward_ms"] > 0
    assert "BENCHMARK_RESULT_JSON_START" in out and "BENCHMARK_RESULT_JSON_END" in out


def test_cli_torch_path_agrees_on_loss_scale(tmp_path):
    _, nat = _cli(tmp_path / "n", "--dropout", "0")
    _, ref = _cli(tmp_path / "t", "--dropout", "0", "--impl", "torch", "--dtype", "fp32")
    assert ref["impl"] == "TorchEngine"
    # same data order, same init distribution (not the same RNG stream): mean losses land close together
    assert abs(nat["mean_loss"] - ref["mean_loss"]) < 0.25
