"""Result schema, metric formulas, marker round-trip, efficiency formula vs the published rows."""
import argparse
import json
import subprocess

import pandas as pd
import pytest

from b200bench.analysis.parse_metrics import add_scaling_efficiency
from b200bench.utils import result as R


def _args(**kw):
    base = dict(strategy="ddp", world_size=2, rank=0, seq_len=2048, tier="A", steps=100, per_device_batch=1,
                grad_accum=4)
    base.update(kw)
    return argparse.Namespace(**base)


def test_metric_formulas_match_published_row():
    # results/example_output/README.md:8-22 of the reference: DDP 2 GPU
    m = R.compute_metrics([0.489398612] * 3, [6.133937] * 3, 1, 2048, 2, int(13.96808448e9))
    assert m["tokens_per_sec"] == pytest.approx(8369.455699, rel=1e-6)
    assert m["h2d_gbps_per_gpu"] == pytest.approx(1.6739e-05, rel=1e-3)
    assert m["peak_vram_gb"] == pytest.approx(13.96808448)


def test_result_key_order_and_filename(tmp_path):
    m = R.compute_metrics([0.5], [1.0], 1, 2048, 2, 0)
    res = R.make_result(_args(), m, {"impl": "x"})
    assert list(res.keys())[:13] == R.RESULT_KEYS
    p = R.write_result(res, str(tmp_path))
    assert p.endswith("result_ddp_ws2_seq2048_tierA.json")
    assert json.load(open(p))["tokens_per_sec"] == res["tokens_per_sec"]


def test_marker_round_trip_python_and_sed(tmp_path):
    res = R.make_result(_args(), R.compute_metrics([0.5], [1.0], 1, 2048, 2, 0))
    log = "noise\n" + "=" * 80 + "\n" + R.format_marked(res) + "\n" + "=" * 80 + "\ntrailer\n"
    assert R.extract_marked(log) == res
    # the reference collector's extraction (scripts/collect_results.sh:50-52)
    f = tmp_path / "job.log"
    f.write_text(log)
    out = subprocess.run(
        f"sed -n '/BENCHMARK_RESULT_JSON_START/,/BENCHMARK_RESULT_JSON_END/p' {f} | sed '1d;$d'",
        shell=True, capture_output=True, text=True).stdout
    assert json.loads(out) == res


def test_efficiency_formula_reproduces_published_csv():
    # README.md:216-217 of the reference: DDP 2 GPU 8369.4557 -> 50 %, 4 GPU 12220.3415 -> 36.5028 %
    df = pd.DataFrame([
        dict(strategy="ddp", world_size=2, seq_len=2048, tier="A", tokens_per_sec=8369.455699),
        dict(strategy="ddp", world_size=4, seq_len=2048, tier="A", tokens_per_sec=12220.341464),
        dict(strategy="fsdp", world_size=4, seq_len=2048, tier="A", tokens_per_sec=9424.0),
    ])
    out = add_scaling_efficiency(df)
    assert out.loc[0, "scaling_efficiency_pct"] == pytest.approx(50.0)
    assert out.loc[1, "scaling_efficiency_pct"] == pytest.approx(36.5028, abs=1e-3)
    assert out.loc[2, "scaling_efficiency_pct"] == pytest.approx(100.0)       # lone row
    assert out.loc[1, "scaling_efficiency_true_pct"] == pytest.approx(73.0056, abs=1e-3)
