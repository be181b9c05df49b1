"""parse_metrics -> plot -> make_report on synthetic result files (golden structure)."""
import json

import pandas as pd

from b200bench.analysis import make_report, parse_metrics, plot


def _write(root, job, **kw):
    d = root / f"{job}_results"
    d.mkdir(parents=True)
    rec = dict(strategy="ddp", world_size=1, rank=0, seq_len=2048, tier="A", steps=100, per_device_batch=1,
               grad_accum=4, tokens_per_sec=1000.0, mean_step_time_sec=2.0, mean_loss=6.0, peak_vram_gb=10.0,
               h2d_gbps_per_gpu=1e-5)
    rec.update(kw)
    (d / "result.json").write_text(json.dumps(rec))


def test_pipeline(tmp_path):
    res = tmp_path / "results"
    for s, base in (("ddp", 1000.0), ("zero2", 1500.0)):
        for ws in (1, 2, 4, 8):
            _write(res, f"{s}-{ws}", strategy=s, world_size=ws, tokens_per_sec=base * ws * 0.9 ** (ws > 1),
                   peak_vram_gb=10.0 / (ws if s == "zero2" else 1))
    _write(res, "zero2-8-long", strategy="zero2", world_size=8, seq_len=4096, tokens_per_sec=9000.0, peak_vram_gb=3.0)
    csv = parse_metrics.parse_results(str(res), str(tmp_path / "summary"))
    df = pd.read_csv(csv)
    assert len(df) == 9
    assert list(df.columns[:13]) == ["strategy", "world_size", "rank", "seq_len", "tier", "steps",
                                     "per_device_batch", "grad_accum", "tokens_per_sec", "mean_step_time_sec",
                                     "mean_loss", "peak_vram_gb", "h2d_gbps_per_gpu"]
    assert "scaling_efficiency_pct" in df.columns
    row = df[(df.strategy == "ddp") & (df.world_size == 8)].iloc[0]
    assert abs(row.scaling_efficiency_pct - 90.0) < 1e-6
    made = plot.plot_metrics(str(csv), str(tmp_path / "summary" / "plots"))
    names = sorted(p.name for p in made)
    assert names == sorted(["tokens_per_sec_vs_gpu.png", "step_time_vs_gpu.png", "vram_vs_seqlen.png",
                            "scaling_efficiency.png", "gbps_vs_gpu.png"])
    for p in made:
        assert p.stat().st_size > 2000
    rep = make_report.generate_report(str(csv), str(tmp_path / "summary")).read_text()
    for section in ("# Distributed Training Benchmark Report", "## Summary", "## Strategy Comparison",
                    "### ZERO2", "## Key Findings", "**Best Throughput:**", "## Strategy Trade-offs",
                    "## Visualizations", "plots/vram_vs_seqlen.png"):
        assert section in rep
    assert "A10" not in rep


def test_single_seq_len_skips_vram_plot(tmp_path):
    res = tmp_path / "results"
    _write(res, "a", world_size=2)
    _write(res, "b", world_size=4, tokens_per_sec=1800.0)
    csv = parse_metrics.parse_results(str(res), str(tmp_path / "s"))
    made = plot.plot_metrics(str(csv), str(tmp_path / "s" / "plots"))
    assert "vram_vs_seqlen.png" not in [p.name for p in made]
    rep = make_report.generate_report(str(csv), str(tmp_path / "s")).read_text()
    assert "vram_vs_seqlen.png" not in rep


def test_hardware_efficiency_section_appears_with_roofline_columns(tmp_path):
    res = tmp_path / "results"
    _write(res, "a", world_size=1, model_tflops_per_gpu=428.0, step_roofline_ms=3.59, fraction_of_step_roofline=0.42,
           mean_step_time_sec=0.00864)
    _write(res, "b", world_size=8, tokens_per_sec=7000.0)                 # an older result without the columns
    csv = parse_metrics.parse_results(str(res), str(tmp_path / "s"))
    rep = make_report.generate_report(str(csv), str(tmp_path / "s")).read_text()
    assert "## Hardware Efficiency" in rep and "428" in rep and "0.42" in rep
    assert rep.index("## Hardware Efficiency") < rep.index("## Key Findings")
