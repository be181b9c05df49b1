"""results/**/result.json -> metrics.csv.

CLI, input discovery (``rglob("result.json")``), sort order, the ``scaling_efficiency_pct`` column
and the output file name are those of the reference's ``scripts/parse_metrics.py:12-83``.  The
reference's efficiency divides by ``tps(min_ws) * world_size`` (not ``world_size / min_ws``,
``parse_metrics.py:56-62``); that column is reproduced as is, and
``scaling_efficiency_true_pct`` (relative to the smallest world size actually present) is added.
Harness-native ``result_*.json`` files are accepted too.
"""
from __future__ import annotations

import argparse
import json
from pathlib import Path
from typing import List, Optional

import pandas as pd

SUMMARY_COLS = ["strategy", "world_size", "seq_len", "tier", "tokens_per_sec", "mean_step_time_sec",
                "peak_vram_gb", "h2d_gbps_per_gpu", "scaling_efficiency_pct"]


def add_scaling_efficiency(df: pd.DataFrame) -> pd.DataFrame:
    df = df.copy()
    grp = df.groupby(["strategy", "seq_len"], sort=False)
    min_ws = grp["world_size"].transform("min")
    # tokens/s of the smallest world size in each (strategy, seq_len) group
    base = df[df["world_size"] == min_ws].drop_duplicates(["strategy", "seq_len"]).set_index(
        ["strategy", "seq_len"])["tokens_per_sec"]
    base_tps = pd.Series([base[(s, q)] for s, q in zip(df["strategy"], df["seq_len"])], index=df.index)
    n_in_group = grp["world_size"].transform("size")
    ideal_ref = base_tps * df["world_size"]
    eff = (df["tokens_per_sec"] / ideal_ref * 100.0).where(ideal_ref > 0, 0.0)
    df["scaling_efficiency_pct"] = eff.where(n_in_group > 1, 100.0)      # lone row -> 100 %
    ideal_true = base_tps * df["world_size"] / min_ws
    df["scaling_efficiency_true_pct"] = (df["tokens_per_sec"] / ideal_true * 100.0).where(ideal_true > 0, 0.0)
    return df


def collect_results(results_dir: str) -> Optional[pd.DataFrame]:
    root = Path(results_dir)
    files: List[Path] = sorted(root.rglob("result.json")) + sorted(root.rglob("result_*.json"))
    if not files:
        print(f"WARNING: No result.json files found in {results_dir}")
        print("Expected files like: <job>_results/result.json or result_<strategy>_ws<N>_seq<T>_tier<X>.json")
        return None
    print(f"Found {len(files)} result files")
    records = []
    for jf in files:
        try:
            rec = json.loads(jf.read_text())
            rec = {k: v for k, v in rec.items() if not isinstance(v, (dict, list))}
            records.append(rec)
        except Exception as e:  # a truncated log must not kill the whole report
            print(f"WARNING: Failed to parse {jf}: {e}")
    if not records:
        print("ERROR: No valid records parsed")
        return None
    df = pd.DataFrame(records).drop_duplicates(["strategy", "world_size", "seq_len", "tier"], keep="last")
    df = df.sort_values(by=["strategy", "world_size", "seq_len"]).reset_index(drop=True)
    return add_scaling_efficiency(df)


def parse_results(results_dir: str, output_dir: str) -> Optional[Path]:
    df = collect_results(results_dir)
    if df is None:
        return None
    out = Path(output_dir)
    out.mkdir(parents=True, exist_ok=True)
    csv_file = out / "metrics.csv"
    df.to_csv(csv_file, index=False)
    print(f"\nMetrics saved to: {csv_file}")
    print("\n=== Summary ===")
    print(df[SUMMARY_COLS].to_string(index=False))
    return csv_file


def main(argv=None) -> None:
    ap = argparse.ArgumentParser(description="Parse benchmark results to CSV")
    ap.add_argument("--results-dir", required=True, help="Directory containing result JSON files")
    ap.add_argument("--out", required=True, help="Output directory for metrics.csv")
    a = ap.parse_args(argv)
    parse_results(a.results_dir, a.out)


if __name__ == "__main__":
    main()
