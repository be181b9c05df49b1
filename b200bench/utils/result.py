"""Result schema, file name and stdout marker protocol.

The 13 keys, their order, the file name ``result_{strategy}_ws{N}_seq{T}_tier{X}.json`` and the
``BENCHMARK_RESULT_JSON_START/END`` markers are those of the reference
(``benchmarking/train_harness.py:415-456``); metric formulas follow ``TH:399-413`` including the
"4 bytes per token" H2D figure.  Extra keys are appended after the 13 so that
reference tooling that indexes by name keeps working.
"""
from __future__ import annotations

import json
import os
import re
from typing import Any, Dict, List, Optional

RESULT_KEYS = [
    "strategy", "world_size", "rank", "seq_len", "tier", "steps", "per_device_batch", "grad_accum",
    "tokens_per_sec", "mean_step_time_sec", "mean_loss", "peak_vram_gb", "h2d_gbps_per_gpu",
]
MARK_START = "BENCHMARK_RESULT_JSON_START"
MARK_END = "BENCHMARK_RESULT_JSON_END"


def compute_metrics(step_times: List[float], losses: List[float], per_device_batch: int, seq_len: int,
                    world_size: int, peak_vram_bytes: int) -> Dict[str, float]:
    mean_step = sum(step_times) / len(step_times) if step_times else 0.0
    mean_loss = sum(losses) / len(losses) if losses else 0.0
    tokens_per_step = per_device_batch * seq_len * world_size          # TH:403 (no grad_accum factor)
    tps = tokens_per_step / mean_step if mean_step > 0 else 0.0
    bytes_per_step = per_device_batch * seq_len * 4                     # TH:412
    h2d = (bytes_per_step / mean_step) / 1e9 if mean_step > 0 else 0.0
    return dict(tokens_per_sec=tps, mean_step_time_sec=mean_step, mean_loss=mean_loss,
                peak_vram_gb=peak_vram_bytes / 1e9, h2d_gbps_per_gpu=h2d)


def make_result(args: Any, metrics: Dict[str, float], extra: Optional[Dict[str, Any]] = None) -> Dict[str, Any]:
    res: Dict[str, Any] = {
        "strategy": args.strategy, "world_size": args.world_size, "rank": args.rank,
        "seq_len": args.seq_len, "tier": args.tier, "steps": args.steps,
        "per_device_batch": args.per_device_batch, "grad_accum": args.grad_accum,
    }
    for k in RESULT_KEYS[8:]:
        res[k] = metrics[k]
    if extra:
        for k, v in extra.items():
            if k not in res:
                res[k] = v
    return res


def result_filename(strategy: str, world_size: int, seq_len: int, tier: str) -> str:
    return f"result_{strategy}_ws{world_size}_seq{seq_len}_tier{tier}.json"


def write_result(result: Dict[str, Any], results_dir: str) -> str:
    os.makedirs(results_dir, exist_ok=True)
    path = os.path.join(results_dir, result_filename(result["strategy"], result["world_size"],
                                                     result["seq_len"], result["tier"]))
    with open(path, "w") as f:
        json.dump(result, f, indent=2)
    return path


def format_marked(result: Dict[str, Any]) -> str:
    return "\n".join([MARK_START, json.dumps(result, indent=2), MARK_END])


def extract_marked(log_text: str) -> Optional[Dict[str, Any]]:
    """Inverse of ``format_marked``; equivalent to the reference collector's
    ``sed -n '/START/,/END/p' | sed '1d;$d'`` (``scripts/collect_results.sh:50-52``)."""
    m = re.search(re.escape(MARK_START) + r"\s*\n(.*?)\n\s*" + re.escape(MARK_END), log_text, re.S)
    if not m:
        return None
    return json.loads(m.group(1))
