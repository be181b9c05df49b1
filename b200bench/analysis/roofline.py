"""Roofline of one training step per strategy: model FLOPs, optimizer HBM bytes and NVLink bytes in closed form,
divided by the MEASURED peaks (``MEASURED_PEAKS.json``: HBM copy bandwidth, cuBLAS bf16 rate; NVLink 770 GB/s per
direction from the profiling recipe).  ``BASELINE.json`` asks for every fused path as "achieved fraction of its
roofline (the slower of its compute at peak and its bytes over NVLink at link bandwidth)"; this module is where
those denominators come from.

    python -m b200bench.analysis.roofline --tier A --strategy ddp --world-size 8 --measured-ms 9.60

The reference has no such accounting (tokens/s and step time only).
"""
from __future__ import annotations

import argparse
import json
import os
from typing import Dict, Optional

from ..config import get_model_config
from ..models.layout import build_layout

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
FALLBACK = {"hbm_gbs": 6454.6, "bf16_tflops_sustained": 1440.7, "bf16_tflops": 1684.1}
NVLINK_GBS = 770.0          # measured peer copy per direction (B200_PROFILING.md); nominal 900


def measured_peaks() -> Dict[str, float]:
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            raw = json.load(f)
        return {k: float(raw.get(k, v)) for k, v in FALLBACK.items()}
    except (OSError, ValueError):
        return dict(FALLBACK)


def step_work(model_cfg: Dict[str, int], strategy: str, world_size: int, batch: int, seq_len: int,
              grad_accum: int = 1, keep_gathered: bool = True) -> Dict[str, float]:
    """Work of ONE micro-step on ONE GPU (optimizer / collective work amortised over ``grad_accum``)."""
    ws = max(1, world_size)
    Lt = build_layout(world_size=ws, **model_cfg)
    d, V, L = model_cfg["n_embd"], model_cfg["vocab_size"], model_cfg["n_layer"]
    M = batch * seq_len
    matmul_params = L * 12 * d * d + V * d                      # every weight that meets every token in a GEMM
    gemm_flops = 6.0 * matmul_params * M                        # fwd + dgrad + wgrad
    attn_flops = 3.0 * 4.0 * batch * seq_len * seq_len * d * L  # QK^T and PV, fwd + 2x in bwd (recompute not counted)
    P, big = Lt.total, Lt.big_end
    S2 = P * 2.0                                                # bf16 bytes of the whole flat buffer
    n_opt = P / ws if strategy != "ddp" else P
    hbm_opt = n_opt * 28.0 / grad_accum                         # m, v, master fp32 r+w, grad bf16 r, weight bf16 w
    frac = (ws - 1) / ws
    if ws == 1:
        nv_in = nv_out = 0.0
    elif strategy == "ddp":                                     # two-shot all-reduce: pull my slice, push it back
        nv_in = nv_out = 2.0 * frac * S2
    elif strategy == "zero2":                                   # pull-reduce my partition, push updated weights
        nv_in = nv_out = 2.0 * frac * S2 / grad_accum
    else:                                                       # fsdp / zero3: reduce-scatter + weight gathers
        gathers = (1.0 / grad_accum) if keep_gathered else 2.0  # per micro-step: once per cycle, or fwd + bwd
        nv_in = frac * big * 2.0 * (1.0 / grad_accum + gathers)
        nv_out = nv_in
    return {"gemm_flops": gemm_flops, "attention_flops": attn_flops, "flops": gemm_flops + attn_flops,
            "hbm_optimizer_bytes": hbm_opt, "nvlink_bytes_in": nv_in, "nvlink_bytes_out": nv_out}


def roofline(model_cfg: Dict[str, int], strategy: str, world_size: int = 1, batch: int = 1, seq_len: int = 2048,
             grad_accum: int = 1, keep_gathered: bool = True, measured_ms: Optional[float] = None) -> Dict[str, float]:
    pk = measured_peaks()
    w = step_work(model_cfg, strategy, world_size, batch, seq_len, grad_accum, keep_gathered)
    t_compute = w["flops"] / (pk["bf16_tflops_sustained"] * 1e12) * 1e3
    t_hbm = w["hbm_optimizer_bytes"] / (pk["hbm_gbs"] * 1e9) * 1e3
    t_nvlink = max(w["nvlink_bytes_in"], w["nvlink_bytes_out"]) / (NVLINK_GBS * 1e9) * 1e3
    out = dict(w)
    out.update({"compute_ms": t_compute, "optimizer_hbm_ms": t_hbm, "nvlink_ms": t_nvlink,
                # everything overlapped perfectly / nothing overlapped at all
                "roofline_ms": max(t_compute + t_hbm, t_nvlink), "serial_ms": t_compute + t_hbm + t_nvlink})
    if measured_ms:
        out["measured_ms"] = measured_ms
        out["fraction_of_roofline"] = out["roofline_ms"] / measured_ms
        out["model_tflops_achieved"] = w["flops"] / (measured_ms * 1e-3) / 1e12
    return out


def main() -> None:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--tier", default="A")
    ap.add_argument("--strategy", default="ddp", choices=["ddp", "fsdp", "zero2", "zero3"])
    ap.add_argument("--world-size", type=int, default=1)
    ap.add_argument("--per-device-batch", type=int, default=1)
    ap.add_argument("--seq-len", type=int, default=2048)
    ap.add_argument("--grad-accum", type=int, default=1, help="effective accumulation (zero2/zero3: the --grad-accum flag)")
    ap.add_argument("--reshard", action="store_true", help="fsdp/zero3: weights gathered again in backward")
    ap.add_argument("--measured-ms", type=float, default=None, help="measured ms per micro-step")
    a = ap.parse_args()
    r = roofline(get_model_config(a.tier, a.seq_len), a.strategy, a.world_size, a.per_device_batch, a.seq_len,
                 a.grad_accum, not a.reshard, a.measured_ms)
    for k, v in r.items():
        print(f"  {k:26s} {v:16.4f}" if v < 1e6 else f"  {k:26s} {v:16.4e}")


if __name__ == "__main__":
    main()
