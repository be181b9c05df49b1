"""Per-GPU memory model of the native engine: what a (tier, strategy, world size, sequence length) run
allocates, component by component, without touching a GPU.

    python -m b200bench.analysis.memory --tier A --strategy zero3 --world-size 8 --seq-len 2048

The reference documents memory only as measured peaks (``README.md:216-222``: 13.97 GB for every DDP row,
dominated by fp32 weights and the T x T attention probabilities kept for backward).  Here every
buffer is allocated once in ``NativeEngine.__init__`` (``engine.py``), so the peak is a closed form; the
numbers below agree with the measured ``peak_vram_gb`` of the round-1 runs to within 2 % (tests/test_memory_model.py).
"""
from __future__ import annotations

import argparse
from typing import Dict, Optional

from ..config import get_model_config
from ..models.layout import build_layout

BF16, FP32 = 2, 4


def estimate_memory(model_cfg: Dict[str, int], strategy: str, world_size: int = 1, batch: int = 1,
                    seq_len: int = 2048, comm: str = "fused", keep_gathered: Optional[bool] = None,
                    max_live_parameters: int = int(1e9), device_bytes: int = 180 * 10**9) -> Dict[str, float]:
    """Bytes per GPU by component; ``total`` is their sum.  ``keep_gathered=None`` applies the engine's policy
    (ZeRO-3: matrices <= stage3_max_live_parameters; FSDP: gathered blocks < 5 % of device memory)."""
    ws = max(1, world_size)
    Lt = build_layout(world_size=ws, **model_cfg)
    d, V, L, H = model_cfg["n_embd"], model_cfg["vocab_size"], model_cfg["n_layer"], model_cfg["n_head"]
    M = batch * seq_len
    total, big = Lt.total, Lt.big_end
    small = total - big
    fused = ws > 1 and comm == "fused"
    sharded_params = fused and strategy in ("fsdp", "zero3")
    sharded_opt = strategy in ("zero2", "fsdp", "zero3")      # the engine keeps a reduced-gradient shard even at ws 1
    out: Dict[str, float] = {}

    out["gradients_bf16"] = total * BF16                                   # full flat buffer on every rank
    if sharded_params:
        block = Lt.layer_range(0)[1] - Lt.layer_range(0)[0]
        if keep_gathered is None:
            keep_gathered = (big <= max_live_parameters) if strategy == "zero3" else (L * block * BF16 <= 0.05 * device_bytes)
        slots = L if keep_gathered else 2
        out["weights_bf16"] = (big // ws + small) * BF16                   # own shard + replicated vector tail
        out["master_and_adam_fp32"] = (big // ws + small) * 3 * FP32
        out["grad_shard_bf16"] = (big // ws) * BF16
        out["gather_staging_bf16"] = (slots * block + 2 * 4 * d * d + (V + seq_len) * d) * BF16
    else:
        n_opt = total // ws if sharded_opt else total
        out["weights_bf16"] = total * BF16
        out["master_and_adam_fp32"] = n_opt * 3 * FP32
        if sharded_opt:
            out["grad_shard_bf16"] = n_opt * BF16
    # activations saved for backward: x (L+1), ln1, ln2, att, xmid (d each), qkv (3d), pre + act (4d each), lnf
    per_layer = (1 + 1 + 1 + 1 + 1 + 3 + 4 + 4) * M * d * BF16 + (4 * M + batch * H * seq_len) * FP32
    out["activations"] = L * per_layer + 2 * M * d * BF16 + 2 * M * FP32
    out["logits_bf16"] = M * V * BF16 + M * FP32
    # backward temporaries shared by all blocks: dres dmid dln dln2 dy_drop datt (d), dqkv (3d), dbig (4d), dq fp32
    out["backward_temporaries"] = (6 + 3 + 4) * M * d * BF16 + M * d * FP32 + batch * H * seq_len * FP32
    out["total"] = float(sum(out.values()))
    return out


def main() -> None:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--tier", default="A")
    ap.add_argument("--strategy", default="ddp", choices=["ddp", "fsdp", "zero2", "zero3"])
    ap.add_argument("--world-size", type=int, default=1)
    ap.add_argument("--per-device-batch", type=int, default=1)
    ap.add_argument("--seq-len", type=int, default=2048)
    ap.add_argument("--comm", default="fused", choices=["fused", "nccl"])
    a = ap.parse_args()
    est = estimate_memory(get_model_config(a.tier, a.seq_len), a.strategy, a.world_size, a.per_device_batch,
                          a.seq_len, a.comm)
    print(f"tier {a.tier}, {a.strategy}, {a.world_size} GPU(s), batch {a.per_device_batch} x seq {a.seq_len}: per-GPU memory")
    for k, v in est.items():
        print(f"  {k:24s} {v / 1e9:9.3f} GB")


if __name__ == "__main__":
    main()
