#!/usr/bin/env python3
"""Per-op device-time table of one eager training step (1 GPU): which kernels the step is made of.

    python scripts/op_timing.py [--tier A] [--seq-len 2048] [--steps 3] > profiles/round2/op_timing.md

Every front-end call is bracketed by CUDA events (b200bench.ops.enable_op_timing); the engine runs without CUDA
graph and without side-stream overlap, so the table is the serial sum of kernel times, not the step time.
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200bench import ops as _ops  # noqa: E402
from b200bench.config import StrategyConfig, get_model_config  # noqa: E402
from b200bench.engine import NativeEngine  # noqa: E402
from b200bench.parallel.collectives import DistCtx  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tier", default="A")
    ap.add_argument("--seq-len", type=int, default=2048)
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--dropout", type=float, default=0.1)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    o = _ops.load()
    cfg = get_model_config(a.tier, a.seq_len)
    eng = NativeEngine(cfg, "ddp", StrategyConfig(), DistCtx(), dev, dropout=a.dropout, batch=a.batch,
                       seq_len=a.seq_len, use_graph=False)
    eng.overlap_wgrad = False
    tok = torch.randint(0, cfg["vocab_size"], (a.batch, a.seq_len), device=dev, dtype=torch.int32)
    for _ in range(3):
        eng.train_step(tok, tok)
    torch.cuda.synchronize()
    _ops.enable_op_timing(o)
    for _ in range(a.steps):
        eng.train_step(tok, tok)
    print(f"# Per-op device time of one training step, serial (tier {a.tier}, T {a.seq_len}, batch {a.batch}, "
          f"dropout {a.dropout}; {torch.cuda.get_device_name(0)})\n")
    print(_ops.op_timing_report(o, a.steps))


if __name__ == "__main__":
    main()
