#!/usr/bin/env python3
"""Phase breakdown of one native training step (forward / backward / optimizer), graph-replayed.
Usage: python scripts/profile_step.py [--tier A] [--seq-len 2048] [--batch 1] [--dropout 0.1]"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200bench.config import StrategyConfig, get_model_config  # noqa: E402
from b200bench.engine import NativeEngine  # noqa: E402
from b200bench.parallel.collectives import DistCtx  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--tier", default="A")
ap.add_argument("--seq-len", type=int, default=2048)
ap.add_argument("--batch", type=int, default=1)
ap.add_argument("--dropout", type=float, default=0.1)
a = ap.parse_args()
dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
cfg = get_model_config(a.tier, a.seq_len)
eng = NativeEngine(cfg, "ddp", StrategyConfig(), DistCtx(), dev, dropout=a.dropout, batch=a.batch, seq_len=a.seq_len)
tok = torch.randint(0, cfg["vocab_size"], (a.batch, a.seq_len), device=dev, dtype=torch.int32)
for _ in range(4):
    eng.train_step(tok, tok)
torch.cuda.synchronize()
print(json.dumps({"tier": a.tier, "seq_len": a.seq_len, "batch": a.batch, "dropout": a.dropout,
                  **eng.profile_breakdown()}))
