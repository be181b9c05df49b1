"""Command line of the harness.  Flags, defaults and validation are those of the reference
(``benchmarking/train_harness.py:465-504``); the extra flags below them are additions.

    python -m b200bench.cli --strategy ddp --world-size 1 --rank 0 --tier A --seq-len 2048 \
        --steps 50 --per-device-batch 1 --grad-accum 4 --results-dir results
"""
from __future__ import annotations

import argparse
from typing import List, Optional


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="Distributed Training Benchmark (B200-native)")
    # Strategy
    p.add_argument("--strategy", type=str, required=True, choices=["ddp", "fsdp", "zero2", "zero3"],
                   help="Distributed strategy")
    # Distributed
    p.add_argument("--world-size", type=int, required=True, help="Total number of GPUs")
    p.add_argument("--rank", type=int, required=True, help="Global rank")
    p.add_argument("--local-rank", type=int, default=0, help="Local rank")
    p.add_argument("--master-addr", type=str, default="localhost", help="Master address")
    p.add_argument("--master-port", type=int, default=29500, help="Master port")
    # Model & Data
    p.add_argument("--tier", type=str, required=True, choices=["A", "B", "7B", "T"], help="Model tier")
    p.add_argument("--seq-len", type=int, required=True, help="Sequence length")
    p.add_argument("--synthetic", action="store_true", help="Use synthetic data (always on)")
    # Training
    p.add_argument("--steps", type=int, required=True, help="Training steps")
    p.add_argument("--warmup-steps", type=int, default=5, help="Warmup steps")
    p.add_argument("--per-device-batch", type=int, required=True, help="Batch size per GPU")
    p.add_argument("--grad-accum", type=int, required=True, help="Gradient accumulation steps")
    # Configs
    p.add_argument("--deepspeed-config", type=str, help="DeepSpeed-format config JSON")
    p.add_argument("--fsdp-config", type=str, help="FSDP config YAML")
    # Output
    p.add_argument("--results-dir", type=str, required=True, help="Results directory")

    x = p.add_argument_group("b200bench extensions")
    x.add_argument("--impl", choices=["auto", "native", "torch"], default="auto",
                   help="native = sm_100a kernels + native engine; torch = PyTorch ops + torch.distributed")
    x.add_argument("--comm", choices=["fused", "nccl"], default="fused",
                   help="native path: fused peer-memory collectives, or NCCL for A/B comparison")
    x.add_argument("--attn", choices=["bidir", "causal"], default="bidir",
                   help="bidir = reference behaviour (no mask, TH:127)")
    x.add_argument("--shift-targets", action="store_true", help="next-token targets (reference: unshifted)")
    x.add_argument("--dropout", type=float, default=0.1, help="dropout probability (reference: 0.1)")
    x.add_argument("--dtype", choices=["auto", "bf16", "fp32"], default="auto")
    x.add_argument("--timer", choices=["device", "host"], default="device",
                   help="host = the reference's perf_counter timer without device sync")
    x.add_argument("--grad-accum-mode", choices=["reference", "uniform"], default="reference",
                   help="reference: only zero2/zero3 honour --grad-accum (as the reference does)")
    x.add_argument("--backend", choices=["nccl", "gloo"], default=None)
    x.add_argument("--device", choices=["auto", "cuda", "cpu"], default="auto")
    x.add_argument("--dataset-size", type=int, default=1000)
    x.add_argument("--log-interval", type=int, default=10)
    x.add_argument("--save-state-dict", type=str, default=None)
    x.add_argument("--load-state-dict", type=str, default=None)
    x.add_argument("--profile", action="store_true", help="add a per-phase time breakdown to the result")
    x.add_argument("--nvtx", action="store_true",
                   help="NVTX range per training step (and per phase when the engine runs eagerly) for nsys / ncu")
    x.add_argument("--no-graph", action="store_true", help="run the native engine eagerly instead of CUDA-graph replay")
    return p


def parse_args(argv: Optional[List[str]] = None) -> argparse.Namespace:
    args = build_parser().parse_args(argv)
    if args.strategy in ("zero2", "zero3") and not args.deepspeed_config:
        raise ValueError("DeepSpeed strategy requires --deepspeed-config")
    return args


def main(argv: Optional[List[str]] = None) -> None:
    from .harness import train

    train(parse_args(argv))


if __name__ == "__main__":
    main()
