"""Model tiers and strategy configuration files.

Tier table: reference ``benchmarking/train_harness.py:157-179`` (tiers A and B, verbatim
hyper-parameters).  ``7B`` is the decoder shape sketched in the reference's
``docs/ARCHITECTURE.md:452-461``; ``T`` is a tiny shape used by the CPU/gloo plumbing tests.
DeepSpeed-style JSON: reference ``configs/deepspeed/zero{2,3}.json`` read at ``TH:246-270``.
FSDP YAML: reference ``configs/fsdp/fsdp_config.yaml`` (parsed but never applied there,
``TH:229-237,493``; honoured here).
"""
from __future__ import annotations

import json
from dataclasses import dataclass, field, asdict
from typing import Any, Dict, Optional

TIERS: Dict[str, Dict[str, int]] = {
    "A": dict(vocab_size=32000, n_embd=1024, n_head=16, n_layer=16),
    "B": dict(vocab_size=32000, n_embd=2048, n_head=32, n_layer=32),
    # 7B-shaped decoder (6.58 B params: L=32, d=4096, V=32000).  The reference sketches it with 32 heads of
    # 128 (docs/ARCHITECTURE.md:452-461); the native attention kernels are built for head_dim 64, so the
    # native tier uses 64 heads of 64 — same parameter count, same GEMM shapes, same attention FLOPs.
    "7B": dict(vocab_size=32000, n_embd=4096, n_head=64, n_layer=32),
    "T": dict(vocab_size=512, n_embd=128, n_head=2, n_layer=2),
}


def get_model_config(tier: str, seq_len: int) -> Dict[str, int]:
    """``block_size`` follows ``--seq-len`` exactly as in the reference (``TH:168,176``)."""
    if tier not in TIERS:
        raise ValueError(f"Unknown tier: {tier}")
    cfg = dict(TIERS[tier])
    cfg["block_size"] = int(seq_len)
    return cfg


@dataclass
class OptimConfig:
    """AdamW hyper-parameters.  Defaults = ``torch.optim.AdamW(lr=1e-4, weight_decay=0.01)``
    of the DDP/FSDP path (``TH:329``); the ZeRO JSON overrides them (``zero2.json:27-44``)."""

    lr: float = 1e-4
    beta1: float = 0.9
    beta2: float = 0.999
    eps: float = 1e-8
    weight_decay: float = 0.01
    grad_clip: float = 0.0          # 0 = off (DDP/FSDP reference paths do not clip)
    warmup_steps: int = 0           # WarmupLR 0 -> lr over this many optimizer steps
    warmup_min_lr: float = 0.0


@dataclass
class StrategyConfig:
    """Everything a strategy object needs that is not a CLI flag."""

    optim: OptimConfig = field(default_factory=OptimConfig)
    # ZeRO / DDP bucket sizes in ELEMENTS (DeepSpeed convention, ``zero2.json:19-23``)
    reduce_bucket_size: int = int(5e8)
    allgather_bucket_size: int = int(5e8)
    overlap_comm: bool = True
    # DDP bucket cap in MiB (torch DDP default 25)
    ddp_bucket_cap_mb: int = 25
    # ZeRO-3 knobs (``zero3.json:20-26``)
    stage3_prefetch_bucket_size: int = int(5e8)
    stage3_param_persistence_threshold: int = int(1e5)
    stage3_max_live_parameters: int = int(1e9)
    # FSDP knobs (``configs/fsdp/fsdp_config.yaml:4-11``)
    fsdp_sharding_strategy: str = "full_shard"
    fsdp_auto_wrap_policy: str = "transformer_block"
    fsdp_backward_prefetch: str = "backward_pre"
    fsdp_forward_prefetch: bool = True
    fsdp_sync_module_states: bool = True
    # "auto": keep the gathered blocks from forward to backward when they take < 5 % of device memory
    fsdp_reshard_after_forward: Any = "auto"
    # engine knobs settable from the strategy files ("b200bench" section of the JSON / keys of the YAML);
    # None / 0 = engine default.  Environment variables B200BENCH_* override them (engine.py).
    param_gather: Optional[str] = None      # "stream" | "gemm": fsdp / zero3 weight all-gather schedule
    gather_ctas: int = 0                    # CTAs of one block gather
    comm_ctas: int = 0                      # CTAs of one K13 all-reduce
    # gradient accumulation: "reference" = only ZeRO strategies honour --grad-accum
    grad_accum_mode: str = "reference"
    raw: Dict[str, Any] = field(default_factory=dict)

    def to_dict(self) -> Dict[str, Any]:
        return asdict(self)


def _num(x: Any) -> int:
    return int(float(x))


def _apply_engine_knobs(cfg: "StrategyConfig", section: Dict[str, Any]) -> None:
    pg = section.get("param_gather")
    if pg is not None:
        if pg not in ("stream", "gemm"):
            raise ValueError(f"param_gather must be 'stream' or 'gemm', got {pg!r}")
        cfg.param_gather = pg
    for key in ("gather_ctas", "comm_ctas"):
        if section.get(key) is not None:
            v = int(section[key])
            if not 0 <= v <= 512:
                raise ValueError(f"{key} must be between 0 (engine default) and 512, got {v}")
            setattr(cfg, key, v)


def load_deepspeed_config(path: str, micro_batch: int, grad_accum: int, world_size: int) -> StrategyConfig:
    """Parse a DeepSpeed-format JSON.  The three batch keys are replaced by integers computed
    from the CLI, as the reference does at ``TH:251-262``."""
    with open(path, "r") as f:
        raw = json.load(f)
    for k in ("train_batch_size", "train_micro_batch_size_per_gpu", "gradient_accumulation_steps"):
        raw.pop(k, None)
    raw["train_micro_batch_size_per_gpu"] = int(micro_batch)
    raw["gradient_accumulation_steps"] = int(grad_accum)
    raw["train_batch_size"] = int(micro_batch) * int(grad_accum) * int(world_size)

    cfg = StrategyConfig(raw=raw)
    cfg.optim.grad_clip = float(raw.get("gradient_clipping", 0.0))
    opt = raw.get("optimizer", {}).get("params", {})
    cfg.optim.lr = float(opt.get("lr", cfg.optim.lr))
    betas = opt.get("betas", [cfg.optim.beta1, cfg.optim.beta2])
    cfg.optim.beta1, cfg.optim.beta2 = float(betas[0]), float(betas[1])
    cfg.optim.eps = float(opt.get("eps", cfg.optim.eps))
    cfg.optim.weight_decay = float(opt.get("weight_decay", cfg.optim.weight_decay))
    sched = raw.get("scheduler", {})
    if sched.get("type") == "WarmupLR":
        sp = sched.get("params", {})
        cfg.optim.warmup_steps = int(sp.get("warmup_num_steps", 0))
        cfg.optim.warmup_min_lr = float(sp.get("warmup_min_lr", 0.0))
        cfg.optim.lr = float(sp.get("warmup_max_lr", cfg.optim.lr))
    z = raw.get("zero_optimization", {})
    cfg.reduce_bucket_size = _num(z.get("reduce_bucket_size", cfg.reduce_bucket_size))
    cfg.allgather_bucket_size = _num(z.get("allgather_bucket_size", cfg.allgather_bucket_size))
    cfg.overlap_comm = bool(z.get("overlap_comm", True))
    cfg.stage3_prefetch_bucket_size = _num(z.get("stage3_prefetch_bucket_size", cfg.stage3_prefetch_bucket_size))
    cfg.stage3_param_persistence_threshold = _num(
        z.get("stage3_param_persistence_threshold", cfg.stage3_param_persistence_threshold))
    cfg.stage3_max_live_parameters = _num(z.get("stage3_max_live_parameters", cfg.stage3_max_live_parameters))
    _apply_engine_knobs(cfg, raw.get("b200bench", {}))
    for dev_key in ("offload_optimizer", "offload_param"):
        dev = z.get(dev_key, {}).get("device", "none")
        if dev not in ("none", None):
            raise ValueError(f"{dev_key}.device={dev!r}: offload is not supported (reference uses 'none')")
    return cfg


def load_fsdp_config(path: Optional[str]) -> StrategyConfig:
    cfg = StrategyConfig()
    if not path:
        return cfg
    import yaml

    with open(path, "r") as f:
        raw = yaml.safe_load(f) or {}
    body = raw.get("fsdp_config", raw)
    cfg.raw = raw
    cfg.fsdp_sharding_strategy = str(body.get("sharding_strategy", cfg.fsdp_sharding_strategy))
    cfg.fsdp_auto_wrap_policy = str(body.get("auto_wrap_policy", cfg.fsdp_auto_wrap_policy))
    cfg.fsdp_backward_prefetch = str(body.get("backward_prefetch", cfg.fsdp_backward_prefetch))
    cfg.fsdp_forward_prefetch = bool(body.get("forward_prefetch", cfg.fsdp_forward_prefetch))
    cfg.fsdp_sync_module_states = bool(body.get("sync_module_states", cfg.fsdp_sync_module_states))
    r = body.get("reshard_after_forward", cfg.fsdp_reshard_after_forward)
    if not (isinstance(r, bool) or r == "auto"):
        raise ValueError(f"reshard_after_forward must be true, false or auto, got {r!r}")
    cfg.fsdp_reshard_after_forward = r
    _apply_engine_knobs(cfg, body)
    if body.get("cpu_offload", False):
        raise ValueError("cpu_offload is not supported (reference: CPUOffload(False), TH:234)")
    return cfg
