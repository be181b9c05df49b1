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
    ddp_bucket_cap_mb: int = 0       # 0 = engine default (see NativeEngine.blocks_per_bucket); torch DDP's default is 25
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
    notes: list = field(default_factory=list)     # how non-native values of a strategy file were interpreted

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


# ---- strict key handling: every key of a strategy file is either applied, or accepted only with the one value
# that matches what the engine does anyway, or rejected.  Nothing is parsed and then silently ignored.
_DS_TOP_HONOURED = {"zero_optimization", "bf16", "optimizer", "scheduler", "gradient_clipping", "b200bench",
                    "train_batch_size", "train_micro_batch_size_per_gpu", "gradient_accumulation_steps", "_doc"}
# reporting-only keys of DeepSpeed: no effect on the computation, any value accepted
_DS_TOP_INERT = {"steps_per_print", "wall_clock_breakdown", "flops_profiler", "zero_allow_untested_optimizer"}
_DS_ZERO_HONOURED = {"stage", "overlap_comm", "reduce_bucket_size", "stage3_prefetch_bucket_size",
                     "stage3_param_persistence_threshold", "stage3_max_live_parameters", "offload_optimizer",
                     "offload_param"}
# keys whose only supported value is the one the flat-buffer engine implements by construction
_DS_ZERO_FIXED = {
    "contiguous_gradients": (True, "gradients always live in one flat buffer"),
    "reduce_scatter": (True, "gradients are always reduce-scattered into the owning partition (K15)"),
    "allgather_partitions": (True, "updated partitions are always pushed to every rank by K15"),
    "round_robin_gradients": (None, "partitions are contiguous ranges of the flat buffer; accepted, no effect"),
    "stage3_gather_16bit_weights_on_model_save": (True, "state_dict() always gathers the bf16 weights"),
}
# size hints that are upper bounds in DeepSpeed and are satisfied by construction here (one fused launch)
_DS_ZERO_BOUNDS = {"allgather_bucket_size", "sub_group_size", "stage3_max_reuse_distance"}


def _check_deepspeed_keys(raw: Dict[str, Any]) -> None:
    for k in raw:
        if k not in _DS_TOP_HONOURED and k not in _DS_TOP_INERT:
            raise ValueError(f"DeepSpeed config key {k!r} is not supported by b200bench (it would be ignored)")
    if raw.get("bf16", {}).get("enabled", True) is not True:
        raise ValueError("bf16.enabled must be true: the native engine computes in bf16 with fp32 master weights")
    if raw.get("fp16", {}).get("enabled", False):
        raise ValueError("fp16 is not supported (bf16 needs no loss scaling)")
    opt_type = raw.get("optimizer", {}).get("type", "AdamW")
    if opt_type not in ("AdamW", "Adam"):
        raise ValueError(f"optimizer.type {opt_type!r} is not supported (AdamW)")
    sched = raw.get("scheduler", {}).get("type")
    if sched not in (None, "WarmupLR"):
        raise ValueError(f"scheduler.type {sched!r} is not supported (WarmupLR or none)")
    z = raw.get("zero_optimization", {})
    if int(z.get("stage", 0)) not in (2, 3):
        raise ValueError(f"zero_optimization.stage must be 2 or 3, got {z.get('stage')!r}")
    for k, v in z.items():
        if k in _DS_ZERO_HONOURED or k in _DS_ZERO_BOUNDS:
            continue
        if k in _DS_ZERO_FIXED:
            want, why = _DS_ZERO_FIXED[k]
            if want is not None and v is not want:
                raise ValueError(f"zero_optimization.{k}={v!r} is not supported: {why}")
            continue
        raise ValueError(f"zero_optimization.{k} is not supported by b200bench (it would be ignored)")


_FSDP_KEYS = {"sharding_strategy", "auto_wrap_policy", "backward_prefetch", "forward_prefetch", "sync_module_states",
              "reshard_after_forward", "param_gather", "gather_ctas", "comm_ctas", "mixed_precision", "cpu_offload",
              "ddp_bucket_cap_mb"}


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
    _check_deepspeed_keys(raw)

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
    for k in body:
        if k not in _FSDP_KEYS:
            raise ValueError(f"FSDP config key {k!r} is not supported by b200bench (it would be ignored)")
    cfg.fsdp_sharding_strategy = str(body.get("sharding_strategy", cfg.fsdp_sharding_strategy)).lower()
    if cfg.fsdp_sharding_strategy not in ("full_shard", "shard_grad_op", "no_shard"):
        raise ValueError(f"sharding_strategy must be full_shard, shard_grad_op or no_shard, got "
                         f"{cfg.fsdp_sharding_strategy!r} (hybrid sharding needs more than one node)")
    awp = str(body.get("auto_wrap_policy", cfg.fsdp_auto_wrap_policy)).lower()
    if awp == "size_based":
        # the reference's stock file (configs/fsdp/fsdp_config.yaml:6 there): its policy wraps nothing at these
        # model sizes, i.e. ONE flat unit without overlap.  Accepted so that the file loads; the
        # engine's gather unit is the transformer block, and that is said out loud.
        cfg.notes.append("auto_wrap_policy: size_based read as transformer_block (the engine's gather unit is one "
                         "transformer block; the reference's size-based policy yields a single flat unit)")
        awp = "transformer_block"
    if awp != "transformer_block":
        raise ValueError(f"auto_wrap_policy must be transformer_block (or the reference's size_based), got {awp!r}")
    cfg.fsdp_auto_wrap_policy = awp
    cfg.fsdp_backward_prefetch = str(body.get("backward_prefetch", cfg.fsdp_backward_prefetch)).lower()
    if cfg.fsdp_backward_prefetch not in ("backward_pre", "backward_post", "none"):
        raise ValueError(f"backward_prefetch must be backward_pre, backward_post or none, got "
                         f"{cfg.fsdp_backward_prefetch!r}")
    mp = body.get("mixed_precision", "bf16")
    if mp is False or str(mp).lower() in ("false", "none"):
        # reference file: "false # handled by autocast" - there fp16 autocast over fp32 weights (TH:334-376)
        cfg.notes.append("mixed_precision: false read as bf16 (the reference leaves precision to fp16 autocast; the "
                         "native engine always computes in bf16 with fp32 master weights and optimizer state)")
    elif str(mp).lower() != "bf16":
        raise ValueError(f"mixed_precision must be bf16 (or the reference's false), got {mp!r}")
    if body.get("ddp_bucket_cap_mb") is not None:
        cfg.ddp_bucket_cap_mb = int(body["ddp_bucket_cap_mb"])
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
