"""Training harness: strategy switch, hot loop, timers, metrics, result emission.

Reproduces the observable behaviour of the reference's ``train(args)``
(``benchmarking/train_harness.py:278-458``): banners, ``[Step 0010] Loss: ..., Time: ...s`` every
10 steps, metrics over post-warm-up steps only, the 13-key result JSON written to
``result_{strategy}_ws{N}_seq{T}_tier{X}.json`` and echoed between stdout markers.

Deliberate differences (each selectable back): device-event timing with the max over
ranks (``--timer host`` restores ``TH:361-385``), bf16 weights + fp32 master instead of fp16
autocast + GradScaler, identical initial weights on every rank.
"""
from __future__ import annotations

from typing import Any, Dict

import torch

from .config import StrategyConfig, get_model_config, load_deepspeed_config, load_fsdp_config
from .parallel.collectives import DistCtx
from .utils.data import PinnedBatchSource, SyntheticDataset
from .utils.result import compute_metrics, format_marked, make_result, write_result
from .utils.timer import StepTimer


def pick_device(args: Any) -> torch.device:
    want = getattr(args, "device", "auto")
    if want == "cpu" or (want == "auto" and not torch.cuda.is_available()):
        return torch.device("cpu")
    return torch.device(f"cuda:{args.local_rank}")


def load_strategy_config(args: Any) -> StrategyConfig:
    """wrap_model's config handling (TH:246-262); the FSDP YAML is honoured here."""
    if args.strategy in ("zero2", "zero3"):
        if not args.deepspeed_config:
            raise ValueError("DeepSpeed strategy requires --deepspeed-config")
        cfg = load_deepspeed_config(args.deepspeed_config, args.per_device_batch, args.grad_accum,
                                    args.world_size)
        stage = cfg.raw.get("zero_optimization", {}).get("stage")
        want = 2 if args.strategy == "zero2" else 3
        if stage is not None and int(stage) != want:
            print(f"WARNING: --strategy {args.strategy} but config says stage {stage}; using stage {want}")
    elif args.strategy == "fsdp":
        cfg = load_fsdp_config(getattr(args, "fsdp_config", None))
    else:
        cfg = StrategyConfig()
    cfg.grad_accum_mode = getattr(args, "grad_accum_mode", "reference")
    if cfg.notes and getattr(args, "rank", 0) == 0:
        for n in cfg.notes:                     # values of a strategy file that were read as something else
            print(f"NOTE ({args.strategy} config): {n}")
    return cfg


def build_engine(args: Any, model_cfg: Dict[str, int], cfg: StrategyConfig, ctx: DistCtx,
                 device: torch.device):
    impl = getattr(args, "impl", "auto")
    dtype = {"bf16": torch.bfloat16, "fp32": torch.float32}[getattr(args, "dtype", "auto") if
                                                               getattr(args, "dtype", "auto") != "auto"
                                                               else ("bf16" if device.type == "cuda" else "fp32")]
    causal = getattr(args, "attn", "bidir") == "causal"
    dropout = float(getattr(args, "dropout", 0.1))
    if impl == "auto":
        impl = "native" if device.type == "cuda" else "torch"
    if impl == "native":
        if device.type != "cuda":
            raise RuntimeError("--impl native needs a CUDA device")
        from .engine import NativeEngine  # raises loudly if libb200bench.so is missing

        return NativeEngine(model_cfg, args.strategy, cfg, ctx, device, grad_accum=args.grad_accum,
                            dropout=dropout, causal=causal, batch=args.per_device_batch,
                            seq_len=args.seq_len, comm=getattr(args, "comm", "fused"),
                            use_graph=not getattr(args, "no_graph", False), nvtx=getattr(args, "nvtx", False))
    from .parallel.torch_engine import TorchEngine

    native_ops = None
    return TorchEngine(model_cfg, args.strategy, cfg, ctx, device, dtype, grad_accum=args.grad_accum,
                       dropout=dropout, causal=causal, native_ops=native_ops)


def train(args: Any) -> Dict[str, Any]:
    device = pick_device(args)
    if device.type == "cuda":
        torch.cuda.set_device(device)
    ctx = DistCtx.setup(args.rank, args.world_size, args.master_addr, args.master_port, device,
                        backend=getattr(args, "backend", None))
    torch.manual_seed(42 + args.rank)                                   # TH:283-284
    if device.type == "cuda":
        torch.cuda.manual_seed(42 + args.rank)
    is_main = args.rank == 0

    if is_main:
        print("\n" + "=" * 80)
        print(f"Benchmark Config: {args.strategy.upper()} | Tier {args.tier} | WS={args.world_size} | "
              f"SeqLen={args.seq_len}")
        print("=" * 80 + "\n")

    model_cfg = get_model_config(args.tier, args.seq_len)
    cfg = load_strategy_config(args)
    engine = build_engine(args, model_cfg, cfg, ctx, device)
    if getattr(args, "load_state_dict", None):
        engine.load_state_dict(torch.load(args.load_state_dict, map_location=device))

    dataset = SyntheticDataset(vocab_size=model_cfg["vocab_size"], seq_len=args.seq_len,
                               size=int(getattr(args, "dataset_size", 1000)), seed=42, verbose=is_main)
    # ZeRO path of the reference uses no DistributedSampler: every rank draws the same batches (TH:313)
    same_data = args.strategy in ("zero2", "zero3") and getattr(args, "grad_accum_mode", "reference") == "reference"
    source = PinnedBatchSource(dataset, args.per_device_batch, 0 if same_data else args.rank,
                               1 if same_data else args.world_size, device)

    if device.type == "cuda":
        torch.cuda.reset_peak_memory_stats(device)                      # TH:338
    if is_main:
        print(f"Starting training: {args.steps} steps, warmup={args.warmup_steps}")
        print(f"Per-device batch: {args.per_device_batch}, Grad accum: {args.grad_accum}")
        print("")

    timer = StepTimer(getattr(args, "timer", "device"), device)
    shift = bool(getattr(args, "shift_targets", False))
    log_every = int(getattr(args, "log_interval", 10))
    loss_slots = torch.zeros(max(1, args.steps), dtype=torch.float32,
                             pin_memory=(device.type == "cuda"))
    profile = getattr(args, "profile", False)
    nvtx = bool(getattr(args, "nvtx", False)) and device.type == "cuda"

    sampler = None
    if device.type == "cuda":             # SM clock / throttle-reason record of the run (rank-local GPU, NVML thread)
        from .utils.clocks import ClockSampler
        sampler = ClockSampler(device.index if device.index is not None else 0).start()

    for step in range(args.steps):
        if nvtx:
            torch.cuda.nvtx.range_push(f"step {step}" + (" (warm-up)" if step < args.warmup_steps else ""))
        timer.start()
        tokens = source.next()                                          # H2D from pinned memory
        if shift:
            targets = torch.cat([tokens[:, 1:], torch.full_like(tokens[:, :1], -1)], dim=1)
        else:
            targets = tokens                                            # TH:359 (unshifted)
        loss = engine.train_step(tokens, targets)
        loss_slots[step].copy_(loss.detach().float().reshape(()), non_blocking=True)   # D2H of the step result
        dt = timer.stop(keep=step >= args.warmup_steps)
        if nvtx:
            torch.cuda.nvtx.range_pop()
        if is_main and log_every > 0 and step % log_every == 0:
            if device.type == "cuda":
                torch.cuda.current_stream().synchronize()
            print(f"[Step {step:04d}] Loss: {float(loss_slots[step]):.4f}, Time: {dt:.3f}s")

    if device.type == "cuda":
        torch.cuda.synchronize(device)
    clocks = sampler.stop() if sampler is not None else None
    ctx.barrier()                                                       # TH:396-397

    step_times = timer.times()
    losses = [float(x) for x in loss_slots[args.warmup_steps:args.steps]]
    mean_local = sum(step_times) / len(step_times) if step_times else 0.0
    mean_max = ctx.max_float(mean_local, device)                        # max over ranks
    peak_local = torch.cuda.max_memory_allocated(device) if device.type == "cuda" else 0
    peak_local += engine.extra_memory_bytes()
    peak_max = int(ctx.max_float(float(peak_local), device))
    scale = (mean_max / mean_local) if mean_local > 0 else 1.0
    metrics = compute_metrics([t * scale for t in step_times], losses, args.per_device_batch, args.seq_len,
                              args.world_size, peak_local)
    host_times = timer.host_times()
    extra = {
        "impl": type(engine).__name__,
        "timer": timer.mode,
        "mean_step_time_host_sec": (sum(host_times) / len(host_times)) if host_times else 0.0,
        "peak_vram_max_over_ranks_gb": peak_max / 1e9,
        "peak_vram_reserved_gb": (torch.cuda.max_memory_reserved(device) / 1e9) if device.type == "cuda" else 0.0,
        "grad_accum_effective": getattr(engine, "accum", 1),
        "h2d_bytes_per_step": source.bytes_per_step,
        "dtype": str(getattr(engine, "dtype", torch.bfloat16)).replace("torch.", ""),
        "attn": getattr(args, "attn", "bidir"),
        "gpu_launches_per_step": engine.kernel_launches_per_step(),
    }
    if clocks is not None:
        extra["clocks"] = {k: clocks.get(k) for k in ("sm_mhz", "sm_max_mhz", "reasons", "power_w_max", "samples")}
    try:                                  # closed-form work of one micro-step vs the measured peaks (analysis/roofline.py)
        from .analysis.roofline import roofline

        rf = roofline(model_cfg, args.strategy, args.world_size, args.per_device_batch, args.seq_len,
                      extra["grad_accum_effective"], bool(getattr(engine, "keep_gathered", True)),
                      metrics["mean_step_time_sec"] * 1e3 if metrics["mean_step_time_sec"] > 0 else None)
        extra["model_tflops_per_gpu"] = rf.get("model_tflops_achieved")
        extra["step_roofline_ms"] = rf["roofline_ms"]
        extra["fraction_of_step_roofline"] = rf.get("fraction_of_roofline")
    except Exception as e:  # never let the accounting break a benchmark run
        extra["roofline_error"] = repr(e)
    if profile and hasattr(engine, "profile_breakdown"):
        extra["profile"] = engine.profile_breakdown()
    result = make_result(args, metrics, extra)

    if is_main:
        print("\n" + "=" * 80)
        print("Benchmark Results:")
        print(f"  Tokens/sec:       {result['tokens_per_sec']:,.0f}")
        print(f"  Mean step time:   {result['mean_step_time_sec']:.4f}s")
        print(f"  Peak VRAM/GPU:    {result['peak_vram_gb']:.2f} GB")
        print(f"  H2D GB/s/GPU:     {result['h2d_gbps_per_gpu']:.3f}")
        print(f"  Mean loss:        {result['mean_loss']:.4f}")
        print("=" * 80 + "\n")
        path = write_result(result, args.results_dir)
        print(f"Results saved to: {path}")
        print("\n" + "=" * 80)
        print(format_marked(result))
        print("=" * 80 + "\n")

    if getattr(args, "save_state_dict", None):
        sd = engine.state_dict()                                        # collective for sharded strategies
        if is_main:
            torch.save({k: v.cpu() for k, v in sd.items()}, args.save_state_dict)
            print(f"state_dict saved to: {args.save_state_dict}")
    engine.close()
    ctx.cleanup()                                                       # TH:458
    return result
