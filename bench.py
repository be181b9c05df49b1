#!/usr/bin/env python3
"""Headline benchmark: TinyGPT tier A (236.4 M params), seq 2048, batch 1 per GPU, DDP, bf16,
synthetic tokens, random-init weights — the configuration behind the reference's published table
(README.md:195-223 of the reference).  Weak scaling: per-GPU work is fixed.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29500 bench.py --gpus 8 --steps 20 --warmup 5
    python bench.py --impl reference --gpus 1 --steps 20 --warmup 5      # unmodified reference arm

Prints ONE JSON line on rank 0.  `value` = whole-job tokens/s, device-timed (CUDA events, max over
ranks) over exactly --steps optimizer steps with tokens already resident on the device; `e2e` is
the same metric through the public harness path: every step copies its tokens from pinned host
memory and reads the loss back to the host.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "tokens_per_sec"
TIER, SEQ, BATCH, GRAD_ACCUM = "A", 2048, 1, 4
# BASELINE.md §1 (published, 4xA10 over TCP): DDP tokens/s by world size
PUBLISHED_DDP = {2: 8369.455699, 4: 12220.341464}


def env_int(name: str, default: int) -> int:
    return int(os.environ.get(name, default))


_RESULT_FD = None      # the original stdout, once quiet_stdout() has pointed fd 1 at stderr


def quiet_stdout() -> None:
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner with
    printf), so for the duration of the run file descriptor 1 is pointed at stderr and the result line alone is
    written to the saved original."""
    global _RESULT_FD
    if _RESULT_FD is None:
        sys.stdout.flush()
        _RESULT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(obj: dict) -> None:
    line = (json.dumps(obj) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(line.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_RESULT_FD, line)


# ------------------------------------------------------------------------------------------
# reference arm
# ------------------------------------------------------------------------------------------
def run_reference(a: argparse.Namespace) -> None:
    rank, local_rank, ws = env_int("RANK", 0), env_int("LOCAL_RANK", 0), env_int("WORLD_SIZE", 1)
    harness = os.path.join(ROOT, "baseline", "_ref", "benchmarking", "train_harness.py")
    if not os.path.exists(harness) and os.path.isdir("/root/reference") and rank == 0:
        subprocess.run(["bash", os.path.join(ROOT, "baseline", "install_reference.sh")], check=False,
                       capture_output=True)       # git-ignored directory: re-create it from the read-only reference
    if not os.path.exists(harness):
        if rank == 0:
            emit({"impl": "reference", "unavailable": "baseline/_ref missing: run baseline/install_reference.sh"})
        return
    if a.strategy in ("zero2", "zero3"):
        if rank == 0:
            emit({"impl": "reference", "unavailable": f"reference --strategy {a.strategy} needs DeepSpeed, which is not "
                  "installed in this image and cannot be (no package index)"})
        return
    port = env_int("MASTER_PORT", 29500) + 1 + (os.getpid() % 1 if ws == 1 else 0)
    outdir = tempfile.mkdtemp(prefix="refbench_")
    cmd = [sys.executable, "-u", harness, "--strategy", a.strategy, "--world-size", str(ws), "--rank", str(rank),
           "--local-rank", str(local_rank), "--master-addr", "127.0.0.1", "--master-port", str(port),
           "--tier", TIER, "--seq-len", str(SEQ), "--synthetic", "--steps", str(a.steps + a.warmup),
           "--warmup-steps", str(a.warmup), "--per-device-batch", str(BATCH), "--grad-accum", str(GRAD_ACCUM),
           "--results-dir", outdir]
    t0 = time.time()
    # torchrun exports TORCHELASTIC_USE_AGENT_STORE=True; with it c10d's tcp:// rendezvous makes EVERY rank a
    # store client and the reference (which hosts its own store on rank 0, TH:190-195) waits forever.
    env = {k: v for k, v in os.environ.items() if not k.startswith(("TORCHELASTIC", "TORCH_NCCL", "NCCL_ASYNC"))}
    env["OMP_NUM_THREADS"] = "8"          # torchrun pins it to 1, which makes the reference's CPU-side init crawl
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=a.ref_timeout)
    except subprocess.TimeoutExpired as te:
        if rank == 0:
            tail = ((te.stdout or b"")[-300:]).decode("utf-8", "replace").replace("\n", " | ") if te.stdout else ""
            emit({"impl": "reference", "unavailable": f"reference harness did not finish within {a.ref_timeout}s "
                  f"at world size {ws}; last output: {tail}"})
        return
    wall = time.time() - t0
    if rank != 0:
        return
    if r.returncode != 0:
        emit({"impl": "reference", "unavailable": f"reference harness exited {r.returncode}: "
              + (r.stderr.strip().splitlines() or ['?'])[-1][:200]})
        return
    from b200bench.utils.result import extract_marked

    res = extract_marked(r.stdout)
    if res is None:
        emit({"impl": "reference", "unavailable": "no result JSON between the reference's stdout markers"})
        return
    tps = float(res["tokens_per_sec"])
    emit({
        "impl": "reference", "metric": METRIC, "value": tps, "unit": "tokens/s", "n_gpus": ws,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": float(res["mean_step_time_sec"]) * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": (tps / PUBLISHED_DDP[ws]) if ws in PUBLISHED_DDP else None,
        "dtype": "fp16 autocast over fp32 weights (reference default, TH:334-376)", "data": "synthetic",
        "config": {"model": "TinyGPT tier A (236.4M)", "global_batch": BATCH * ws, "seq_len": SEQ,
                   "parallelism": f"dp{ws}", "strategy": a.strategy, "grad_accum_flag": GRAD_ACCUM,
                   "timer": "reference host perf_counter, rank 0 (TH:361-385)"},
        "e2e": {"value": tps, "unit": "tokens/s", "h2d_bytes_per_step": BATCH * SEQ * 8, "d2h_bytes_per_step": 4,
                "note": "the reference loop itself copies int64 tokens from pinned memory and reads loss.item() every step"},
        "gpu_launches": 0, "peak_vram_gb": res.get("peak_vram_gb"), "mean_loss": res.get("mean_loss"),
        "wall_s": round(wall, 1),
    })


# ------------------------------------------------------------------------------------------
# native arm
# ------------------------------------------------------------------------------------------
def run_extra(a, tier, strategy, ctx, dev, rank, local_rank, ws, steps):
    """One more (tier, strategy) measured like the headline's device-timed arm: fresh engine, warm-up over two
    accumulation cycles, `steps` timed steps between CUDA events, max over ranks.  Returns a dict for the
    `strategies` object of the result line.  The shipped strategy files are used (DeepSpeed JSON for zero2 / zero3,
    YAML for fsdp), so zero2 / zero3 accumulate over 4 micro-steps and clip like the reference's configs."""
    import torch
    import torch.distributed as dist

    from b200bench.config import StrategyConfig, get_model_config, load_deepspeed_config, load_fsdp_config
    from b200bench.engine import NativeEngine
    from b200bench.utils.clocks import ClockSampler

    model_cfg = get_model_config(tier, a.seq_len)
    if strategy in ("zero2", "zero3"):
        cfg = load_deepspeed_config(os.path.join(ROOT, "configs", "deepspeed", f"{strategy}.json"), a.batch,
                                    GRAD_ACCUM, ws)
    elif strategy == "fsdp":
        cfg = load_fsdp_config(os.path.join(ROOT, "configs", "fsdp", "fsdp_config.yaml"))
    else:
        cfg = StrategyConfig()
    torch.cuda.reset_peak_memory_stats(dev)
    eng = NativeEngine(model_cfg, strategy, cfg, ctx, dev, grad_accum=GRAD_ACCUM, dropout=a.dropout, causal=False,
                       batch=a.batch, seq_len=a.seq_len, comm=a.comm, use_graph=not a.no_graph)
    try:
        gen = torch.Generator(device="cpu").manual_seed(1234 + rank)
        toks = [torch.randint(0, model_cfg["vocab_size"], (a.batch, a.seq_len), generator=gen, dtype=torch.int32).to(dev)
                for _ in range(4)]

        def sync_all():
            torch.cuda.synchronize(dev)
            if ws > 1:
                dist.barrier()
            torch.cuda.synchronize(dev)

        n_warm = max(3, 2 * eng.accum + 2 if eng.accum > 1 else 3)
        n_warm += (-n_warm) % eng.accum
        steps += (-steps) % eng.accum                     # whole accumulation cycles
        for i in range(n_warm):
            eng.train_step(toks[i % 4], toks[i % 4])
        sync_all()
        sampler = ClockSampler(local_rank).start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            loss = eng.train_step(toks[i % 4], toks[i % 4])
        e1.record()
        sync_all()
        clocks = sampler.stop()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if ws > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
        peak = torch.cuda.max_memory_allocated(dev) + eng.extra_memory_bytes()
        out = {"tier": tier, "value": a.batch * a.seq_len * ws * steps / (ms / 1e3), "unit": "tokens/s",
               "ms_per_step": ms / steps, "steps": steps, "warmup": n_warm, "grad_accum": eng.accum,
               "peak_vram_gb": peak / 1e9, "final_loss": float(loss[0]),
               "param_gather": eng.param_gather if getattr(eng, "sharded_params", False) else None,
               "params_m": round(eng.layout.num_params() / 1e6, 1),
               "clocks": {"sm_mhz": clocks.get("sm_mhz"), "reasons": clocks.get("reasons")}}
    finally:
        eng.close()
        del eng
        torch.cuda.empty_cache()
    return out


def run_native(a: argparse.Namespace) -> None:
    import torch
    import torch.distributed as dist

    from b200bench.config import StrategyConfig, get_model_config
    from b200bench.engine import NativeEngine
    from b200bench.parallel.collectives import DistCtx
    from b200bench.utils.clocks import ClockSampler
    from b200bench.utils.data import PinnedBatchSource, SyntheticDataset

    rank, local_rank, ws = env_int("RANK", 0), env_int("LOCAL_RANK", 0), env_int("WORLD_SIZE", 1)
    if ws != a.gpus and rank == 0:
        print(f"[bench] note: --gpus {a.gpus} but WORLD_SIZE={ws}; using WORLD_SIZE", file=sys.stderr)
    dev = torch.device(f"cuda:{local_rank}")
    torch.cuda.set_device(dev)
    ctx = DistCtx(rank, ws, "nccl")
    if ws > 1:
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(42 + rank)

    model_cfg = get_model_config(a.tier, a.seq_len)
    cfg = StrategyConfig()
    eng = NativeEngine(model_cfg, a.strategy, cfg, ctx, dev, grad_accum=GRAD_ACCUM, dropout=a.dropout,
                       causal=False, batch=a.batch, seq_len=a.seq_len, comm=a.comm, use_graph=not a.no_graph)
    ds = SyntheticDataset(model_cfg["vocab_size"], a.seq_len, size=1000, seed=42, verbose=False)
    src = PinnedBatchSource(ds, a.batch, rank, ws, dev)

    def sync_all() -> None:
        torch.cuda.synchronize(dev)
        if ws > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- warm-up (untimed): builds tensor maps, captures the CUDA graph -----------------------
    resident = [src.next() for _ in range(8)]            # device-resident batches for the kernel-only arm
    # warm-up must cover every micro-step variant of an accumulation cycle twice (eager, then graph capture)
    n_warm = max(3, a.warmup, 2 * eng.accum + 2 if eng.accum > 1 else 0)
    n_warm += (-n_warm) % eng.accum          # stop on an accumulation boundary
    for i in range(n_warm):
        eng.train_step(resident[i % 8], resident[i % 8])
    sync_all()

    # ---- device-timed arm -------------------------------------------------------------------------
    sampler = ClockSampler(local_rank).start()
    launches0 = eng.ops.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    e0.record()
    for i in range(a.steps):
        eng.train_step(resident[i % 8], resident[i % 8])
    e1.record()
    sync_all()
    dev_ms = e0.elapsed_time(e1)
    eager_launches = eng.ops.launches - launches0

    # ---- end-to-end arm: pinned H2D of the tokens + D2H of the loss every step ---------------------
    loss_host = torch.zeros(a.steps, dtype=torch.float32).pin_memory()
    sync_all()
    t0 = time.perf_counter()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    for i in range(a.steps):
        tok = src.next()                                  # cudaMemcpyAsync from pinned host memory
        loss = eng.train_step(tok, tok)
        loss_host[i].copy_(loss[0], non_blocking=True)    # device -> pinned host
    e3.record()
    sync_all()
    e2e_wall_ms = (time.perf_counter() - t0) * 1e3
    e2e_ms = max(e2.elapsed_time(e3), e2e_wall_ms)
    clocks = sampler.stop()

    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    if ws > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)         # max over ranks
    dev_ms, e2e_ms = float(t[0]), float(t[1])
    tokens = a.batch * a.seq_len * ws * a.steps
    value = tokens / (dev_ms / 1e3)
    e2e_value = tokens / (e2e_ms / 1e3)
    peak = torch.cuda.max_memory_allocated(dev) + eng.extra_memory_bytes()
    result = None
    if rank == 0:
        per_step = eng.kernel_launches_per_step()
        result = ({
            "metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": ws, "steps": a.steps,
            "warmup": n_warm, "ms_per_step": dev_ms / a.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": (value / PUBLISHED_DDP[ws]) if ws in PUBLISHED_DDP else None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"model": f"TinyGPT tier {a.tier} ({eng.layout.num_params()/1e6:.1f}M)",
                       "global_batch": a.batch * ws, "seq_len": a.seq_len, "parallelism": f"dp{ws}",
                       "strategy": a.strategy, "comm": a.comm if ws > 1 else "none", "dropout": a.dropout,
                       "attention": "bidirectional (reference passes no mask)",
                       "optimizer_step_every_iteration": eng.accum == 1,
                       "l2": "no flush: each step streams >4 GB (weights, grads, fp32 Adam state, activations) >> 126 MB L2",
                       "cuda_graph": not a.no_graph},
            "clocks": {"sm_mhz": clocks.get("sm_mhz"), "sm_max_mhz": clocks.get("sm_max_mhz"),
                       "reasons": clocks.get("reasons"), "power_w_max": clocks.get("power_w_max")},
            "e2e": {"value": e2e_value, "unit": "tokens/s", "ms_per_step": e2e_ms / a.steps,
                    "h2d_bytes_per_step": src.bytes_per_step, "d2h_bytes_per_step": 4},
            "gpu_launches": per_step * a.steps, "gpu_launches_per_step": per_step,
            "peak_vram_gb": peak / 1e9, "final_loss": float(loss_host[-1]), "impl": "native",
        })
    if a.diag:
        # host-side enqueue cost of a step (no sync inside) and graph-replayed phase times, every rank
        sync_all()
        h0 = time.perf_counter()
        for i in range(a.steps):
            eng.train_step(resident[i % 8], resident[i % 8])
        host_ms = (time.perf_counter() - h0) * 1e3 / a.steps
        sync_all()
        phases = eng.profile_breakdown() if eng.accum == 1 else {}
        print(f"[diag] rank {rank}: host enqueue {host_ms:.3f} ms/step, phases {phases}", file=sys.stderr, flush=True)
    eng.close()
    del eng
    torch.cuda.empty_cache()
    # ---- the other strategies of the reference's matrix (scripts/run_all_benchmarks.sh:32-52 there), same box,
    # same launch, in the SAME result line under "strategies"; the headline above stays ddp.  A failing extra is
    # recorded as {"error": ...} and never costs the headline.
    extras = {}
    if rank == 0 and ws > 1 and a.extras:
        # torchrun sends SIGTERM to the surviving ranks when one rank of an extra dies, and this rank may sit in a
        # blocking CUDA / NCCL call at that moment: the C-level handler writes to a wake-up pipe, a watcher thread
        # prints the headline (with the extras finished so far) and exits.
        import signal
        import threading
        r_fd, w_fd = os.pipe()
        os.set_blocking(w_fd, False)
        signal.signal(signal.SIGTERM, lambda signum, frame: None)
        signal.set_wakeup_fd(w_fd, warn_on_full_buffer=False)

        def _watch():
            os.read(r_fd, 1)
            extras["aborted"] = "signal from the launcher (another rank exited)"
            result["strategies"] = extras
            emit(result)
            os._exit(0)
        threading.Thread(target=_watch, daemon=True).start()
    try:
        if a.extras and a.strategy == "ddp":
            todo = [(a.tier, "zero2"), (a.tier, "fsdp"), (a.tier, "zero3")]
            if ws >= 2 and os.environ.get("B200BENCH_BENCH_7B", "1") != "0":
                todo.append(("7B", "zero3"))
            for tier, strat in todo:
                key = strat if tier == a.tier else f"{strat}_{tier}"
                ok = 1
                try:
                    extras[key] = run_extra(a, tier, strat, ctx, dev, rank, local_rank, ws, min(a.steps, 12))
                except Exception as e:      # noqa: BLE001
                    extras[key] = {"error": f"{type(e).__name__}: {str(e)[:200]}"}
                    ok = 0
                if ws > 1:              # ranks agree whether to go on (a rank that failed may have left a collective)
                    flag = torch.tensor([ok], dtype=torch.int32, device=dev)
                    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
                    ok = int(flag.item())
                if not ok:
                    extras.setdefault(key, {})["aborted_remaining"] = True
                    break
    except BaseException as e:          # noqa: BLE001  (a dead CUDA context after a failed extra: keep the headline)
        extras["aborted"] = f"{type(e).__name__}: {str(e)[:200]}"
    finally:
        if rank == 0:
            result["strategies"] = extras
            emit(result)
    if ws > 1:
        dist.destroy_process_group()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["native", "reference"], default="native")
    ap.add_argument("--strategy", choices=["ddp", "fsdp", "zero2", "zero3"], default="ddp")
    ap.add_argument("--comm", choices=["fused", "nccl"], default="fused")
    ap.add_argument("--tier", default=TIER)
    ap.add_argument("--seq-len", type=int, default=SEQ)
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--dropout", type=float, default=0.1)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--diag", action="store_true", help="after the timed arms, print a per-phase breakdown to stderr")
    ap.add_argument("--no-extras", dest="extras", action="store_false",
                    help="headline only: skip the zero2 / fsdp / zero3 (and 7B zero3) runs reported under 'strategies'")
    ap.add_argument("--ref-timeout", type=float, default=180.0, help="seconds before the reference arm is abandoned")
    a = ap.parse_args()
    quiet_stdout()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_native(a)


if __name__ == "__main__":
    main()
