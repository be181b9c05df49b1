"""Thin layer over ``torch.distributed`` used by the torch-path strategies and for rendezvous.

The reference only ever calls ``init_process_group('nccl', tcp://...)``, ``barrier`` and
``destroy_process_group`` itself (``benchmarking/train_harness.py:186-204,397``); every other
collective is implicit in DDP/FSDP/DeepSpeed.  Here the collectives are explicit so that the
native path can replace each one with a fused peer-memory kernel (``csrc/comm_*.cu``), and the
backend is selectable (``gloo`` for the CPU plumbing tests).
"""
from __future__ import annotations

import datetime
from typing import Optional

import torch
import torch.distributed as dist


class DistCtx:
    def __init__(self, rank: int = 0, world_size: int = 1, backend: Optional[str] = None):
        self.rank, self.world_size, self.backend = rank, world_size, backend
        self.owns_pg = False

    # ---- lifecycle (TH:186-204) ---------------------------------------------------------
    @classmethod
    def setup(cls, rank: int, world_size: int, master_addr: str, master_port: int, device: torch.device,
              backend: Optional[str] = None, timeout_s: int = 600) -> "DistCtx":
        if backend is None:
            backend = "nccl" if device.type == "cuda" else "gloo"
        ctx = cls(rank, world_size, backend)
        if world_size > 1:
            if not dist.is_initialized():
                kw = {}
                if backend == "nccl":
                    kw["device_id"] = device
                dist.init_process_group(backend=backend, init_method=f"tcp://{master_addr}:{master_port}",
                                        world_size=world_size, rank=rank,
                                        timeout=datetime.timedelta(seconds=timeout_s), **kw)
                ctx.owns_pg = True
            print(f"[Rank {rank}/{world_size}] Distributed initialized ({backend})")
        else:
            print("Single-GPU mode (no distributed)")
        return ctx

    def cleanup(self) -> None:
        if self.owns_pg and dist.is_initialized():
            dist.destroy_process_group()

    @property
    def active(self) -> bool:
        return self.world_size > 1 and dist.is_initialized()

    # ---- collectives --------------------------------------------------------------------
    def barrier(self) -> None:
        if self.active:
            dist.barrier()

    def broadcast(self, t: torch.Tensor, src: int = 0) -> None:
        if self.active:
            dist.broadcast(t, src)

    def all_reduce_sum(self, t: torch.Tensor) -> None:
        if self.active:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)

    def all_reduce_max(self, t: torch.Tensor) -> None:
        if self.active:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)

    def reduce_scatter_sum(self, full: torch.Tensor, out_shard: torch.Tensor) -> None:
        """``out_shard = sum_r full_r[rank*S:(rank+1)*S]``."""
        s = out_shard.numel()
        if not self.active:
            out_shard.copy_(full[:s])
            return
        if self.backend == "gloo":  # gloo has no reduce_scatter
            dist.all_reduce(full, op=dist.ReduceOp.SUM)
            out_shard.copy_(full[self.rank * s:(self.rank + 1) * s])
        else:
            dist.reduce_scatter_tensor(out_shard, full, op=dist.ReduceOp.SUM)

    def all_gather(self, shard: torch.Tensor, full: torch.Tensor) -> None:
        """``full[r*S:(r+1)*S] = shard_r`` (``shard`` may alias its own slot in ``full``)."""
        if not self.active:
            if shard.data_ptr() != full.data_ptr():
                full[:shard.numel()].copy_(shard)
            return
        if self.backend == "gloo":
            parts = [torch.empty_like(shard) for _ in range(self.world_size)]
            dist.all_gather(parts, shard.clone())
            s = shard.numel()
            for r, p in enumerate(parts):
                full[r * s:(r + 1) * s].copy_(p)
        else:
            dist.all_gather_into_tensor(full, shard)

    def max_float(self, x: float, device: torch.device) -> float:
        if not self.active:
            return x
        t = torch.tensor([x], dtype=torch.float64 if device.type == "cpu" else torch.float32, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
