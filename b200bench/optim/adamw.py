"""Flat AdamW over one contiguous (possibly partitioned) parameter range.

State is three fp32 vectors (master weights, exp_avg, exp_avg_sq) over the rank's slice of the
flat layout; the working weights (bf16 on GPU) are rewritten from the master after every update.
Semantics = ``torch.optim.AdamW`` (decoupled weight decay, bias correction) as used by the
reference at ``benchmarking/train_harness.py:329`` and DeepSpeed's FusedAdam in AdamW mode
configured by ``configs/deepspeed/zero2.json:27-35``; ``WarmupLR`` follows
``zero2.json:37-44`` (linear 0 -> lr over ``warmup_num_steps``; DeepSpeed's default warmup type is
log, the reference sets no ``warmup_type`` so the log ramp is reproduced when ``log=True``).

The torch implementation here is the oracle and the CPU path; on a GPU the same update runs as
one fused kernel (``adamw_kernel`` in ``csrc/elementwise.cu``; fused with the gradient reduce-scatter as K15 in
``csrc/comm.cu``) through ``b200bench.ops``.
"""
from __future__ import annotations

import math
from typing import Optional

import torch

from ..config import OptimConfig


def warmup_lr(step: int, cfg: OptimConfig, log: bool = True) -> float:
    """Learning rate for optimizer step ``step`` (0-based)."""
    if cfg.warmup_steps <= 0:
        return cfg.lr
    if step >= cfg.warmup_steps:
        return cfg.lr
    if log:
        gamma = math.log(step + 1) / math.log(max(2, cfg.warmup_steps))
    else:
        gamma = step / cfg.warmup_steps
    gamma = min(1.0, gamma)
    return cfg.warmup_min_lr + (cfg.lr - cfg.warmup_min_lr) * gamma


class FlatAdamW:
    def __init__(self, master_init: torch.Tensor, cfg: OptimConfig):
        """``master_init``: 1-D tensor (any float dtype) holding this rank's slice of the weights."""
        self.cfg = cfg
        self.master = master_init.detach().to(torch.float32).clone()
        self.exp_avg = torch.zeros_like(self.master)
        self.exp_avg_sq = torch.zeros_like(self.master)
        self.step_count = 0

    def current_lr(self) -> float:
        return warmup_lr(self.step_count, self.cfg)

    @torch.no_grad()
    def step(self, grad: torch.Tensor, out_weights: torch.Tensor, grad_scale: float = 1.0,
             native: Optional[object] = None) -> None:
        """One update.  ``grad``: this rank's slice of the (already reduced) gradient;
        ``out_weights``: the slice of the working-weight buffer to rewrite; ``grad_scale``
        multiplies the gradient first (1/world_size, 1/grad_accum, clip coefficient)."""
        c = self.cfg
        lr = self.current_lr()
        self.step_count += 1
        t = self.step_count
        bc1 = 1.0 - c.beta1 ** t
        bc2 = 1.0 - c.beta2 ** t
        if native is not None and grad.is_cuda:
            native.adamw(self.master, self.exp_avg, self.exp_avg_sq, grad, out_weights, lr, c.beta1,
                         c.beta2, c.eps, c.weight_decay, bc1, bc2, grad_scale)
            return
        g = grad.to(torch.float32)
        if grad_scale != 1.0:
            g = g * grad_scale
        self.master.mul_(1.0 - lr * c.weight_decay)
        self.exp_avg.mul_(c.beta1).add_(g, alpha=1.0 - c.beta1)
        self.exp_avg_sq.mul_(c.beta2).addcmul_(g, g, value=1.0 - c.beta2)
        denom = (self.exp_avg_sq.sqrt() / math.sqrt(bc2)).add_(c.eps)
        self.master.addcdiv_(self.exp_avg, denom, value=-lr / bc1)
        out_weights.copy_(self.master)

    def state_dict(self) -> dict:
        return dict(master=self.master, exp_avg=self.exp_avg, exp_avg_sq=self.exp_avg_sq,
                    step=self.step_count)

    def load_state_dict(self, sd: dict) -> None:
        self.master.copy_(sd["master"])
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        self.step_count = int(sd["step"])
