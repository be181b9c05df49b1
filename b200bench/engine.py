"""Native training engine: TinyGPT forward / backward / optimizer as an explicit schedule of
sm_100a kernels over flat buffers — no autograd, no ATen or cuBLAS on the hot path.

What the reference does through ``nn.Module`` + autograd + ``torch.optim`` / DeepSpeed
(``benchmarking/train_harness.py:80-131,364-382``) is written out here op by op (SURVEY.md §2.6
K1-K12, §3.4).  Parameters, gradients and optimizer state live in the flat layout of
``models/layout.py``; activations are pre-allocated once for the fixed (batch, seq_len) of the
run, which makes the whole micro-step a static launch sequence that is captured into a CUDA
graph and replayed (launch-bound at tier A, batch 1: ~600 kernels per step).

Strategies differ only in what happens at the accumulation boundary:
    ddp    all-reduce(grads) -> AdamW on the full flat buffer
    zero2  reduce-scatter(grads) -> clip -> AdamW on the local shard -> all-gather(weights)
    fsdp / zero3   same arithmetic as zero2 but the bf16 weights are owned shard-wise and gathered
           before use (per step here; per block inside the consuming GEMM with the fused path)
``comm="fused"`` uses the peer-memory kernels of ``csrc/comm_*.cu``; ``comm="nccl"`` keeps the same
schedule on ``torch.distributed`` collectives for A/B comparison.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import torch

from . import ops as _ops
from .config import StrategyConfig
from .models.layout import ParamLayout, build_layout
from .optim.adamw import warmup_lr
from .parallel.collectives import DistCtx
from .parallel.torch_engine import SHARDED_OPT, SHARDED_PARAMS, effective_grad_accum

BF16 = torch.bfloat16


def _mix(*xs: int) -> int:
    h = 0x9E3779B9
    for x in xs:
        h = (h ^ (x + 0x7F4A7C15 + ((h << 6) & 0xFFFFFFFF) + (h >> 2))) & 0xFFFFFFFF
    return h


class NativeEngine:
    dtype = BF16

    def __init__(self, model_cfg: Dict[str, int], strategy: str, cfg: StrategyConfig, ctx: DistCtx,
                 device: torch.device, grad_accum: int = 1, dropout: float = 0.1, causal: bool = False,
                 batch: int = 1, seq_len: int = 2048, comm: str = "fused", use_graph: bool = True,
                 seed: int = 42):
        self.ops = _ops.load()
        self.strategy, self.cfg, self.ctx, self.device = strategy, cfg, ctx, device
        self.accum = effective_grad_accum(strategy, grad_accum, cfg.grad_accum_mode)
        self.dropout, self.causal = float(dropout), bool(causal)
        self.B, self.T = int(batch), int(seq_len)
        self.V, self.d = model_cfg["vocab_size"], model_cfg["n_embd"]
        self.H, self.L = model_cfg["n_head"], model_cfg["n_layer"]
        self.hd = self.d // self.H
        if self.hd != 64:
            raise NotImplementedError("native attention kernels are built for head_dim 64")
        if self.T > model_cfg["block_size"]:
            raise ValueError(f"Sequence {self.T} exceeds block size {model_cfg['block_size']}")
        self.M = self.B * self.T
        ws, rank = ctx.world_size, ctx.rank
        self.layout: ParamLayout = build_layout(world_size=ws, **model_cfg)
        Lt = self.layout
        dev = device

        # ---- parameters / gradients / optimizer state ---------------------------------------
        self.comm_mode = comm if ws > 1 else "none"
        self.comm = None
        if ws > 1 and comm == "fused":
            from .parallel.peer import PeerComm

            self.comm = PeerComm(ctx, device)
            # symmetric buffers: every rank can address every peer's weights and gradients
            self.params_pb = self.comm.alloc("params", Lt.total, BF16)
            self.grads_pb = self.comm.alloc("grads", Lt.total, BF16)
            self.params, self.grads = self.params_pb.tensor, self.grads_pb.tensor
        else:
            self.params = torch.zeros(Lt.total, dtype=BF16, device=dev)
            self.grads = torch.zeros(Lt.total, dtype=BF16, device=dev)
        self._init_params(seed)
        if ws > 1:
            ctx.broadcast(self.params, 0)
        self.lo, self.hi = Lt.shard_range(rank) if strategy in SHARDED_OPT else (0, Lt.total)
        n_opt = self.hi - self.lo
        self.master = self.params[self.lo:self.hi].float()
        self.exp_avg = torch.zeros(n_opt, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(n_opt, dtype=torch.float32, device=dev)
        self.grad_shard = torch.zeros(n_opt, dtype=BF16, device=dev) if strategy in SHARDED_OPT else None
        self.hp = torch.zeros(16, dtype=torch.float32, device=dev)
        self.gnorm_sq = torch.zeros(1, dtype=torch.float32, device=dev)
        self.opt_step = 0
        self.micro = 0
        self.params_stale = False
        self.P: Dict[str, torch.Tensor] = {s.name: self.params[s.offset:s.offset + s.numel].view(s.shape)
                                           for s in Lt}
        self.G: Dict[str, torch.Tensor] = {s.name: self.grads[s.offset:s.offset + s.numel].view(s.shape)
                                           for s in Lt}

        # ---- activations (allocated once) -----------------------------------------------------
        M, d, V, L = self.M, self.d, self.V, self.L
        e = lambda *shape, dt=BF16: torch.empty(*shape, dtype=dt, device=dev)  # noqa: E731
        self.tokens = torch.zeros(self.B, self.T, dtype=torch.int32, device=dev)
        self.targets = torch.zeros(self.B, self.T, dtype=torch.int32, device=dev)
        self.x = [e(M, d) for _ in range(L + 1)]
        self.ln1 = [e(M, d) for _ in range(L)]
        self.ln2 = [e(M, d) for _ in range(L)]
        self.stats = [e(4, M, dt=torch.float32) for _ in range(L)]      # mean1, rstd1, mean2, rstd2
        self.qkv = [e(M, 3 * d) for _ in range(L)]
        self.att = [e(M, d) for _ in range(L)]
        self.lse = [e(self.B * self.H * self.T, dt=torch.float32) for _ in range(L)]
        self.xmid = [e(M, d) for _ in range(L)]
        self.pre = [e(M, 4 * d) for _ in range(L)]
        self.act = [e(M, 4 * d) for _ in range(L)]
        self.lnf = e(M, d)
        self.statf = e(2, M, dt=torch.float32)
        self.logits = e(M, V)
        self.loss_rows = e(M, dt=torch.float32)
        self.n_valid = torch.zeros(1, dtype=torch.int32, device=dev)
        self.loss = torch.zeros(1, dtype=torch.float32, device=dev)
        # backward temporaries, shared by all layers
        self.dres = e(M, d)
        self.dmid = e(M, d)
        self.dln = e(M, d)
        self.dy_drop = e(M, d)
        self.dbig = e(M, 4 * d)
        self.dqkv = e(M, 3 * d)
        self.datt = e(M, d)
        self.dq_acc = e(M, d, dt=torch.float32)
        self.attD = e(self.B * self.H * self.T, dt=torch.float32)
        sbytes = max(self.ops.colreduce_scratch_bytes(M, 4 * d, 1), self.ops.colreduce_scratch_bytes(M, d, 2),
                     self.ops.colreduce_scratch_bytes(M, 3 * d, 1))
        self.scratch = torch.zeros(sbytes // 4 + 64, dtype=torch.float32, device=dev)   # counters start at 0
        self.seed_base = seed
        self.step_seed = torch.full((1,), seed, dtype=torch.int32, device=dev)
        self.ops.step_seed_register(self.step_seed)

        self.use_graph = use_graph
        self._graphs: Dict[tuple, torch.cuda.CUDAGraph] = {}
        self._warm = 0
        self._launches_per_step: Optional[int] = None
        # DDP gradient buckets = one per transformer block + (embeddings, final LN)   [SURVEY N3]
        l0 = Lt.layer_range(0)[0]
        l1 = Lt.layer_range(self.L - 1)[1]
        self._bucket_head = (0, l0)
        self._bucket_tail = (l1, Lt.total)

    # ---------------------------------------------------------------------------------------
    def _init_params(self, seed: int) -> None:
        """Reference initialisation (TH:69-78): N(0, 0.02) for Linear / Embedding weights, zero
        biases, LayerNorm (1, 0); MultiheadAttention.in_proj keeps xavier_uniform (not an nn.Linear)."""
        g = torch.Generator(device=self.device)
        g.manual_seed(seed)
        for s in self.layout:
            view = self.params[s.offset:s.offset + s.numel].view(s.shape)
            n = s.name
            if n.endswith("ln_1.weight") or n.endswith("ln_2.weight") or n.endswith("ln_f.weight"):
                view.fill_(1.0)
            elif n.endswith("bias"):
                view.zero_()
            elif n.endswith("in_proj_weight"):
                bound = math.sqrt(6.0 / (s.shape[0] + s.shape[1]))
                view.copy_((torch.rand(s.shape, generator=g, device=self.device) * 2 - 1) * bound)
            else:
                view.copy_(torch.randn(s.shape, generator=g, device=self.device) * 0.02)

    # ---------------------------------------------------------------------------------------
    def _seed(self, layer: int, site: int) -> int:
        # The per-step component lives in device memory (graph-replay safe): see ops.set_step_seed.
        return _mix(self.seed_base, layer, site)

    def _forward(self) -> None:
        o, P, p = self.ops, self.P, self.dropout
        B, T, d, H, hd = self.B, self.T, self.d, self.H, self.hd
        o.embedding_fwd(self.tokens, P["transformer.wte.weight"], P["transformer.wpe.weight"], self.x[0],
                        B, T, d, p, self._seed(-1, 0))
        for l in range(self.L):
            pre = f"transformer.h.{l}."
            st = self.stats[l]
            o.layernorm_fwd(self.x[l], P[pre + "ln_1.weight"], P[pre + "ln_1.bias"], self.ln1[l], st[0], st[1])
            o.linear_fwd(self.ln1[l], P[pre + "attn.in_proj_weight"], self.qkv[l], bias=P[pre + "attn.in_proj_bias"])
            o.attention_fwd(self.qkv[l], self.att[l], self.lse[l], B, T, H, hd, self.causal, p, self._seed(l, 1))
            o.linear_fwd(self.att[l], P[pre + "attn.out_proj.weight"], self.xmid[l],
                         bias=P[pre + "attn.out_proj.bias"], epi=_ops.EPI_BIAS_RESID, aux=self.x[l])
            o.layernorm_fwd(self.xmid[l], P[pre + "ln_2.weight"], P[pre + "ln_2.bias"], self.ln2[l], st[2], st[3])
            o.linear_fwd(self.ln2[l], P[pre + "mlp.0.weight"], self.pre[l], bias=P[pre + "mlp.0.bias"],
                         epi=_ops.EPI_BIAS_GELU, out2=self.act[l])
            o.linear_fwd(self.act[l], P[pre + "mlp.2.weight"], self.x[l + 1], bias=P[pre + "mlp.2.bias"],
                         epi=_ops.EPI_BIAS_RESID, aux=self.xmid[l], drop_p=p, seed=self._seed(l, 2))
        o.layernorm_fwd(self.x[self.L], P["transformer.ln_f.weight"], P["transformer.ln_f.bias"], self.lnf,
                        self.statf[0], self.statf[1])
        o.linear_fwd(self.lnf, P["transformer.wte.weight"], self.logits)

    def _loss_and_backward(self, first_micro: bool, last_micro: bool) -> None:
        o, P, G, p = self.ops, self.P, self.G, self.dropout
        reduce_now = self.comm is not None and self.strategy == "ddp" and last_micro
        B, T, d, H, hd, M, V = self.B, self.T, self.d, self.H, self.hd, self.M, self.V
        acc = not first_micro
        o.cross_entropy(self.logits, self.targets, self.loss_rows, self.n_valid, self.loss, M, V,
                        grad_scale=1.0 / self.accum, write_grad=True)
        wte = P["transformer.wte.weight"]
        # lm_head: dlnf = dlogits @ wte ; dwte (+)= dlogits^T @ lnf      (tied weight)
        o.linear_dgrad(self.logits, wte, self.dln)
        o.linear_wgrad(self.logits, self.lnf, G["transformer.wte.weight"], accumulate=acc)
        o.layernorm_bwd(self.dln, self.x[self.L], P["transformer.ln_f.weight"], self.statf[0], self.statf[1],
                        self.dres)
        o.layernorm_bwd_params(self.dln, self.x[self.L], self.statf[0], self.statf[1],
                               G["transformer.ln_f.weight"], G["transformer.ln_f.bias"], self.scratch, acc)
        for l in reversed(range(self.L)):
            pre = f"transformer.h.{l}."
            st = self.stats[l]
            # ---- MLP ----
            dy = self.dres
            if p > 0:
                o.dropout_apply(self.dres, self.dy_drop, M, d, p, self._seed(l, 2))
                dy = self.dy_drop
            o.colsum(dy, G[pre + "mlp.2.bias"], self.scratch, acc)
            o.linear_wgrad(dy, self.act[l], G[pre + "mlp.2.weight"], accumulate=acc)
            o.linear_dgrad(dy, P[pre + "mlp.2.weight"], self.dbig, epi=_ops.EPI_DGELU, aux=self.pre[l])
            o.colsum(self.dbig, G[pre + "mlp.0.bias"], self.scratch, acc)
            o.linear_wgrad(self.dbig, self.ln2[l], G[pre + "mlp.0.weight"], accumulate=acc)
            o.linear_dgrad(self.dbig, P[pre + "mlp.0.weight"], self.dln)
            o.layernorm_bwd(self.dln, self.xmid[l], P[pre + "ln_2.weight"], st[2], st[3], self.dmid, dres=self.dres)
            o.layernorm_bwd_params(self.dln, self.xmid[l], st[2], st[3], G[pre + "ln_2.weight"],
                                   G[pre + "ln_2.bias"], self.scratch, acc)
            # ---- attention ----
            o.colsum(self.dmid, G[pre + "attn.out_proj.bias"], self.scratch, acc)
            o.linear_wgrad(self.dmid, self.att[l], G[pre + "attn.out_proj.weight"], accumulate=acc)
            o.linear_dgrad(self.dmid, P[pre + "attn.out_proj.weight"], self.datt)
            o.attention_bwd(self.qkv[l], self.att[l], self.datt, self.lse[l], self.dqkv, self.attD, self.dq_acc,
                            B, T, H, hd, self.causal, p, self._seed(l, 1))
            o.colsum(self.dqkv, G[pre + "attn.in_proj_bias"], self.scratch, acc)
            o.linear_wgrad(self.dqkv, self.ln1[l], G[pre + "attn.in_proj_weight"], accumulate=acc)
            o.linear_dgrad(self.dqkv, P[pre + "attn.in_proj_weight"], self.dln)
            o.layernorm_bwd(self.dln, self.x[l], P[pre + "ln_1.weight"], st[0], st[1], self.dres, dres=self.dmid)
            o.layernorm_bwd_params(self.dln, self.x[l], st[0], st[1], G[pre + "ln_1.weight"],
                                   G[pre + "ln_1.bias"], self.scratch, acc)
            if reduce_now:
                # K13: this block's gradients are final -> all-reduce them on the comm stream while
                # the compute stream continues with block l-1 (overlap with backward)
                a, b = self.layout.layer_range(l)
                self.comm.fork()
                self.comm.allreduce(self.grads_pb, a, b - a, 1.0 / self.ctx.world_size)
        o.embedding_bwd(self.tokens, self.dres, G["transformer.wte.weight"], G["transformer.wpe.weight"],
                        B, T, d, p, self._seed(-1, 0), accumulate_wpe=acc)
        if reduce_now:
            self.comm.fork()
            for a, b in (self._bucket_tail, self._bucket_head):
                self.comm.allreduce(self.grads_pb, a, b - a, 1.0 / self.ctx.world_size)
            self.comm.join()

    # ---------------------------------------------------------------------------------------
    def _micro_step(self, first_micro: bool, last_micro: bool) -> None:
        if self.dropout > 0:
            self.ops.step_seed_advance()      # new dropout masks every micro-step, also under graph replay
        self._forward()
        self._loss_and_backward(first_micro, last_micro)

    def _run_micro(self, first_micro: bool, last_micro: bool) -> None:
        if not self.use_graph:
            self._micro_step(first_micro, last_micro)
            return
        key = (first_micro, last_micro)
        g = self._graphs.get(key)
        if g is None:
            if self._warm < 2:              # eager warm-up: function attributes, tensor maps, lazy init
                self._warm += 1
                self._micro_step(first_micro, last_micro)
                return
            before = self.ops.launches
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._micro_step(first_micro, last_micro)
            self._graph_launches = self.ops.launches - before
            self._graphs[key] = g
        g.replay()

    def _set_hp(self, grad_scale: float) -> None:
        c = self.cfg.optim
        lr = warmup_lr(self.opt_step, c)
        self.opt_step += 1
        t = self.opt_step
        bc1 = 1.0 - c.beta1 ** t
        bc2 = 1.0 - c.beta2 ** t
        self.ops.adamw_set_hp(self.hp, lr, c.beta1, c.beta2, c.eps, c.weight_decay, 1.0 / bc1,
                              1.0 / math.sqrt(bc2), grad_scale, c.grad_clip)

    def _optimizer_boundary(self) -> None:
        ctx, ws = self.ctx, self.ctx.world_size
        scale = 1.0 / ws
        clip = self.cfg.optim.grad_clip > 0
        if self.comm is not None:
            if self.strategy == "ddp":          # gradients were all-reduced and scaled during backward (K13)
                self._set_hp(1.0)
                gn = None
                if clip:
                    self.gnorm_sq.zero_()
                    self.ops.sumsq(self.grads, self.gnorm_sq, 1.0)
                    gn = self.gnorm_sq
                self.ops.adamw_flat(self.master, self.exp_avg, self.exp_avg_sq, self.grads, self.params, self.hp, gn)
                return
            # K15: pull + reduce my partition, clip, AdamW, store updated weights (to all ranks for ZeRO-2)
            self._set_hp(scale)
            self.comm.rs_adamw(self.grads_pb, self.params_pb, self.lo, self.hi - self.lo, self.master,
                               self.exp_avg, self.exp_avg_sq, self.grad_shard, self.hp,
                               push_all=(self.strategy == "zero2"))
            if self.strategy in SHARDED_PARAMS:
                ctx.all_gather(self.params[self.lo:self.hi], self.params)   # until K14 gathers per GEMM
            return
        if self.strategy == "ddp":
            if ws > 1:
                cap = max(1, self.cfg.ddp_bucket_cap_mb * 1024 * 1024 // 2)
                for a, b in self.layout.buckets(cap):
                    ctx.all_reduce_sum(self.grads[a:b])
            g = self.grads
        else:
            ctx.reduce_scatter_sum(self.grads, self.grad_shard)
            g = self.grad_shard
        self._set_hp(scale)
        gn = None
        if clip:
            self.gnorm_sq.zero_()
            self.ops.sumsq(g, self.gnorm_sq, scale)
            if self.strategy != "ddp":
                ctx.all_reduce_sum(self.gnorm_sq)
            gn = self.gnorm_sq
        self.ops.adamw_flat(self.master, self.exp_avg, self.exp_avg_sq, g, self.params[self.lo:self.hi], self.hp, gn)
        if self.strategy in SHARDED_OPT and ws > 1:
            # zero2: all-gather updated partitions (N7); fsdp/zero3: gather before the next forward
            # (N4/N8) — with NCCL both are the same collective, issued here.
            ctx.all_gather(self.params[self.lo:self.hi], self.params)

    def train_step(self, tokens: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
        self.tokens.copy_(tokens, non_blocking=True)
        if targets is tokens:
            self.targets.copy_(self.tokens, non_blocking=True)
        else:
            self.targets.copy_(targets, non_blocking=True)
        first = (self.micro % self.accum) == 0
        last = ((self.micro + 1) % self.accum) == 0
        before = self.ops.launches
        self._run_micro(first, last)
        self.micro += 1
        if self.micro % self.accum == 0:
            self._optimizer_boundary()
        eager = self.ops.launches - before
        self._launches_per_step = eager + (getattr(self, "_graph_launches", 0) if self._graphs else 0)
        return self.loss

    # ---------------------------------------------------------------------------------------
    def state_dict(self) -> Dict[str, torch.Tensor]:
        sd = {s.name: self.params[s.offset:s.offset + s.numel].view(s.shape).detach().clone().float()
              for s in self.layout}
        sd["lm_head.weight"] = sd["transformer.wte.weight"]
        return sd

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        with torch.no_grad():
            for s in self.layout:
                self.params[s.offset:s.offset + s.numel].view(s.shape).copy_(sd[s.name])
            self.master.copy_(self.params[self.lo:self.hi])
            self.exp_avg.zero_()
            self.exp_avg_sq.zero_()

    def extra_memory_bytes(self) -> int:
        return self.comm.extra_memory_bytes() if self.comm is not None else 0

    def kernel_launches_per_step(self) -> int:
        return int(self._launches_per_step or 0)

    def close(self) -> None:
        self._graphs.clear()
        if self.comm is not None:
            self.comm.close()
