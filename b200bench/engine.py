"""Native training engine: TinyGPT forward / backward / optimizer as an explicit schedule of
sm_100a kernels over flat buffers — no autograd, no ATen or cuBLAS on the hot path.

What the reference does through ``nn.Module`` + autograd + ``torch.optim`` / DeepSpeed
(``benchmarking/train_harness.py:80-131,364-382``) is written out here op by op (SURVEY.md §2.6
K1-K12, §3.4).  Parameters, gradients and optimizer state live in the flat layout of
``models/layout.py``; activations are pre-allocated once for the fixed (batch, seq_len) of the
run, which makes the whole micro-step a static launch sequence that is captured into a CUDA
graph and replayed (launch-bound at tier A, batch 1: ~500 kernels per step).

Strategies and their communication (``--comm fused``, kernels in ``csrc/comm.cu`` and
``csrc/gemm_sm100.cu``; ``--comm nccl`` keeps the same schedule on NCCL collectives for A/B runs):

ddp     full replica; K13 all-reduce(+1/N scale) per transformer block on a side stream while
        backward continues; AdamW over the whole flat buffer                       [TH:217-222]
zero2   weights replicated, gradients + optimizer state partitioned; K15 = pull-reduce of the
        partition + global-norm clip + AdamW + bf16 weight push to every peer, one kernel
                                                            [configs/deepspeed/zero2.json:10-44]
fsdp /  weights, gradients and optimizer state partitioned (matrix region of the layout).  Weight
zero3   all-gather, param_gather="gemm" (K14): no gather pass at all - the TMA producer of every consuming
        GEMM (forward and dgrad) fetches each B tile straight from the owning rank's shard over NVLink and
        multicasts it to the other CTAs of its thread-block cluster; "stream": one peer-gather kernel per
        transformer block on the comm stream into local staging, one block ahead of the compute stream.
        K15 updates the local shard; the vector tail (biases, LayerNorm) stays replicated like DeepSpeed's
        persistence threshold                                    [TH:231-237, zero3.json:10-27]
"""
from __future__ import annotations

import math
import os
from typing import Dict, Optional

import torch

from . import ops as _ops
from .config import StrategyConfig
from .models.layout import ParamLayout, build_layout
from .optim.adamw import warmup_lr
from .parallel.collectives import DistCtx
from .parallel.torch_engine import SHARDED_OPT, SHARDED_PARAMS, comm_strategy, effective_grad_accum

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
                 seed: int = 42, param_gather: Optional[str] = None, nvtx: bool = False):
        self.ops = _ops.load()
        self.nvtx = bool(nvtx)                # host-side ranges: meaningful for eager steps and graph capture only
        # fsdp / zero3 weight all-gather: "gemm" = K14, every consuming GEMM fetches its B tiles from the owners'
        # shards by TMA (cluster multicast), "stream" = one peer-gather kernel per transformer block on the comm
        # stream, one block ahead of the compute stream
        # precedence: constructor argument, environment, strategy config file, default
        self.param_gather = (param_gather or os.environ.get("B200BENCH_PARAM_GATHER")
                             or getattr(cfg, "param_gather", None) or "stream")
        if self.param_gather not in ("gemm", "stream"):
            raise ValueError(f"param_gather must be 'gemm' or 'stream', got {self.param_gather!r}")
        self.n_slots = 2
        self._regather_tables = True
        self._invalidate_staging()
        self.strategy, self.cfg, self.ctx, self.device = strategy, cfg, ctx, device
        self.comm_strategy = cs = comm_strategy(strategy, cfg)     # fsdp follows the YAML's sharding_strategy
        self.accum = effective_grad_accum(strategy, grad_accum, cfg.grad_accum_mode)
        self.dropout, self.causal = float(dropout), bool(causal)
        self.B, self.T = int(batch), int(seq_len)
        self.V, self.d = model_cfg["vocab_size"], model_cfg["n_embd"]
        self.H, self.L = model_cfg["n_head"], model_cfg["n_layer"]
        self.hd = self.d // self.H
        if self.hd not in self.ops.attention_head_dims():
            raise NotImplementedError(f"native attention kernels are built for head_dim "
                                      f"{self.ops.attention_head_dims()}, got {self.hd}")
        if self.T > model_cfg["block_size"]:
            raise ValueError(f"Sequence {self.T} exceeds block size {model_cfg['block_size']}")
        self.M = self.B * self.T
        ws, rank = ctx.world_size, ctx.rank
        self.layout: ParamLayout = build_layout(world_size=ws, **model_cfg)
        Lt = self.layout
        dev = device

        # ---- communication runtime ------------------------------------------------------------
        self.comm_mode = comm if ws > 1 else "none"
        self.comm = None
        if ws > 1 and comm == "fused":
            from .parallel.peer import PeerComm

            # K13 CTA count: more peers per CTA means more loads in flight per CTA, so fewer CTAs (= fewer SMs
            # taken from the backward GEMMs) reach the same NVLink rate.  Measured: N=8 16 CTAs 9.34 ms/step,
            # 32: 9.41, 64: 9.60 (profiles/n8*); N=2 needs 64 (profiles/comm_n2_v2.log).
            ctas = (int(os.environ.get("B200BENCH_COMM_CTAS", "0")) or int(getattr(cfg, "comm_ctas", 0) or 0)
                    or {2: 64, 4: 32}.get(ws, 16))
            self.comm = PeerComm(ctx, device, comm_ctas=ctas)
        self.sharded_params = self.comm is not None and cs in SHARDED_PARAMS

        # ---- parameters / gradients ---------------------------------------------------------------
        if self.comm is not None:
            self.grads_pb = self.comm.alloc("grads", Lt.total, BF16)      # every peer can read them
            self.grads = self.grads_pb.tensor
        else:
            self.grads = torch.zeros(Lt.total, dtype=BF16, device=dev)
        if self.comm is not None and not self.sharded_params:
            self.params_pb = self.comm.alloc("params", Lt.total, BF16)    # every peer can write them
            full = self.params_pb.tensor
        else:
            full = torch.zeros(Lt.total, dtype=BF16, device=dev)           # temporary when sharded
        self._init_params(full, seed)
        # N2 / N9: rank 0's initial weights to everyone.  fsdp_config.yaml `sync_module_states: false` skips it (every
        # rank initialises from the same seed here, so the replicas agree either way)
        if ws > 1 and not (strategy == "fsdp" and not cfg.fsdp_sync_module_states):
            ctx.broadcast(full, 0)
        self.hp = torch.zeros(16, dtype=torch.float32, device=dev)
        self.hp_small = torch.zeros(16, dtype=torch.float32, device=dev)
        self.gnorm_sq = torch.zeros(1, dtype=torch.float32, device=dev)
        self.gnorm2 = torch.zeros(2, dtype=torch.float32, device=dev)      # K15: norm, norm^2
        self.opt_step = 0
        self.micro = 0
        self.G: Dict[str, torch.Tensor] = {s.name: self.grads[s.offset:s.offset + s.numel].view(s.shape)
                                           for s in Lt}
        if self.sharded_params:
            self._setup_sharded(full)
            del full
        else:
            self.params = full
            self.lo, self.hi = Lt.shard_range(rank) if cs in SHARDED_OPT else (0, Lt.total)
            self.master = self.params[self.lo:self.hi].float()
            self.P = {s.name: self.params[s.offset:s.offset + s.numel].view(s.shape) for s in Lt}
        n_opt = self.hi - self.lo
        self.exp_avg = torch.zeros(n_opt, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(n_opt, dtype=torch.float32, device=dev)
        self.grad_shard = torch.zeros(n_opt, dtype=BF16, device=dev) if cs in SHARDED_OPT else None

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
        self.dln2 = e(M, d)          # LN2's input gradient gets its own buffer (the side stream reads it late)
        self.dy_drop = e(M, d)
        self.dbig = e(M, 4 * d)
        self.dqkv = e(M, 3 * d)
        self.datt = e(M, d)
        self.dq_acc = torch.zeros(M, d, dtype=torch.float32, device=dev)   # attention bwd leaves it zeroed again
        self.attD = e(self.B * self.H * self.T, dt=torch.float32)
        # fused LayerNorm backward: per-CTA fp32 partial column sums of the two LayerNorms of a block; the folds
        # into dgamma / dbeta run on the side stream
        self.ln_part = [torch.empty(self.ops.layernorm_bwd_fused_scratch_bytes(M, d) // 4, dtype=torch.float32,
                                    device=dev) for _ in range(2)]
        self._ab = set(filter(None, os.environ.get("B200BENCH_AB", "").split(",")))   # A/B switches (diagnostics)
        sbytes = max(self.ops.colreduce_scratch_bytes(M, 4 * d, 1), self.ops.colreduce_scratch_bytes(M, d, 2))
        self.scratch_side = torch.zeros(sbytes // 4 + 64, dtype=torch.float32, device=dev)   # counters start at 0
        # weight-gradient GEMMs and bias / LayerNorm-parameter reductions do not feed the backward chain:
        # they run on a side stream and fill the tail bubbles of the dgrad -> attention -> LN kernels
        self.side = torch.cuda.Stream(device=dev)
        self.overlap_wgrad = True
        # ddp without clipping: AdamW of a block runs as soon as its gradients are final (memory-bound, it
        # hides behind the compute-bound backward of the earlier blocks); the boundary has nothing left to do
        self.opt_in_backward = (cs == "ddp" and not (cfg.optim.grad_clip > 0)
                                and os.environ.get("B200BENCH_OPT_IN_BACKWARD", "1") != "0")
        self.seed_base = seed
        self.step_seed = torch.full((1,), seed, dtype=torch.int32, device=dev)
        self.ops.step_seed_register(self.step_seed)

        self.use_graph = use_graph
        self._graphs: Dict[tuple, torch.cuda.CUDAGraph] = {}
        self._warm = 0
        self._graph_launches = 0
        self._launches_per_step: Optional[int] = None
        # DDP gradient buckets = the matrices of one transformer block each + (embeddings) + (vector tail)
        self._bucket_head = (0, Lt.layer_range(0)[0])
        self._bucket_tail = (Lt.layer_range(self.L - 1)[1], Lt.total)
        # K13 bucket = this many consecutive transformer blocks per all-reduce launch (torch DDP's bucket_cap_mb,
        # TH:217-222; 25 MiB = one tier-A block of bf16 gradients).  overlap_comm: false keeps K13 on the compute
        # stream (DeepSpeed's knob, zero2.json:19).
        a0_, a1_ = Lt.layer_range(0)
        if cfg.ddp_bucket_cap_mb > 0:
            self.blocks_per_bucket = max(1, int(cfg.ddp_bucket_cap_mb * (1 << 20) // max(1, (a1_ - a0_) * 2)))
        else:
            # measured (profiles/round2): the in-switch all-reduce (>= 4 ranks) is best launched per block (8 GPUs:
            # 8.89 vs 9.03 ms per step); the peer load/store kernel at 2 ranks per two blocks (9.00 vs 9.19 ms)
            self.blocks_per_bucket = 1 if (self.comm is not None and self.comm.use_nvls) else 2
        self.blocks_per_bucket = int(os.environ.get("B200BENCH_BLOCKS_PER_BUCKET", "0")) or self.blocks_per_bucket
        self.overlap_comm = bool(cfg.overlap_comm)
        if ws > 1:
            torch.cuda.synchronize(dev)
            ctx.barrier()

    # ---------------------------------------------------------------------------------------
    def _init_params(self, flat: torch.Tensor, seed: int) -> None:
        """Reference initialisation (TH:69-78): N(0, 0.02) for Linear / Embedding weights, zero
        biases, LayerNorm (1, 0); MultiheadAttention.in_proj keeps xavier_uniform (not an nn.Linear)."""
        g = torch.Generator(device=self.device)
        g.manual_seed(seed)
        for n in self.layout.state_dict_order():
            s = self.layout[n]
            view = flat[s.offset:s.offset + s.numel].view(s.shape)
            if n.endswith("ln_1.weight") or n.endswith("ln_2.weight") or n.endswith("ln_f.weight"):
                view.fill_(1.0)
            elif n.endswith("bias"):
                view.zero_()
            elif n.endswith("in_proj_weight"):
                bound = math.sqrt(6.0 / (s.shape[0] + s.shape[1]))
                view.copy_((torch.rand(s.shape, generator=g, device=self.device) * 2 - 1) * bound)
            else:
                view.copy_(torch.randn(s.shape, generator=g, device=self.device) * 0.02)

    def _setup_sharded(self, full: torch.Tensor) -> None:
        """FSDP / ZeRO-3 storage: this rank's partition of the matrix region (peer-visible), a
        replicated copy of the vector tail, and staging buffers the gathers write into."""
        Lt, dev, rank = self.layout, self.device, self.ctx.rank
        S = Lt.big_shard_size
        self.S = S
        self.lo, self.hi = Lt.big_shard_range(rank)
        self.shard_pb = self.comm.alloc("pshard", S, BF16)
        self.shard_pb.tensor.copy_(full[self.lo:self.hi])
        self.adj_table = self.comm.adjusted_table(self.shard_pb, S)
        self.master = self.shard_pb.tensor.float()
        b0, b1 = Lt.small_range()
        self.small = full[b0:b1].clone()
        self.master_small = self.small.float()
        self.m_small = torch.zeros_like(self.master_small)
        self.v_small = torch.zeros_like(self.master_small)
        wte, wpe = Lt["transformer.wte.weight"], Lt["transformer.wpe.weight"]
        # "stream" mode: whole-block staging slots, block l lives in slot l % n_slots.  Two slots = gather again
        # in backward (full re-sharding after forward).  L slots = every block stays resident until the end of the
        # micro-step: ZeRO-3 when the model fits stage3_max_live_parameters (zero3.json:11, DeepSpeed keeps such
        # parameters live), FSDP per reshard_after_forward (the reference's single flat unit never re-shards
        # between forward and backward either, TH:229-237).
        a0, a1 = Lt.layer_range(0)
        if self.strategy == "zero3":
            keep = Lt.big_end <= self.cfg.stage3_max_live_parameters
        else:
            r = self.cfg.fsdp_reshard_after_forward
            keep = (not r) if isinstance(r, bool) else \
                (self.L * (a1 - a0) * 2 <= 0.05 * torch.cuda.get_device_properties(dev).total_memory)
        self.keep_gathered = bool(keep) and self.param_gather == "stream"
        self.n_slots = self.L if self.keep_gathered else 2
        # gather-ahead depth in blocks, per direction.  zero3: stage3_prefetch_bucket_size (elements, zero3.json:21)
        # in units of one block; fsdp: forward_prefetch / backward_prefetch of the YAML (:8-10).  Bounded by the
        # staging slots that exist (2 slots = at most one block ahead) and by 4.
        blk = max(1, a1 - a0)
        if self.strategy == "zero3":
            fwd = bwd = int(self.cfg.stage3_prefetch_bucket_size // blk)
        else:
            fwd = 1 if self.cfg.fsdp_forward_prefetch else 0
            bwd = 1 if self.cfg.fsdp_backward_prefetch == "backward_pre" else 0
        cap = min(4, self.n_slots - 1)
        self.prefetch_fwd, self.prefetch_bwd = min(fwd, cap), min(bwd, cap)
        # CTAs of one block gather: 32 are enough for tier A's 25 MB blocks (measured, profiles/README.md); big
        # blocks (403 MB at the 7B shape) need the full ~650 GB/s of a 128-CTA copy to stay ahead of the compute
        self.gather_ctas = (int(os.environ.get("B200BENCH_GATHER_CTAS", "0"))
                            or int(getattr(self.cfg, "gather_ctas", 0) or 0)
                            or min(128, max(32, (a1 - a0) * 2 // (2 << 20))))
        # Gather kernel.  From 4 ranks on: K16, ONE collective per range - every rank fetches 1/N of it from the
        # owner and writes it into every rank's staging copy through the switch (a block lies inside one or two
        # partitions: with every rank pulling the whole block, N-1 readers share one owner's links).  At 2 ranks
        # there is one reader per owner either way: plain peer pull, no barriers.  B200BENCH_GATHER_COLL=1/0 forces.
        gc = os.environ.get("B200BENCH_GATHER_COLL", "")
        self.gather_coll = (self.ctx.world_size >= 4) if gc == "" else (gc != "0")
        n_tab = wte.numel + wpe.numel
        n_stage = n_tab + (self.n_slots * (a1 - a0) if self.param_gather == "stream" else 0)
        if self.gather_coll:                  # K16 writes into the peers' staging buffers: symmetric allocation
            self.stage_pb = self.comm.alloc("stage", n_stage, BF16)
            self._stage = self.stage_pb.tensor
        else:
            self.stage_pb, self._stage = None, torch.empty(n_stage, dtype=BF16, device=dev)
        self.stage_wte, self.stage_wpe = self._stage[:wte.numel], self._stage[wte.numel:n_tab]
        self._stage_off = {"transformer.wte.weight": 0, "transformer.wpe.weight": wte.numel, "slots": n_tab}
        self.stage_layer = (self._stage[n_tab:].view(self.n_slots, a1 - a0)
                            if self.param_gather == "stream" else None)
        self._invalidate_staging()
        # _slot_layer: block resident (or in flight) in each slot during this micro-step; _slot_done: comm-stream
        # event of the gather filling it; _slot_pending: the compute stream has not waited for that event yet
        self.params = None
        self.P = {}
        for s in Lt:
            if s.offset >= b0:
                self.P[s.name] = self.small[s.offset - b0:s.offset - b0 + s.numel].view(s.shape)
            elif s.name == "transformer.wte.weight":
                self.P[s.name] = self.stage_wte.view(s.shape)
            elif s.name == "transformer.wpe.weight":
                self.P[s.name] = self.stage_wpe.view(s.shape)
            # block matrices have no resident view: _w() hands out the staging slot per GEMM

    def _w(self, name: str):
        """Weight matrix for the next GEMM: (tensor or shape, K14 peer descriptor).  Resident weights:
        (P[name], None).  Sharded weights, "stream": the block's staging slot.  Sharded weights, "gemm" (K14):
        only the shape - the GEMM's TMA producer fetches every tile from the owning rank's shard."""
        if not self.sharded_params:
            return self.P[name], None
        s = self.layout[name]
        if self.param_gather == "stream":
            a, _ = self.layout.layer_range(s.layer)
            slot = s.layer % self.n_slots
            assert self._slot_layer[slot] == s.layer and not self._slot_pending[slot], (name, self._slot_layer)
            return self.stage_layer[slot][s.offset - a:s.offset - a + s.numel].view(s.shape), None
        return tuple(s.shape), dict(table=self.shard_pb.table, ws=self.ctx.world_size, flat_off=s.offset, shard=self.S)

    def _gather_block(self, l: int) -> None:
        """Queue the peer gather of block l's four matrices into its slot on the comm stream.  The comm
        stream first waits for the compute stream's current position: everything that read the slot's previous
        occupant (block l +- 2, or l itself before an optimizer step) has been queued before this call."""
        slot = l % self.n_slots
        a, b = self.layout.layer_range(l)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self.comm.stream.wait_event(ev)
        with torch.cuda.stream(self.comm.stream):
            self._gather(self._stage_off["slots"] + slot * (b - a), a, b - a, self.gather_ctas)
        done = torch.cuda.Event()
        done.record(self.comm.stream)
        self._slot_layer[slot], self._slot_done[slot], self._slot_pending[slot] = l, done, True

    def _gather(self, stage_off: int, flat_off: int, n: int, ctas: int = 0) -> None:
        """Flat range [flat_off, flat_off + n) of the partitioned weights -> staging[stage_off:], current stream."""
        if self.gather_coll:
            self.comm.gather_bcast(self.adj_table, self.stage_pb, stage_off, flat_off, n, self.S)
        else:
            self.comm.gather_range(self.adj_table, self._stage[stage_off:stage_off + n], flat_off, n, self.S,
                                   ctas=ctas or 128)

    def _layer_begin(self, l: int, step: int) -> None:
        """Stream-mode prefetch schedule: make block l resident for the compute stream, then start the gather of
        the block that comes next in this direction (step = +1 forward, -1 backward)."""
        if not (self.sharded_params and self.param_gather == "stream"):
            return
        slot = l % self.n_slots
        if self._slot_layer[slot] != l:
            self._gather_block(l)
        if self._slot_pending[slot]:
            torch.cuda.current_stream().wait_event(self._slot_done[slot])
            self._slot_pending[slot] = False
        for k in range(1, (self.prefetch_fwd if step > 0 else self.prefetch_bwd) + 1):
            nxt = l + step * k
            if 0 <= nxt < self.L and self._slot_layer[nxt % self.n_slots] != nxt:
                self._gather_block(nxt)

    def _invalidate_staging(self) -> None:
        n = getattr(self, "n_slots", 2)
        self._slot_layer, self._slot_done, self._slot_pending = [-1] * n, [None] * n, [False] * n

    # ---------------------------------------------------------------------------------------
    def _seed(self, layer: int, site: int) -> int:
        # The per-step component lives in device memory (graph-replay safe): see ops.step_seed_advance.
        return _mix(self.seed_base, layer, site)

    def _forward(self) -> None:
        o, P, p = self.ops, self.P, self.dropout
        B, T, d, H, hd = self.B, self.T, self.d, self.H, self.hd
        if self.sharded_params and self._regather_tables:   # embedding tables: plain peer gather up front
            for name in ("transformer.wte.weight", "transformer.wpe.weight"):
                s = self.layout[name]
                self._gather(self._stage_off[name], s.offset, s.numel)
        o.embedding_fwd(self.tokens, P["transformer.wte.weight"], P["transformer.wpe.weight"], self.x[0],
                        B, T, d, p, self._seed(-1, 0))
        for l in range(self.L):
            pre = f"transformer.h.{l}."
            st = self.stats[l]
            self._layer_begin(l, +1)
            o.layernorm_fwd(self.x[l], P[pre + "ln_1.weight"], P[pre + "ln_1.bias"], self.ln1[l], st[0], st[1])
            w, g = self._w(pre + "attn.in_proj_weight")
            o.linear_fwd(self.ln1[l], w, self.qkv[l], bias=P[pre + "attn.in_proj_bias"], peer=g)
            o.attention_fwd(self.qkv[l], self.att[l], self.lse[l], B, T, H, hd, self.causal, p, self._seed(l, 1))
            w, g = self._w(pre + "attn.out_proj.weight")
            o.linear_fwd(self.att[l], w, self.xmid[l], bias=P[pre + "attn.out_proj.bias"], epi=_ops.EPI_BIAS_RESID,
                         aux=self.x[l], peer=g)
            o.layernorm_fwd(self.xmid[l], P[pre + "ln_2.weight"], P[pre + "ln_2.bias"], self.ln2[l], st[2], st[3])
            w, g = self._w(pre + "mlp.0.weight")
            # the forward keeps gelu'(pre) instead of pre: the backward epilogue is then a multiply
            o.linear_fwd(self.ln2[l], w, self.pre[l], bias=P[pre + "mlp.0.bias"], epi=_ops.EPI_BIAS_GELU_GRAD,
                         out2=self.act[l], peer=g)
            w, g = self._w(pre + "mlp.2.weight")
            o.linear_fwd(self.act[l], w, self.x[l + 1], bias=P[pre + "mlp.2.bias"], epi=_ops.EPI_BIAS_RESID,
                         aux=self.xmid[l], drop_p=p, seed=self._seed(l, 2), peer=g)
        o.layernorm_fwd(self.x[self.L], P["transformer.ln_f.weight"], P["transformer.ln_f.bias"], self.lnf,
                        self.statf[0], self.statf[1])
        o.linear_fwd(self.lnf, P["transformer.wte.weight"], self.logits)

    def _ln_bwd(self, dy, x, gamma, mean, rstd, dx, dres, drop_out, seed, part, dg, db, acc):
        """LayerNorm backward + parameter gradients (+ dropout-masked copy for the block below)."""
        o, p = self.ops, self.dropout
        M, d = x.shape
        if "oldln" in self._ab:         # diagnostics: round-1 schedule (three kernels)
            o.layernorm_bwd(dy, x, gamma, mean, rstd, dx, dres=dres)
            self._on_side(lambda: o.layernorm_bwd_params(dy, x, mean, rstd, dg, db, self.scratch_side, acc))
            if drop_out is not None:
                o.dropout_apply(dx, drop_out, M, d, p, seed)
            return
        if "lnpart" in self._ab:        # column sums inside the main-stream kernel + fold on the side stream
            o.layernorm_bwd_fused(dy, x, gamma, mean, rstd, dx, dres=dres, dx_drop=drop_out, drop_p=p, seed=seed,
                                  partials=part)
            self._on_side(lambda: self._fold(part, M, d, dg, db, acc))
            return
        # Default: the main stream gets the lightest possible kernel (dx + the masked copy: no shared memory, no
        # block barriers, < 128 registers, so it slots in next to side-stream CTAs); dgamma / dbeta only feed the
        # optimizer and are reduced on the side stream.  Measured 0.15 ms per step faster than the in-kernel sums.
        o.layernorm_bwd_fused(dy, x, gamma, mean, rstd, dx, dres=dres, dx_drop=drop_out, drop_p=p, seed=seed)
        self._on_side(lambda: o.layernorm_bwd_params(dy, x, mean, rstd, dg, db, self.scratch_side, acc))

    def _fold(self, part, M, d, dg, db, acc):
        if "nofold" not in self._ab:
            self.ops.layernorm_param_fold(part, M, d, dg, db, acc)

    def _on_side(self, fn):
        """Run ``fn`` (kernels that only produce parameter gradients) on the side stream, ordered after
        everything issued so far on the current stream."""
        if not self.overlap_wgrad:
            fn()
            return None
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self.side.wait_event(ev)
        with torch.cuda.stream(self.side):
            fn()
        done = torch.cuda.Event()
        done.record(self.side)
        return done

    def _adamw_range(self, a: int, b: int) -> None:
        self.ops.adamw_flat(self.master[a:b], self.exp_avg[a:b], self.exp_avg_sq[a:b], self.grads[a:b],
                            self.params[a:b], self.hp)

    def _join_side(self) -> None:
        if self.overlap_wgrad:
            ev = torch.cuda.Event()
            ev.record(self.side)
            torch.cuda.current_stream().wait_event(ev)

    def _loss_and_backward(self, first_micro: bool, last_micro: bool) -> None:
        o, P, G, p = self.ops, self.P, self.G, self.dropout
        reduce_now = self.comm is not None and self.comm_strategy == "ddp" and last_micro
        opt_now = self.opt_in_backward and last_micro and (self.comm is not None or self.ctx.world_size == 1)
        B, T, d, H, hd, M, V = self.B, self.T, self.d, self.H, self.hd, self.M, self.V
        acc = not first_micro
        o.cross_entropy(self.logits, self.targets, self.loss_rows, self.n_valid, self.loss, M, V,
                        grad_scale=1.0 / self.accum, write_grad=True)
        wte = P["transformer.wte.weight"]
        drop = p > 0
        if "colsum" in self._ab:        # diagnostics: separate column-sum kernels instead of the row-sum MMA
            def wgrad(dy_, x_, dw_, accumulate, dbias):
                o.colsum(dy_, dbias, self.scratch_side, accumulate)
                o.linear_wgrad(dy_, x_, dw_, accumulate=accumulate)
        else:
            def wgrad(dy_, x_, dw_, accumulate, dbias):
                o.linear_wgrad(dy_, x_, dw_, accumulate=accumulate, dbias=dbias)
        # lm_head: dlnf = dlogits @ wte ; dwte (+)= dlogits^T @ lnf      (tied weight)
        self._on_side(lambda: o.linear_wgrad(self.logits, self.lnf, G["transformer.wte.weight"], accumulate=acc))
        o.linear_dgrad(self.logits, wte, self.dln)
        # every LayerNorm backward is ONE launch: input gradient (+ residual gradient), gamma / beta gradients, and
        # - when the block below ends in a dropout - the masked copy of the gradient that its fc2 GEMMs consume
        self._ln_bwd(self.dln, self.x[self.L], P["transformer.ln_f.weight"], self.statf[0], self.statf[1], self.dres,
                     None, self.dy_drop if drop else None, self._seed(self.L - 1, 2), self.ln_part[0],
                     G["transformer.ln_f.weight"], G["transformer.ln_f.bias"], acc)
        for l in reversed(range(self.L)):
            pre = f"transformer.h.{l}."
            st = self.stats[l]
            # every buffer the side stream read for block l+1 is about to be overwritten
            self._join_side()
            self._layer_begin(l, -1)
            # ---- MLP ----  (weight-gradient GEMMs also emit the bias gradient: row sums of dY^T in the same loop)
            dy = self.dy_drop if drop else self.dres
            dy_read = self._on_side(lambda dy=dy: wgrad(dy, self.act[l], G[pre + "mlp.2.weight"],
                                                        accumulate=acc, dbias=G[pre + "mlp.2.bias"]))
            w, g = self._w(pre + "mlp.2.weight")
            o.linear_dgrad(dy, w, self.dbig, epi=_ops.EPI_MUL, aux=self.pre[l], peer=g)
            self._on_side(lambda: wgrad(self.dbig, self.ln2[l], G[pre + "mlp.0.weight"], accumulate=acc,
                                        dbias=G[pre + "mlp.0.bias"]))
            w, g = self._w(pre + "mlp.0.weight")
            o.linear_dgrad(self.dbig, w, self.dln2, peer=g)
            self._ln_bwd(self.dln2, self.xmid[l], P[pre + "ln_2.weight"], st[2], st[3], self.dmid, self.dres, None, 0,
                         self.ln_part[0], G[pre + "ln_2.weight"], G[pre + "ln_2.bias"], acc)
            # ---- attention ----
            self._on_side(lambda: wgrad(self.dmid, self.att[l], G[pre + "attn.out_proj.weight"],
                                        accumulate=acc, dbias=G[pre + "attn.out_proj.bias"]))
            w, g = self._w(pre + "attn.out_proj.weight")
            # the out-proj dgrad writes dO and, from the same accumulator tile, D = rowsum(dO o O) per head
            if "attn4" in self._ab:
                o.linear_dgrad(self.dmid, w, self.datt, peer=g)
                o.attention_bwd(self.qkv[l], self.att[l], self.datt, self.lse[l], self.dqkv, self.attD, self.dq_acc,
                                B, T, H, hd, self.causal, p, self._seed(l, 1))
            else:
                o.linear_dgrad(self.dmid, w, self.datt, epi=_ops.EPI_ROWDOT, aux=self.att[l], peer=g,
                               rowdot=self.attD, rd_T=T, rd_hd=hd)
                o.attention_bwd(self.qkv[l], self.att[l], self.datt, self.lse[l], self.dqkv, self.attD, self.dq_acc,
                                B, T, H, hd, self.causal, p, self._seed(l, 1), have_d=True, acc_is_zero=True)
            self._on_side(lambda: wgrad(self.dqkv, self.ln1[l], G[pre + "attn.in_proj_weight"],
                                        accumulate=acc, dbias=G[pre + "attn.in_proj_bias"]))
            w, g = self._w(pre + "attn.in_proj_weight")
            o.linear_dgrad(self.dqkv, w, self.dln, peer=g)
            if dy_read is not None:      # the fc2 weight gradient on the side stream is still entitled to read dy
                torch.cuda.current_stream().wait_event(dy_read)
            self._ln_bwd(self.dln, self.x[l], P[pre + "ln_1.weight"], st[0], st[1], self.dres, self.dmid,
                         self.dy_drop if (drop and l > 0) else None, self._seed(l - 1, 2), self.ln_part[1],
                         G[pre + "ln_1.weight"], G[pre + "ln_1.bias"], acc)
            a, b = self.layout.layer_range(l)
            if reduce_now:
                # K13: a bucket (blocks_per_bucket consecutive blocks, contiguous in the flat buffer) is final ->
                # all-reduce it on the comm stream while the compute stream continues with the blocks below
                if (self.L - 1 - l) % self.blocks_per_bucket == 0:
                    bucket_end = b
                if (self.L - l) % self.blocks_per_bucket == 0 or l == 0:
                    self._join_side()
                    if self.overlap_comm:
                        self.comm.fork()
                    self.comm.allreduce(self.grads_pb, a, bucket_end - a, 1.0 / self.ctx.world_size,
                                        side_stream=self.overlap_comm)
                    if opt_now:
                        with torch.cuda.stream(self.comm.stream if self.overlap_comm else torch.cuda.current_stream()):
                            self._adamw_range(a, bucket_end)
            elif opt_now:
                # single GPU: the block's wgrads were just queued on the side stream; its AdamW follows them
                self._on_side(lambda: self._adamw_range(a, b))
        self._join_side()       # lm_head wgrad must land before the embedding scatter-add onto the tied weight
        o.embedding_bwd(self.tokens, self.dres, G["transformer.wte.weight"], G["transformer.wpe.weight"],
                        B, T, d, p, self._seed(-1, 0), accumulate_wpe=acc)
        if reduce_now:
            if self.overlap_comm:
                self.comm.fork()
            for a, b in (self._bucket_tail, self._bucket_head):
                self.comm.allreduce(self.grads_pb, a, b - a, 1.0 / self.ctx.world_size, side_stream=self.overlap_comm)
                if opt_now:
                    with torch.cuda.stream(self.comm.stream if self.overlap_comm else torch.cuda.current_stream()):
                        self._adamw_range(a, b)
            self.comm.join()
        elif opt_now:
            for a, b in (self._bucket_tail, self._bucket_head):
                self._adamw_range(a, b)
            self._join_side()

    # ---------------------------------------------------------------------------------------
    def _micro_step(self, first_micro: bool, last_micro: bool) -> None:
        if self.dropout > 0:
            self.ops.step_seed_advance()      # new dropout masks every micro-step, also under graph replay
        self._regather_tables = first_micro or not getattr(self, "keep_gathered", False)
        if self._regather_tables:
            self._invalidate_staging()      # weights moved at the optimizer boundary / slots were recycled
        else:
            # every block was gathered earlier in this accumulation cycle and nothing has updated the weights
            # since.  Set explicitly (not inherited) so that a captured graph never depends on host-side history.
            self._slot_layer, self._slot_pending = list(range(self.L)), [False] * self.L
        if self.nvtx:
            torch.cuda.nvtx.range_push("forward")
        self._forward()
        if self.nvtx:
            torch.cuda.nvtx.range_pop()
            torch.cuda.nvtx.range_push("loss+backward" + ("+reduce" if last_micro else ""))
        self._loss_and_backward(first_micro, last_micro)
        if self.nvtx:
            torch.cuda.nvtx.range_pop()

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

    def _set_hp(self, grad_scale: float, hp: Optional[torch.Tensor] = None, advance: bool = True) -> None:
        c = self.cfg.optim
        if advance:
            self._lr = warmup_lr(self.opt_step, c)
            self.opt_step += 1
        t = self.opt_step
        bc1 = 1.0 - c.beta1 ** t
        bc2 = 1.0 - c.beta2 ** t
        self.ops.adamw_set_hp(self.hp if hp is None else hp, self._lr, c.beta1, c.beta2, c.eps, c.weight_decay,
                              1.0 / bc1, 1.0 / math.sqrt(bc2), grad_scale, c.grad_clip)

    def _boundary_fused(self, clip: bool) -> None:
        ctx, ws, Lt = self.ctx, self.ctx.world_size, self.layout
        if self.comm_strategy == "ddp":          # gradients were all-reduced and scaled during backward (K13)
            self._set_hp(1.0)
            gn = None
            if clip:
                self.gnorm_sq.zero_()
                self.ops.sumsq(self.grads, self.gnorm_sq, 1.0)
                gn = self.gnorm_sq
            self.ops.adamw_flat(self.master, self.exp_avg, self.exp_avg_sq, self.grads, self.params, self.hp, gn)
            return
        if not self.sharded_params:         # zero2: K15 over this rank's slice of the whole flat buffer
            self._set_hp(1.0 / ws)
            self.comm.rs_adamw(self.grads_pb, self.params_pb, self.lo, self.hi - self.lo, self.master,
                               self.exp_avg, self.exp_avg_sq, self.grad_shard, self.hp, push_all=True)
            return
        # fsdp / zero3: vector tail replicated (K13 + local AdamW), matrix region partitioned (K15)
        b0, b1 = Lt.small_range()
        self.comm.allreduce(self.grads_pb, b0, b1 - b0, 1.0 / ws, side_stream=False)
        extra = None
        if clip:
            self.gnorm_sq.zero_()
            self.ops.sumsq(self.grads[b0:b1], self.gnorm_sq, 1.0)
            extra = self.gnorm_sq
        self._set_hp(1.0 / ws)
        self._set_hp(1.0, hp=self.hp_small, advance=False)
        self.comm.rs_adamw(self.grads_pb, self.shard_pb, self.lo, self.S, self.master, self.exp_avg,
                           self.exp_avg_sq, self.grad_shard, self.hp, push_all=False, gnorm_out=self.gnorm2,
                           extra_norm_sq=extra, weights_table=self.adj_table)
        self.ops.adamw_flat(self.master_small, self.m_small, self.v_small, self.grads[b0:b1], self.small,
                            self.hp_small, self.gnorm2[1:] if clip else None)

    def _optimizer_boundary(self) -> None:
        ctx, ws = self.ctx, self.ctx.world_size
        if self.opt_in_backward and (self.comm is not None or ws == 1):
            return                              # already applied block by block during backward
        scale = 1.0 / ws
        clip = self.cfg.optim.grad_clip > 0
        if self.comm is not None:
            self._boundary_fused(clip)
            return
        if self.comm_strategy == "ddp":
            if ws > 1:
                cap = max(1, (self.cfg.ddp_bucket_cap_mb or 25) * 1024 * 1024 // 2)
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
            if self.comm_strategy != "ddp":
                ctx.all_reduce_sum(self.gnorm_sq)
            gn = self.gnorm_sq
        self.ops.adamw_flat(self.master, self.exp_avg, self.exp_avg_sq, g, self.params[self.lo:self.hi], self.hp, gn)
        if self.comm_strategy in SHARDED_OPT and ws > 1:
            # zero2: all-gather updated partitions (N7); fsdp/zero3: gather before the next forward
            # (N4/N8) — with NCCL both are the same collective, issued here.
            ctx.all_gather(self.params[self.lo:self.hi], self.params)

    def train_step(self, tokens: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
        # the library keeps ONE per-step dropout seed word; with several engines alive in a process (tests build many)
        # each step points it at THIS engine's word, so eager launches never advance another engine's masks (a
        # captured graph has the pointer baked in at capture time, which is the same word)
        self.ops.step_seed_register(self.step_seed)
        self.tokens.copy_(tokens, non_blocking=True)
        if targets is tokens:
            self.targets.copy_(self.tokens, non_blocking=True)
        else:
            self.targets.copy_(targets, non_blocking=True)
        first = (self.micro % self.accum) == 0
        last = ((self.micro + 1) % self.accum) == 0
        before = self.ops.launches
        if last and self.opt_in_backward and (self.comm is not None or self.ctx.world_size == 1):
            self._set_hp(1.0)                   # K13 already scaled the gradients by 1/N
        self._run_micro(first, last)
        self.micro += 1
        if last:
            self._optimizer_boundary()
        eager = self.ops.launches - before
        self._launches_per_step = eager + (self._graph_launches if self._graphs else 0)
        return self.loss

    def profile_breakdown(self, reps: int = 5) -> Dict[str, float]:
        """Per-phase device times (ms) of one optimizer step (the reference has no profiling
        hooks).  Forward, loss+backward and the optimizer boundary are each captured into their own CUDA
        graph and replayed ``reps`` times between CUDA events, so the numbers carry no Python launch
        overhead.  Call between training steps on an accumulation boundary; the replays are real work
        (weights move by ``reps`` optimizer steps) and ranks must call it together."""
        self.ops.step_seed_register(self.step_seed)

        def timed(fn) -> float:
            fn()                                      # warm (also keeps lazy init out of the capture)
            torch.cuda.synchronize(self.device)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                fn()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                g.replay()
            e1.record()
            torch.cuda.synchronize(self.device)
            return e0.elapsed_time(e1) / reps

        def fwd():
            self._regather_tables = True
            self._invalidate_staging()
            self._forward()

        def bwd():
            self._invalidate_staging()
            self._loss_and_backward(True, True)

        if self.opt_in_backward:
            self._set_hp(1.0)                     # the block-wise AdamW launches inside backward read it
        out = {"forward_ms": timed(fwd)}
        if self.sharded_params:
            fwd()                                     # leave the prefetch chain where backward expects it
        out["backward_ms"] = timed(bwd)
        t0 = torch.cuda.Event(enable_timing=True)
        t1 = torch.cuda.Event(enable_timing=True)
        t0.record()
        self._optimizer_boundary()                    # eager: a handful of launches, host cost is negligible
        t1.record()
        torch.cuda.synchronize(self.device)
        out["optimizer_and_comm_ms"] = t0.elapsed_time(t1)
        out["total_ms"] = sum(out.values())
        self.micro = 0
        return {k: round(v, 3) for k, v in out.items()}

    # ---------------------------------------------------------------------------------------
    def _full_params(self) -> torch.Tensor:
        if not self.sharded_params:
            return self.params
        Lt = self.layout
        full = torch.empty(Lt.total, dtype=BF16, device=self.device)
        self.ctx.all_gather(self.shard_pb.tensor, full[:Lt.big_end])
        full[Lt.big_end:].copy_(self.small)
        return full

    def state_dict(self) -> Dict[str, torch.Tensor]:
        """Reference-keyed state_dict.  Collective for the parameter-sharded strategies."""
        flat = self._full_params()
        sd = {s.name: flat[s.offset:s.offset + s.numel].view(s.shape).detach().clone().float()
              for s in (self.layout[n] for n in self.layout.state_dict_order())}
        sd["lm_head.weight"] = sd["transformer.wte.weight"]
        return sd

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        Lt = self.layout
        with torch.no_grad():
            full = self.params if not self.sharded_params else torch.zeros(Lt.total, dtype=BF16, device=self.device)
            for s in Lt:
                full[s.offset:s.offset + s.numel].view(s.shape).copy_(sd[s.name])
            if self.sharded_params:
                self.shard_pb.tensor.copy_(full[self.lo:self.hi])
                self.small.copy_(full[Lt.big_end:])
                self.master_small.copy_(self.small)
                self.m_small.zero_()
                self.v_small.zero_()
                self.master.copy_(self.shard_pb.tensor)
            else:
                self.master.copy_(self.params[self.lo:self.hi])
            self.exp_avg.zero_()
            self.exp_avg_sq.zero_()

    def extra_memory_bytes(self) -> int:
        return self.comm.extra_memory_bytes() if self.comm is not None else 0

    def kernel_launches_per_step(self) -> int:
        return int(self._launches_per_step or 0)

    def close(self) -> None:
        self._graphs.clear()
        if getattr(self.ops, "_seed_word", None) is self.step_seed:
            self.ops.step_seed_unregister()       # the library must not keep pointing at a closed engine's word
        if self.comm is not None:
            self.P, self.G = {}, {}
            self.grads = self.params = None
            self._stage = self.stage_wte = self.stage_wpe = self.stage_layer = self.stage_pb = None
            self.comm.close()
            self.comm = None
