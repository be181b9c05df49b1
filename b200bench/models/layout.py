"""Flat parameter layout shared by every strategy and by the native engine.

All parameters of TinyGPT live in ONE flat buffer: matrices first, vectors last (see
``build_layout``); ``state_dict()`` still lists them in the reference's order.  The same offsets index the bf16 working weights, the bf16 gradient buffer and
the fp32 master / Adam-moment buffers, so a rank's optimizer partition (ZeRO-2/3, FSDP) is simply
a contiguous slice ``[rank*shard, (rank+1)*shard)`` and a DDP bucket is a contiguous range.

This replaces three different mechanisms of the reference stack: the c10d Reducer's buckets
(``TH:217-222``), FSDP's FlatParameter (``TH:231-237``) and DeepSpeed's contiguous partitions
(``configs/deepspeed/zero2.json:18-24``).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Iterator, List, Tuple

ALIGN = 128  # elements; every tensor starts 256-byte aligned in bf16 (TMA needs 16 B)
# K14 (parameter all-gather inside the consuming GEMM, csrc/gemm_sm100.cu): a B tile is fetched by TMA from the
# rank that owns it, so no tile may straddle two owners.  Every matrix therefore starts on a multiple of
# TILE_ROWS of its own rows, and the partition size is a multiple of TILE_ROWS rows of the widest matrix (4d).
TILE_ROWS = 64


@dataclass(frozen=True)
class ParamSpec:
    name: str                 # reference state_dict key
    shape: Tuple[int, ...]
    offset: int               # element offset in the flat buffer
    numel: int
    layer: int                # transformer block index, -1 for embeddings / final LN
    decay: bool               # kept for completeness: the reference decays every tensor (TH:329)


class ParamLayout:
    def __init__(self, specs: List[ParamSpec], world_size: int = 1):
        self.specs = specs
        self.by_name: Dict[str, ParamSpec] = {s.name: s for s in specs}
        end = specs[-1].offset + specs[-1].numel
        # pad so that every rank's shard is a whole number of ALIGN-element groups
        quantum = ALIGN * max(1, world_size)
        self.total = (end + quantum - 1) // quantum * quantum
        self.used = end
        self.world_size = max(1, world_size)
        self.big_end = self.total

    # ---- sharding arithmetic -----------------------------------------------------------
    @property
    def shard_size(self) -> int:
        return self.total // self.world_size

    def shard_range(self, rank: int) -> Tuple[int, int]:
        s = self.shard_size
        return rank * s, (rank + 1) * s

    def owner_of(self, offset: int) -> int:
        return offset // self.shard_size

    def num_params(self) -> int:
        return sum(s.numel for s in self.specs)

    def __iter__(self) -> Iterator[ParamSpec]:
        return iter(self.specs)

    def __getitem__(self, name: str) -> ParamSpec:
        return self.by_name[name]

    def layer_range(self, layer: int) -> Tuple[int, int]:
        """Contiguous [start, end) of the four weight matrices of one transformer block."""
        ss = [s for s in self.specs if s.layer == layer]
        return ss[0].offset, ss[-1].offset + ss[-1].numel

    def small_range(self) -> Tuple[int, int]:
        """The replicated tail: every LayerNorm vector and bias."""
        return self.big_end, self.total

    @property
    def big_shard_size(self) -> int:
        """Partition size when only the matrix region is sharded (FSDP / ZeRO-3)."""
        return self.big_end // self.world_size

    def big_shard_range(self, rank: int) -> Tuple[int, int]:
        s = self.big_shard_size
        return rank * s, (rank + 1) * s

    def state_dict_order(self) -> List[str]:
        """Tensor names in the reference's state_dict order, independent of the flat order."""
        n_layer = 1 + max(s.layer for s in self.specs)
        names = ["transformer.wte.weight", "transformer.wpe.weight"]
        for i in range(n_layer):
            p = f"transformer.h.{i}."
            names += [p + k for k in ("ln_1.weight", "ln_1.bias", "attn.in_proj_weight", "attn.in_proj_bias",
                                      "attn.out_proj.weight", "attn.out_proj.bias", "ln_2.weight", "ln_2.bias",
                                      "mlp.0.weight", "mlp.0.bias", "mlp.2.weight", "mlp.2.bias")]
        return names + ["transformer.ln_f.weight", "transformer.ln_f.bias"]

    def buckets(self, cap_elems: int) -> List[Tuple[int, int]]:
        """Contiguous gradient buckets of at most ``cap_elems`` elements, in REVERSE parameter
        order (the order backward produces them), never splitting a tensor unless the tensor
        itself exceeds the cap."""
        out: List[Tuple[int, int]] = []
        cur_end = self.total
        cur_start = self.total
        for s in reversed(self.specs):
            if cur_end - s.offset > cap_elems and cur_start < cur_end:
                out.append((cur_start, cur_end))
                cur_end = cur_start
            cur_start = s.offset
        if cur_start < cur_end:
            out.append((cur_start, cur_end))
        if out:
            out[-1] = (0, out[-1][1])
        return out


def build_layout(vocab_size: int, n_embd: int, n_head: int, n_layer: int, block_size: int,
                 world_size: int = 1, **_unused) -> ParamLayout:
    """Flat order: every matrix first (embeddings, then per block in_proj / out_proj / fc1 / fc2),
    then every vector (LayerNorm weights and biases, linear biases, final LayerNorm).  Keeping the
    small tensors in one contiguous tail lets the parameter-sharding strategies treat them the way
    DeepSpeed's ``stage3_param_persistence_threshold`` does (``configs/deepspeed/zero3.json:22``:
    biases and LN vectors stay resident on every rank) with ONE replicated region instead of 130
    scattered ones; the matrices in front of it are what gets partitioned."""
    d = n_embd
    specs: List[ParamSpec] = []
    off = 0

    def add(name: str, shape: Tuple[int, ...], layer: int) -> None:
        nonlocal off
        n = 1
        for s in shape:
            n *= s
        if len(shape) == 2:                 # matrices: TILE_ROWS-row alignment in units of their own rows
            q = TILE_ROWS * shape[1]
            off = (off + q - 1) // q * q
        specs.append(ParamSpec(name, shape, off, n, layer, True))
        off = (off + n + ALIGN - 1) // ALIGN * ALIGN

    add("transformer.wte.weight", (vocab_size, d), -1)
    add("transformer.wpe.weight", (block_size, d), -1)
    # blocks start on a TILE_ROWS boundary of the widest matrix, so that (for d a multiple of 256) the per-matrix
    # alignment inside add() never pads between the four matrices of a block
    q = TILE_ROWS * 4 * d
    off = (off + q - 1) // q * q
    for i in range(n_layer):
        p = f"transformer.h.{i}."
        add(p + "attn.in_proj_weight", (3 * d, d), i)
        add(p + "attn.out_proj.weight", (d, d), i)
        add(p + "mlp.0.weight", (4 * d, d), i)
        add(p + "mlp.2.weight", (d, 4 * d), i)
    # the sharded ("big") region ends here, padded so that it splits evenly over the ranks into partitions that
    # are whole TILE_ROWS-row groups of every matrix width (d and 4d)
    quantum = TILE_ROWS * 4 * d * max(1, world_size)
    off = (off + quantum - 1) // quantum * quantum
    big_end = off
    for i in range(n_layer):
        p = f"transformer.h.{i}."
        add(p + "ln_1.weight", (d,), -3)
        add(p + "ln_1.bias", (d,), -3)
        add(p + "attn.in_proj_bias", (3 * d,), -3)
        add(p + "attn.out_proj.bias", (d,), -3)
        add(p + "ln_2.weight", (d,), -3)
        add(p + "ln_2.bias", (d,), -3)
        add(p + "mlp.0.bias", (4 * d,), -3)
        add(p + "mlp.2.bias", (d,), -3)
    add("transformer.ln_f.weight", (d,), -3)
    add("transformer.ln_f.bias", (d,), -3)
    layout = ParamLayout(specs, world_size)
    layout.big_end = big_end
    return layout
