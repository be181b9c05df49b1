"""TinyGPT in plain PyTorch: the in-repo oracle and the ``--impl torch`` path.

Same architecture, initialisation and ``state_dict`` as the reference model
(``benchmarking/train_harness.py:36-131``): GPT-2-style pre-LN decoder,
``transformer.{wte,wpe,drop,h,ln_f}`` + bias-free ``lm_head`` tied to ``wte`` (``TH:51-61``),
N(0, 0.02) init (``TH:69-78``), ``cross_entropy(ignore_index=-1)`` (``TH:99-103``).

Differences that keep the maths but drop the waste: ``ln_1`` is evaluated
once instead of three times, Q/K/V come from one ``[3d, d]`` projection (the parameter is
still called ``in_proj_weight`` with Q|K|V row order), and attention never materialises the
``T x T`` matrix.  Attention is bidirectional by default because the reference passes no mask
(``TH:127``); ``causal=True`` switches the mask on.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F


class SelfAttention(nn.Module):
    """Parameter-compatible with ``nn.MultiheadAttention(batch_first=True)`` (``TH:114-116``)."""

    def __init__(self, n_embd: int, n_head: int, dropout: float = 0.1, causal: bool = False):
        super().__init__()
        assert n_embd % n_head == 0
        self.n_embd, self.n_head, self.dropout, self.causal = n_embd, n_head, dropout, causal
        self.in_proj_weight = nn.Parameter(torch.empty(3 * n_embd, n_embd))
        self.in_proj_bias = nn.Parameter(torch.zeros(3 * n_embd))
        self.out_proj = nn.Linear(n_embd, n_embd)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        b, t, d = x.shape
        h = self.n_head
        qkv = F.linear(x, self.in_proj_weight, self.in_proj_bias)
        q, k, v = qkv.view(b, t, 3, h, d // h).permute(2, 0, 3, 1, 4)
        o = F.scaled_dot_product_attention(
            q, k, v, dropout_p=self.dropout if self.training else 0.0, is_causal=self.causal)
        return self.out_proj(o.transpose(1, 2).reshape(b, t, d))


class TransformerBlock(nn.Module):
    """``x + attn(ln_1 x)``; ``x + mlp(ln_2 x)`` (``TH:108-131``)."""

    def __init__(self, n_embd: int, n_head: int, dropout: float = 0.1, causal: bool = False):
        super().__init__()
        self.ln_1 = nn.LayerNorm(n_embd)
        self.attn = SelfAttention(n_embd, n_head, dropout, causal)
        self.ln_2 = nn.LayerNorm(n_embd)
        self.mlp = nn.Sequential(
            nn.Linear(n_embd, 4 * n_embd), nn.GELU(), nn.Linear(4 * n_embd, n_embd), nn.Dropout(dropout))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = x + self.attn(self.ln_1(x))
        return x + self.mlp(self.ln_2(x))


class TinyGPT(nn.Module):
    def __init__(self, vocab_size: int = 32000, n_embd: int = 768, n_head: int = 12, n_layer: int = 12,
                 block_size: int = 4096, dropout: float = 0.1, causal: bool = False, verbose: bool = True):
        super().__init__()
        self.block_size = block_size
        self.config = dict(vocab_size=vocab_size, n_embd=n_embd, n_head=n_head, n_layer=n_layer,
                           block_size=block_size, dropout=dropout, causal=causal)
        self.transformer = nn.ModuleDict({
            "wte": nn.Embedding(vocab_size, n_embd),
            "wpe": nn.Embedding(block_size, n_embd),
            "drop": nn.Dropout(dropout),
            "h": nn.ModuleList([TransformerBlock(n_embd, n_head, dropout, causal) for _ in range(n_layer)]),
            "ln_f": nn.LayerNorm(n_embd),
        })
        self.lm_head = nn.Linear(n_embd, vocab_size, bias=False)
        self.transformer["wte"].weight = self.lm_head.weight  # weight tying (TH:61)
        self.apply(self._init_weights)
        if verbose:
            n_params = sum(p.numel() for p in self.parameters())
            print(f"Model initialized: {n_params/1e6:.2f}M parameters")

    @staticmethod
    def _init_weights(module: nn.Module) -> None:
        if isinstance(module, nn.Linear):
            nn.init.normal_(module.weight, mean=0.0, std=0.02)
            if module.bias is not None:
                nn.init.zeros_(module.bias)
        elif isinstance(module, nn.Embedding):
            nn.init.normal_(module.weight, mean=0.0, std=0.02)
        elif isinstance(module, nn.LayerNorm):
            nn.init.zeros_(module.bias)
            nn.init.ones_(module.weight)
        elif isinstance(module, SelfAttention):
            # nn.MultiheadAttention is not an nn.Linear, so the reference leaves its in_proj at
            # the module default (xavier_uniform weight, zero bias); out_proj IS visited (TH:69-73).
            nn.init.xavier_uniform_(module.in_proj_weight)
            nn.init.zeros_(module.in_proj_bias)

    def num_params(self) -> int:
        return sum(p.numel() for p in self.parameters())

    def forward(self, idx: torch.Tensor, targets: Optional[torch.Tensor] = None
                ) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        b, t = idx.size()
        assert t <= self.block_size, f"Sequence {t} exceeds block size {self.block_size}"
        pos = torch.arange(0, t, dtype=torch.long, device=idx.device).unsqueeze(0)
        x = self.transformer["drop"](self.transformer["wte"](idx) + self.transformer["wpe"](pos))
        for block in self.transformer["h"]:
            x = block(x)
        x = self.transformer["ln_f"](x)
        logits = self.lm_head(x)
        loss = None
        if targets is not None:
            loss = F.cross_entropy(logits.view(-1, logits.size(-1)).float(), targets.view(-1), ignore_index=-1)
        return logits, loss
