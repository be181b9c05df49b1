"""Synthetic token data.

``SyntheticDataset`` reproduces the reference's dataset (``benchmarking/train_harness.py:138-150``):
``torch.manual_seed(seed)`` then ``randint(0, vocab, (size, seq_len))`` int64 on the host.
``PinnedBatchSource`` is what the harness iterates: it keeps the dataset in pinned host memory
and stages each step's batch to the device with an asynchronous copy on a side stream, so the
measured end-to-end number includes a real H2D transfer per step (the reference does the same
through ``DataLoader(pin_memory=True)`` + ``.to(device, non_blocking=True)``, ``TH:318-325,358``).
Targets equal inputs, unshifted, as in ``TH:359`` unless ``shift_targets``.
"""
from __future__ import annotations

import torch


class SyntheticDataset(torch.utils.data.Dataset):
    def __init__(self, vocab_size: int = 32000, seq_len: int = 2048, size: int = 1000, seed: int = 42,
                 verbose: bool = True):
        g = torch.Generator()
        g.manual_seed(seed)
        self.data = torch.randint(0, vocab_size, (size, seq_len), generator=g)
        if verbose:
            print(f"SyntheticDataset: {size} samples, seq_len={seq_len}")

    def __len__(self) -> int:
        return len(self.data)

    def __getitem__(self, idx: int) -> torch.Tensor:
        return self.data[idx]


class PinnedBatchSource:
    """Deterministic per-rank batch iterator over a SyntheticDataset.

    Sample order: a seeded permutation of the dataset split round-robin over ranks (what
    ``DistributedSampler(shuffle=True)`` does in the reference, ``TH:312-316``), re-shuffled with
    ``seed + epoch`` on wrap-around."""

    def __init__(self, dataset: SyntheticDataset, batch: int, rank: int, world_size: int,
                 device: torch.device, seed: int = 42, int32: bool = True):
        self.ds, self.batch, self.rank, self.ws, self.device = dataset, batch, rank, world_size, device
        self.seed, self.epoch, self.pos = seed, 0, 0
        self.dtype = torch.int32 if int32 else torch.int64
        data = dataset.data.to(self.dtype)
        self.host = data.pin_memory() if device.type == "cuda" else data
        self._order = self._make_order()
        self.bytes_per_step = batch * dataset.data.shape[1] * self.host.element_size()

    def _make_order(self) -> torch.Tensor:
        g = torch.Generator()
        g.manual_seed(self.seed + self.epoch)
        perm = torch.randperm(len(self.ds), generator=g)
        return perm[self.rank::self.ws]

    def next_host(self) -> torch.Tensor:
        if self.pos + self.batch > len(self._order):
            self.epoch += 1
            self.pos = 0
            self._order = self._make_order()
        idx = self._order[self.pos:self.pos + self.batch]
        self.pos += self.batch
        if self.batch == 1:
            return self.host[int(idx[0])].unsqueeze(0)   # a view of pinned memory, no host copy
        out = self.host[idx]
        return out.pin_memory() if self.device.type == "cuda" else out

    def next(self) -> torch.Tensor:
        return self.next_host().to(self.device, non_blocking=True)
