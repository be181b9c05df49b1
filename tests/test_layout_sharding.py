"""Flat layout / shard / bucket index arithmetic for ws in {1,2,4,8} incl. padding and the tied weight."""
import pytest

from b200bench.config import get_model_config
from b200bench.models.layout import ALIGN, build_layout


@pytest.mark.parametrize("ws", [1, 2, 4, 8])
@pytest.mark.parametrize("tier", ["T", "A"])
def test_shards_tile_the_buffer(ws, tier):
    L = build_layout(world_size=ws, **get_model_config(tier, 128))
    assert L.total % (ws * ALIGN) == 0
    assert L.total >= L.used
    covered = 0
    for r in range(ws):
        lo, hi = L.shard_range(r)
        assert lo == covered and hi - lo == L.shard_size
        covered = hi
    assert covered == L.total
    prev_end = 0
    for s in L:                                 # tensors do not overlap and are 16-byte aligned
        assert s.offset >= prev_end and s.offset % 8 == 0
        prev_end = s.offset + s.numel
    assert "lm_head.weight" not in L.by_name    # tied: stored once, under wte


@pytest.mark.parametrize("cap", [1 << 12, 1 << 20, 1 << 30])
def test_buckets_cover_everything_in_reverse(cap):
    L = build_layout(world_size=4, **get_model_config("T", 64))
    b = L.buckets(cap)
    assert b[0][1] == L.total and b[-1][0] == 0
    for (a0, a1), (b0, b1) in zip(b, b[1:]):
        assert b1 == a0 and b0 < b1
    if cap >= L.total:
        assert len(b) == 1


def test_layer_ranges_are_contiguous():
    L = build_layout(world_size=2, **get_model_config("T", 64))
    a0, a1 = L.layer_range(0)
    b0, b1 = L.layer_range(1)
    assert a1 <= b0 and a0 < a1 and b0 < b1
