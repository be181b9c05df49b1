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


# ---- property tests over random model shapes and world sizes -------------------------------------
from hypothesis import given, settings, strategies as st  # noqa: E402


@settings(max_examples=60, deadline=None)
@given(vocab=st.integers(8, 4000), d8=st.integers(1, 48), n_layer=st.integers(1, 6), block=st.integers(4, 300),
       ws=st.sampled_from([1, 2, 3, 4, 8]))
def test_layout_invariants_hold_for_any_shape(vocab, d8, n_layer, block, ws):
    d = 8 * d8
    L = build_layout(vocab_size=vocab, n_embd=d, n_head=1, n_layer=n_layer, block_size=block, world_size=ws)
    # every tensor is ALIGN-aligned, inside the buffer, and tensors never overlap
    prev = 0
    for s in L:
        assert s.offset % ALIGN == 0 and s.offset >= prev and s.offset + s.numel <= L.total
        prev = s.offset + s.numel
    # parameter count is the closed form of TinyGPT (tied head counted once)
    assert L.num_params() == vocab * d + block * d + n_layer * (12 * d * d + 13 * d) + 2 * d
    # both partitions (whole buffer for ZeRO-2, matrix region for FSDP / ZeRO-3) split evenly into aligned shards
    assert L.total % (ws * ALIGN) == 0 and L.big_end % (ws * ALIGN) == 0
    assert L.shard_size * ws == L.total and L.big_shard_size * ws == L.big_end
    assert all(L.big_shard_range(r) == (r * L.big_shard_size, (r + 1) * L.big_shard_size) for r in range(ws))
    assert all(L.owner_of(L.shard_range(r)[0]) == r and L.owner_of(L.shard_range(r)[1] - 1) == r for r in range(ws))
    # matrices sit in front of big_end, every vector behind it
    for s in L:
        assert (len(s.shape) == 2) == (s.offset < L.big_end)
    assert L.small_range() == (L.big_end, L.total)
    # one block's four matrices are contiguous (one DDP bucket / one gather unit) and blocks do not interleave
    ends = 0
    for l in range(n_layer):
        a, b = L.layer_range(l)
        assert a >= ends and b - a >= 12 * d * d and b <= L.big_end
        inside = [s for s in L if a <= s.offset < b]
        assert len(inside) == 4 and all(s.layer == l for s in inside)
        ends = b
    # state_dict order lists every stored tensor exactly once
    assert sorted(L.state_dict_order()) == sorted(s.name for s in L)


@settings(max_examples=40, deadline=None)
@given(d8=st.integers(1, 16), n_layer=st.integers(1, 5), cap=st.integers(1, 1 << 22), ws=st.sampled_from([1, 2, 4]))
def test_buckets_partition_the_buffer_for_any_cap(d8, n_layer, cap, ws):
    L = build_layout(vocab_size=64, n_embd=8 * d8, n_head=1, n_layer=n_layer, block_size=16, world_size=ws)
    b = L.buckets(cap)
    assert b[0][1] == L.total and b[-1][0] == 0
    assert all(lo < hi for lo, hi in b)
    assert all(nxt[1] == cur[0] for cur, nxt in zip(b, b[1:]))           # reverse order, no gaps, no overlap
    starts = {s.offset for s in L} | {0}
    assert all(lo in starts for lo, _ in b)                              # a bucket never starts inside a tensor
