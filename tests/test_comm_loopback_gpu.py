"""The fused collectives on ONE GPU: a single cooperative launch plays all ranks (gridDim.y = ranks) over
per-"rank" buffers, so the flag protocol, the reductions, the clipping norm exchange, AdamW and the weight push
are checked where only one GPU exists (the multi-process path is tests/multi/check_comm.py)."""
import ctypes
import math
from ctypes import c_float, c_int, c_size_t, c_void_p

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"
BF16 = torch.bfloat16


def _lib(native):
    L = native.lib
    vp, i, f, sz = c_void_p, c_int, c_float, c_size_t
    L.b200_comm_flag_bytes.restype = sz
    L.b200_comm_allreduce_loop.argtypes = [vp, vp, i, sz, sz, f, i, vp]
    L.b200_comm_rs_adamw_loop.argtypes = [vp, vp, vp, i, vp, sz, vp, vp, vp, vp, vp, i, vp, vp, i, vp]
    L.b200_comm_gather_range.argtypes = [vp, i, vp, ctypes.c_ulonglong, ctypes.c_ulonglong, ctypes.c_ulonglong, i, vp]
    L.b200_comm_gather_range_from.argtypes = [vp, i, i, vp, ctypes.c_ulonglong, ctypes.c_ulonglong,
                                              ctypes.c_ulonglong, i, vp]
    return L


def _table(tensors):
    return (c_void_p * len(tensors))(*[t.data_ptr() if t is not None else None for t in tensors])


def _stream():
    return c_void_p(torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize("ws", [2, 4, 8])
def test_allreduce_loopback(native, ws):
    L = _lib(native)
    torch.manual_seed(20 + ws)
    n_total, off, n = 1 << 16, 1024, 40000
    flags = [torch.zeros(int(L.b200_comm_flag_bytes()), dtype=torch.uint8, device=DEV) for _ in range(ws)]
    for epoch in range(3):                       # back-to-back epochs reuse the flag words
        grads = [(torch.randn(n_total, device=DEV) * (r + 1)).to(BF16) for r in range(ws)]
        before = [g.clone() for g in grads]
        ref = sum(g.float()[off:off + n] for g in grads) / ws
        rc = L.b200_comm_allreduce_loop(_table(grads), _table(flags), ws, off, n, 1.0 / ws, 4, _stream())
        assert rc == 0, rc
        torch.cuda.synchronize()
        for r in range(ws):
            got = grads[r].float()
            assert torch.allclose(got[off:off + n], ref, atol=0.02 + 0.01 * float(ref.abs().max()), rtol=2e-2), (epoch, r)
            assert torch.equal(grads[r][:off], before[r][:off]) and torch.equal(grads[r][off + n:], before[r][off + n:])
        for r in range(1, ws):                    # every rank holds bit-identical results
            assert torch.equal(grads[r][off:off + n], grads[0][off:off + n])


@pytest.mark.parametrize("ws,clip,push_all", [(2, 0.0, True), (4, 0.0, False), (4, 1.0, True), (8, 0.5, True)])
def test_rs_adamw_loopback(native, ws, clip, push_all):
    """K15: pull-reduce of the own partition + (global-norm clip) + AdamW + bf16 weight push, all ranks in one launch,
    against an fp32 reduction and torch.optim.AdamW."""
    L = _lib(native)
    torch.manual_seed(30 + ws)
    n = 8 * 700                                  # partition size (elements)
    total = n * ws
    flags = [torch.zeros(int(L.b200_comm_flag_bytes()), dtype=torch.uint8, device=DEV) for _ in range(ws)]
    w0 = torch.randn(total, device=DEV) * 0.1
    weights = [w0.to(BF16).clone() for _ in range(ws)]
    master = [w0[r * n:(r + 1) * n].clone() for r in range(ws)]
    m = [torch.zeros(n, device=DEV) for _ in range(ws)]
    v = [torch.zeros(n, device=DEV) for _ in range(ws)]
    gshard = [torch.zeros(n, dtype=BF16, device=DEV) for _ in range(ws)]
    gnorm = [torch.zeros(2, device=DEV) for _ in range(ws)]           # TWO floats per rank
    hp = [torch.zeros(16, device=DEV) for _ in range(ws)]
    p_ref = torch.nn.Parameter(w0.clone())
    opt = torch.optim.AdamW([p_ref], lr=1e-2, weight_decay=0.01)
    lo = (c_size_t * ws)(*[r * n for r in range(ws)])
    for step in range(1, 4):
        grads = [(torch.randn(total, device=DEV) * 0.5).to(BF16) for _ in range(ws)]
        gsum = sum(g.float() for g in grads) / ws
        norm = float(gsum.norm())
        coef = min(1.0, clip / (norm + 1e-6)) if clip > 0 else 1.0
        p_ref.grad = gsum * coef
        opt.step()
        for r in range(ws):
            native.adamw_set_hp(hp[r], 1e-2, 0.9, 0.999, 1e-8, 0.01, 1 / (1 - 0.9 ** step),
                                1 / math.sqrt(1 - 0.999 ** step), 1.0 / ws, clip)
        rc = L.b200_comm_rs_adamw_loop(_table(grads), _table(weights), _table(flags), ws, lo, n, _table(master),
                                       _table(m), _table(v), _table(gshard), _table(hp), int(push_all),
                                       _table(gnorm), None, 4, _stream())
        assert rc == 0, rc
        torch.cuda.synchronize()
        if clip > 0:
            for r in range(ws):
                assert abs(float(gnorm[r][0]) - norm) < 2e-2 * norm, (float(gnorm[r][0]), norm)
    full_master = torch.cat(master)
    assert torch.allclose(full_master, p_ref.data, atol=2e-3, rtol=2e-3), float((full_master - p_ref.data).abs().max())
    for r in range(ws):
        own = slice(r * n, (r + 1) * n)
        assert torch.allclose(weights[r].float()[own], p_ref.data[own], atol=4e-3, rtol=1e-2)
        if push_all:                             # ZeRO-2: every rank received every partition
            assert torch.allclose(weights[r].float(), p_ref.data, atol=4e-3, rtol=1e-2)
            assert torch.equal(weights[r], weights[0])


def test_collectives_stress_loopback(native):
    """Stand-in for racecheck on the flag protocol: many back-to-back epochs of K13 with random ranges and
    CTA counts on two alternating streams; every result is checked, so a lost or early flag shows up as a wrong sum."""
    L = _lib(native)
    torch.manual_seed(99)
    ws, n_total = 4, 1 << 15
    flags = [torch.zeros(int(L.b200_comm_flag_bytes()), dtype=torch.uint8, device=DEV) for _ in range(ws)]
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    gen = torch.Generator().manual_seed(5)
    prev = None
    for it in range(40):
        off = int(torch.randint(0, 64, (1,), generator=gen)) * 8
        n = int(torch.randint(1, (n_total - off) // 8, (1,), generator=gen)) * 8
        ctas = int(torch.randint(1, 9, (1,), generator=gen))
        st = streams[it & 1]
        if prev is not None:
            st.wait_event(prev)                  # collectives sharing flag words must be stream-ordered
        with torch.cuda.stream(st):
            grads = [torch.randn(n_total, device=DEV).to(BF16) for _ in range(ws)]
            ref = sum(g.float()[off:off + n] for g in grads)
            rc = L.b200_comm_allreduce_loop(_table(grads), _table(flags), ws, off, n, 1.0, ctas, _stream())
            assert rc == 0, rc
            ok = torch.stack([(g.float()[off:off + n] - ref).abs().max() for g in grads]).max()
            prev = torch.cuda.Event()
            prev.record(st)
        assert float(ok) < 0.05 + 0.02 * float(ref.abs().max()), (it, off, n, ctas, float(ok))


def test_gather_range_single_process(native):
    """gather_range has no waits: each emulated rank can simply be called in turn."""
    L = _lib(native)
    ws, shard = 4, 8 * 512
    full = torch.randn(ws * shard, device=DEV).to(BF16)
    shards = [full[r * shard:(r + 1) * shard].clone() for r in range(ws)]
    adj = (c_void_p * ws)(*[shards[r].data_ptr() - r * shard * 2 for r in range(ws)])
    off, n = 8 * 100, 8 * 1200
    dst = torch.zeros(n, dtype=BF16, device=DEV)
    rc = L.b200_comm_gather_range(adj, ws, c_void_p(dst.data_ptr()), off, n, shard, 8, _stream())
    assert rc == 0
    assert torch.equal(dst, full[off:off + n])
    for rank in range(ws):                       # each rank starts at a different owner and wraps around
        for n2 in (n, 8 * 1201, 8 * 3):
            dst = torch.zeros(n2, dtype=BF16, device=DEV)
            rc = L.b200_comm_gather_range_from(adj, ws, rank, c_void_p(dst.data_ptr()), off, n2, shard, 3, _stream())
            assert rc == 0
            assert torch.equal(dst, full[off:off + n2]), (rank, n2)


@pytest.mark.parametrize("ws", [2, 4, 8])
def test_gather_bcast_loopback(native, ws):
    """K16 (peer-store path): every rank fetches 1/ws of the range from its owner(s) and writes it into every rank's
    destination copy; back-to-back epochs reuse flag plane 2; bytes outside the destination range stay untouched."""
    L = _lib(native)
    ull = ctypes.c_ulonglong
    L.b200_comm_gather_bcast_loop.argtypes = [c_void_p, c_void_p, c_void_p, c_int, ull, ull, ull, ull, c_int, c_void_p]
    torch.manual_seed(70 + ws)
    shard = 8 * 640
    flags = [torch.zeros(int(L.b200_comm_flag_bytes()), dtype=torch.uint8, device=DEV) for _ in range(ws)]
    for epoch, (off, n, dst_off, ctas) in enumerate([(8 * 100, 8 * 1100, 8 * 16, 4), (0, 8 * 3, 0, 2),
                                                     (8 * 630, 8 * 601, 8 * 64, 8), (8 * 5, shard * ws - 8 * 5, 0, 3)]):
        full = torch.randn(ws * shard, device=DEV).to(BF16)
        shards = [full[r * shard:(r + 1) * shard].clone() for r in range(ws)]
        adj = (c_void_p * ws)(*[shards[r].data_ptr() - r * shard * 2 for r in range(ws)])
        dsts = [torch.full((dst_off + n + 64,), float(r + 1), dtype=BF16, device=DEV) for r in range(ws)]
        rc = L.b200_comm_gather_bcast_loop(adj, _table(dsts), _table(flags), ws, off, dst_off, n, shard, ctas, _stream())
        assert rc == 0, rc
        torch.cuda.synchronize()
        for r in range(ws):
            assert torch.equal(dsts[r][dst_off:dst_off + n], full[off:off + n]), (epoch, r)
            assert bool((dsts[r][:dst_off] == r + 1).all()) and bool((dsts[r][dst_off + n:] == r + 1).all()), (epoch, r)
