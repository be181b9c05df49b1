#!/usr/bin/env python3
"""Device times of the bandwidth-bound kernels and fused GEMM epilogues at the tier-A shapes (one B200).

    python scripts/bench_small_kernels.py > profiles/round2/small_kernels.md

CUDA events around back-to-back launches after warm-up; operands rotate through > 126 MB so nothing is served
from L2; an NVML clock record (20 ms period) is printed per row.
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200bench import ops as _ops  # noqa: E402
from b200bench.utils.clocks import ClockSampler  # noqa: E402

BF16 = torch.bfloat16


def timed(fn, min_ms=200.0):
    for i in range(5):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(20):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    n = max(20, int(min_ms / max(e0.elapsed_time(e1) / 20, 1e-4)))
    s = ClockSampler(0, period=0.02).start()
    e0.record()
    for i in range(n):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    c = s.stop()
    return e0.elapsed_time(e1) / n * 1e3, f"{c.get('sm_mhz')} MHz, {','.join(c.get('reasons') or []) or 'no throttle'}"


def main():
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    o = _ops.load()
    M, d, H, T, B = 2048, 1024, 16, 2048, 1
    NS = 12                                                        # rotating operand sets
    r = lambda *s: [torch.randn(*s, device=dev, dtype=BF16) for _ in range(NS)]  # noqa: E731
    rows = []

    def row(name, fn, bytes_moved=None):
        us, clk = timed(fn)
        bw = f"{bytes_moved / us / 1e6:.2f} TB/s" if bytes_moved else ""
        rows.append(f"| {name} | {us:.1f} | {bw} | {clk} |")

    dy, x, dres, dx, dxd = r(M, d), r(M, d), r(M, d), r(M, d), r(M, d)
    g = torch.randn(d, device=dev, dtype=BF16)
    mean = torch.randn(M, device=dev)
    rstd = torch.rand(M, device=dev) + 0.5
    dg, db = torch.zeros(d, device=dev, dtype=BF16), torch.zeros(d, device=dev, dtype=BF16)
    sc_old = torch.zeros(o.colreduce_scratch_bytes(M, 4 * d, 2) // 4 + 64, device=dev)
    sc_new = torch.zeros(o.layernorm_bwd_fused_scratch_bytes(M, d) // 4, device=dev)
    row("layernorm_bwd (dx + dres)", lambda i: o.layernorm_bwd(dy[i % NS], x[i % NS], g, mean, rstd, dx[i % NS], dres=dres[i % NS]),
        4 * M * d * 2)
    row("layernorm_bwd_params (old, separate)", lambda i: o.layernorm_bwd_params(dy[i % NS], x[i % NS], mean, rstd, dg, db, sc_old),
        2 * M * d * 2)
    row("layernorm_bwd_fused: dx only", lambda i: o.layernorm_bwd_fused(dy[i % NS], x[i % NS], g, mean, rstd, dx[i % NS], dres=dres[i % NS]),
        4 * M * d * 2)
    row("layernorm_bwd_fused: dx + partial column sums", lambda i: o.layernorm_bwd_fused(
        dy[i % NS], x[i % NS], g, mean, rstd, dx[i % NS], dres=dres[i % NS], partials=sc_new), 4 * M * d * 2)
    row("layernorm_bwd_fused: dx + partials + dropout copy", lambda i: o.layernorm_bwd_fused(
        dy[i % NS], x[i % NS], g, mean, rstd, dx[i % NS], dres=dres[i % NS], dx_drop=dxd[i % NS], drop_p=0.1, seed=7,
        partials=sc_new), 5 * M * d * 2)
    row("layernorm_param_fold (side stream)", lambda i: o.layernorm_param_fold(sc_new, M, d, dg, db))
    row("dropout_apply (old, separate)", lambda i: o.dropout_apply(dx[i % NS], dxd[i % NS], M, d, 0.1, 7), 2 * M * d * 2)
    big = r(M, 4 * d)
    bias4 = torch.zeros(4 * d, device=dev, dtype=BF16)
    row("colsum [M, 4d] (old, separate)", lambda i: o.colsum(big[i % NS], bias4, sc_old), M * 4 * d * 2)
    # ---- GEMM epilogues ------------------------------------------------------------------------------------
    w_dd = r(d, d)
    att, datt = r(M, d), r(M, d)
    rd = torch.zeros(B * H * T, device=dev)
    row("proj dgrad, plain", lambda i: o.linear_dgrad(dy[i % NS], w_dd[i % NS], datt[i % NS]))
    row("proj dgrad + ROWDOT (attention D)", lambda i: o.linear_dgrad(dy[i % NS], w_dd[i % NS], datt[i % NS], epi=_ops.EPI_ROWDOT,
                                                                   aux=att[i % NS], rowdot=rd, rd_T=T, rd_hd=64))
    w41 = [torch.randn(4 * d, d, device=dev, dtype=BF16) * 0.02 for _ in range(NS)]
    pre, act = r(M, 4 * d), r(M, 4 * d)
    row("fc1 fwd, bias only", lambda i: o.linear_fwd(x[i % NS], w41[i % NS], pre[i % NS], bias=bias4))
    row("fc1 fwd, bias + GELU dual store", lambda i: o.linear_fwd(x[i % NS], w41[i % NS], pre[i % NS], bias=bias4,
                                                                  epi=_ops.EPI_BIAS_GELU, out2=act[i % NS]))
    w14 = [torch.randn(d, 4 * d, device=dev, dtype=BF16) * 0.02 for _ in range(NS)]
    row("fc2 dgrad, plain", lambda i: o.linear_dgrad(dy[i % NS], w14[i % NS], big[i % NS]))
    row("fc2 dgrad + DGELU", lambda i: o.linear_dgrad(dy[i % NS], w14[i % NS], big[i % NS], epi=_ops.EPI_DGELU, aux=pre[i % NS]))
    row("fc2 fwd, bias + dropout + residual", lambda i: o.linear_fwd(act[i % NS], w14[i % NS], dx[i % NS], bias=g, epi=_ops.EPI_BIAS_RESID,
                                                                     aux=x[i % NS], drop_p=0.1, seed=3))
    dw = [torch.zeros(4 * d, d, device=dev, dtype=BF16) for _ in range(NS)]
    row("fc1 wgrad", lambda i: o.linear_wgrad(big[i % NS], x[i % NS], dw[i % NS]))
    row("fc1 wgrad + bias gradient", lambda i: o.linear_wgrad(big[i % NS], x[i % NS], dw[i % NS], dbias=bias4))
    # ---- attention -----------------------------------------------------------------------------------------
    qkv, out, dqkv = r(M, 3 * d), r(M, d), r(M, 3 * d)
    lse = torch.zeros(B * H * T, device=dev)
    D = torch.zeros(B * H * T, device=dev)
    dq_acc = torch.zeros(M, d, device=dev)
    for p in (0.0, 0.1):
        row(f"attention fwd, dropout {p}", lambda i: o.attention_fwd(qkv[i % NS], out[i % NS], lse, B, T, H, 64, False, p, 5))
        row(f"attention bwd (4 launches), dropout {p}", lambda i: o.attention_bwd(qkv[i % NS], out[i % NS], datt[i % NS], lse, dqkv[i % NS],
                                                                                 D, dq_acc, B, T, H, 64, False, p, 5))
        row(f"attention bwd (2 launches: D given, accumulator pre-zeroed), dropout {p}",
            lambda i: o.attention_bwd(qkv[i % NS], out[i % NS], datt[i % NS], lse, dqkv[i % NS], D, dq_acc, B, T, H, 64, False, p, 5,
                                      have_d=True, acc_is_zero=True))
    # ---- optimizer -----------------------------------------------------------------------------------------
    n = 12 * d * d
    master, m, v = (torch.zeros(n, device=dev) for _ in range(3))
    grad, wout = torch.zeros(n, device=dev, dtype=BF16), torch.zeros(n, device=dev, dtype=BF16)
    hp = torch.zeros(16, device=dev)
    o.adamw_set_hp(hp, 1e-4, 0.9, 0.999, 1e-8, 0.01, 1.0, 1.0, 1.0, 0.0)
    row("AdamW, one transformer block (12.6 M params)", lambda i: o.adamw_flat(master, m, v, grad, wout, hp), n * 28)
    print(f"# Small kernels and fused epilogues at tier-A shapes (M = {M}, d = {d}; {torch.cuda.get_device_name(0)})\n")
    print("| kernel | us | achieved | clocks during the timed region |")
    print("|---|---|---|---|")
    print("\n".join(rows))


if __name__ == "__main__":
    main()
