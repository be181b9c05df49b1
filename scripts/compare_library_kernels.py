#!/usr/bin/env python3
"""Same-shape comparison of the hand-written kernels with the libraries on this box (needs one B200):
cuBLAS through torch.matmul for every GEMM shape of a tier-A training step, and PyTorch's fused
scaled_dot_product_attention (flash / cuDNN back-end) for the attention shape.  Purpose: at batch 1 x 2048
tokens every block GEMM is about one wave of 148 SMs, so "fraction of the 8192^3 cuBLAS rate" understates how
close a kernel is to what the hardware allows at that shape; the library at the SAME shape is the fair yardstick.

    python scripts/compare_library_kernels.py [--tier A] [--seq-len 2048] [--iters 50] > profiles/library_compare.md

Timing: CUDA events around `iters` back-to-back launches after warm-up, operands rotated through enough copies to
exceed the 126 MB L2.  The library calls are measurement-only; nothing in the training path uses them.
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200bench import ops as _ops  # noqa: E402
from b200bench.config import get_model_config  # noqa: E402

BF16 = torch.bfloat16


CLOCKS = []          # one record per timed region: (label, sm_mhz, sm_max_mhz, reasons, power)
_LABEL = ["?"]


def timed(fn, iters, min_ms=250.0):
    """Mean device time per call in us.  The region is stretched to >= min_ms so that the NVML clock sampler
    (20 ms period) sees it under load; its summary is appended to CLOCKS."""
    from b200bench.utils.clocks import ClockSampler

    for i in range(5):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    probe = e0.elapsed_time(e1) / iters                    # ms per call
    n = max(iters, int(min_ms / max(probe, 1e-4)))
    sampler = ClockSampler(0, period=0.02).start()
    e0.record()
    for i in range(n):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    c = sampler.stop()
    CLOCKS.append((_LABEL[0], c.get("sm_mhz"), c.get("sm_max_mhz"), ",".join(c.get("reasons") or []) or "none",
                   c.get("power_w_max")))
    return e0.elapsed_time(e1) / n * 1e3            # us


def copies(bytes_per_set):
    return max(2, int(300e6 // max(1, bytes_per_set)) + 1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tier", default="A")
    ap.add_argument("--seq-len", type=int, default=2048)
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--iters", type=int, default=50)
    a = ap.parse_args()
    cfg = get_model_config(a.tier, a.seq_len)
    d, V, H = cfg["n_embd"], cfg["vocab_size"], cfg["n_head"]
    M = a.batch * a.seq_len
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    o = _ops.load()
    print(f"# Hand-written kernels vs libraries at the same shape (tier {a.tier}, M = {M}, d = {d}; {torch.cuda.get_device_name(0)})\n")
    print("| op | shape (M x N x K) | ours us | ours TFLOP/s | cuBLAS (torch.matmul) us | cuBLAS TFLOP/s | ours / cuBLAS time |")
    print("|---|---|---|---|---|---|---|")
    # (name, kind, rows of x, N = out features, K = in features)
    shapes = [("qkv fwd", "fwd", M, 3 * d, d), ("proj fwd", "fwd", M, d, d), ("fc1 fwd", "fwd", M, 4 * d, d),
              ("fc2 fwd", "fwd", M, d, 4 * d), ("lm_head fwd", "fwd", M, V, d),
              ("qkv dgrad", "dgrad", M, 3 * d, d), ("proj dgrad", "dgrad", M, d, d), ("fc1 dgrad", "dgrad", M, 4 * d, d),
              ("fc2 dgrad", "dgrad", M, d, 4 * d), ("lm_head dgrad", "dgrad", M, V, d),
              ("qkv wgrad", "wgrad", M, 3 * d, d), ("proj wgrad", "wgrad", M, d, d), ("fc1 wgrad", "wgrad", M, 4 * d, d),
              ("fc2 wgrad", "wgrad", M, d, 4 * d), ("lm_head wgrad", "wgrad", M, V, d)]
    for name, kind, m, n, k in shapes:
        nset = copies((m * k + n * k + m * n) * 2)
        x = [torch.randn(m, k, device=dev, dtype=BF16) for _ in range(nset)]
        w = [torch.randn(n, k, device=dev, dtype=BF16) * 0.02 for _ in range(nset)]
        y = [torch.randn(m, n, device=dev, dtype=BF16) for _ in range(nset)]
        dx = [torch.empty(m, k, device=dev, dtype=BF16) for _ in range(nset)]
        dw = [torch.empty(n, k, device=dev, dtype=BF16) for _ in range(nset)]
        _LABEL[0] = name
        if kind == "fwd":
            ours = timed(lambda i: o.linear_fwd(x[i % nset], w[i % nset], y[i % nset]), a.iters)
            lib = timed(lambda i: torch.matmul(x[i % nset], w[i % nset].t(), out=y[i % nset]), a.iters)
        elif kind == "dgrad":
            ours = timed(lambda i: o.linear_dgrad(y[i % nset], w[i % nset], dx[i % nset]), a.iters)
            lib = timed(lambda i: torch.matmul(y[i % nset], w[i % nset], out=dx[i % nset]), a.iters)
        else:
            ours = timed(lambda i: o.linear_wgrad(y[i % nset], x[i % nset], dw[i % nset]), a.iters)
            lib = timed(lambda i: torch.matmul(y[i % nset].t(), x[i % nset], out=dw[i % nset]), a.iters)
        fl = 2.0 * m * n * k
        print(f"| {name} | {m} x {n} x {k} | {ours:.1f} | {fl / ours / 1e6:.0f} | {lib:.1f} | {fl / lib / 1e6:.0f} | {ours / lib:.2f} |")
        del x, w, y, dx, dw
    # ---- attention -------------------------------------------------------------------------------------
    B, T, hd = a.batch, a.seq_len, d // H
    qkv = torch.randn(B * T, 3 * d, device=dev, dtype=BF16)
    out = torch.empty(B * T, d, device=dev, dtype=BF16)
    lse = torch.empty(B * H * T, device=dev, dtype=torch.float32)
    dout = torch.randn(B * T, d, device=dev, dtype=BF16)
    dqkv = torch.empty_like(qkv)
    D = torch.empty(B * H * T, device=dev, dtype=torch.float32)
    dq_acc = torch.empty(B * T, d, device=dev, dtype=torch.float32)
    _LABEL[0] = "attention"
    f_ours = timed(lambda i: o.attention_fwd(qkv, out, lse, B, T, H, hd), a.iters)
    b_ours = timed(lambda i: o.attention_bwd(qkv, out, dout, lse, dqkv, D, dq_acc, B, T, H, hd), a.iters)
    q, k_, v = (t.reshape(B, T, H, hd).transpose(1, 2).contiguous().requires_grad_(True)
                for t in qkv.view(B * T, 3, d).unbind(1))
    sdpa = torch.nn.functional.scaled_dot_product_attention
    f_lib = timed(lambda i: sdpa(q, k_, v), a.iters)
    y = sdpa(q, k_, v)
    g = torch.randn_like(y)
    fb_lib = timed(lambda i: torch.autograd.grad(sdpa(q, k_, v), (q, k_, v), g), a.iters)
    fl = 4.0 * B * H * T * T * hd
    print(f"\n| attention (B {B}, T {T}, H {H}, hd {hd}, no mask, no dropout) | ours us | ours TFLOP/s | PyTorch SDPA us | SDPA TFLOP/s |")
    print("|---|---|---|---|---|")
    print(f"| forward | {f_ours:.1f} | {fl / f_ours / 1e6:.0f} | {f_lib:.1f} | {fl / f_lib / 1e6:.0f} |")
    print(f"| backward (ours) vs forward+backward minus forward (SDPA) | {b_ours:.1f} | {2.5 * fl / b_ours / 1e6:.0f} | "
          f"{fb_lib - f_lib:.1f} | {2.5 * fl / max(1e-3, fb_lib - f_lib) / 1e6:.0f} |")
    print("\nClock record per timed region (NVML, 20 ms period, sampled while the region ran; two regions per GEMM row: "
          "ours, then cuBLAS; four for attention):\n")
    print("| region | sm MHz (median) | sm max MHz | throttle reasons | power W max |")
    print("|---|---|---|---|---|")
    for lab, mhz, mx, rs, pw in CLOCKS:
        print(f"| {lab} | {mhz} | {mx} | {rs} | {pw} |")


if __name__ == "__main__":
    main()
