#!/usr/bin/env python
"""Draws images/architecture_diagram.png: the layer stack of b200bench (counterpart of the reference's
published images/architecture_diagram.png).  Pillow only, no display needed."""
from __future__ import annotations

import argparse

from PIL import Image, ImageDraw, ImageFont

LAYERS = [
    ("Analysis / reporting", "scripts/parse_metrics.py  plot.py  make_report.py  ->  metrics.csv, 5 plots, BENCHMARK_REPORT.md", "#dbe9f6"),
    ("Suite orchestration", "scripts/run_all_benchmarks.sh  validate_multi.sh  launch_multi.sh  collect_results.sh  (k8s/*.yaml, docker/)", "#dbe9f6"),
    ("Launcher", "python -m b200bench.launch: one process per GPU, per-run timeout, rank-failure propagation", "#e3f1df"),
    ("Harness", "b200bench.cli / harness.py: reference CLI + result JSON + stdout markers, device-event timing, --profile", "#e3f1df"),
    ("Engine", "NativeEngine: flat bf16 params + fp32 master, ddp / zero2 / fsdp / zero3, CUDA-graph micro-step, side + comm streams", "#fde9d0"),
    ("Peer-memory runtime", "PeerComm: CUDA-IPC symmetric buffers, flag words, epochs  |  torch.distributed (NCCL) for rendezvous only", "#fde9d0"),
    ("C ABI", "include/b200bench.h  ->  b200bench/lib/libb200bench.so (ctypes)", "#f7dada"),
    ("sm_100a kernels", "GEMM tcgen05+TMEM+TMA (1-CTA and CTA-pair) with fused epilogues and fused all-gather (K14)  |  flash attention fwd/bwd", "#f7dada"),
    ("", "LayerNorm, embedding, cross-entropy, AdamW, reductions  |  K13 all-reduce+scale  |  K15 reduce-scatter+clip+AdamW+push over NVLink", "#f7dada"),
    ("Hardware", "8 x B200 (148 SMs, 180 GB HBM3e each), NVLink 5 / NVSwitch", "#e6e0f3"),
]


def font(size: int):
    try:
        return ImageFont.load_default(size=size)      # Pillow >= 10.1 ships a scalable default face
    except TypeError:
        return ImageFont.load_default()


def main() -> None:
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--out", default="images/architecture_diagram.png")
    a = ap.parse_args()
    W, row_h, pad = 1500, 74, 24
    H = pad * 2 + 60 + row_h * len(LAYERS)
    img = Image.new("RGB", (W, H), "white")
    d = ImageDraw.Draw(img)
    f_title, f_name, f_body = font(30), font(20), font(16)
    d.text((pad, pad), "b200bench: layers (top = user-facing, bottom = hardware)", fill="black", font=f_title)
    y = pad + 60
    for name, body, colour in LAYERS:
        d.rectangle([pad, y, W - pad, y + row_h - 10], fill=colour, outline="#555555", width=2)
        if name:
            d.text((pad + 14, y + 8), name, fill="black", font=f_name)
            d.text((pad + 14, y + 36), body, fill="#222222", font=f_body)
        else:
            d.text((pad + 14, y + 22), body, fill="#222222", font=f_body)
        y += row_h
    img.save(a.out)
    print(f"Saved: {a.out}")


if __name__ == "__main__":
    main()
