"""metrics.csv -> the reference's five PNG charts (``scripts/plot.py:13-109``): same file names,
one line per strategy, 10x6 in at 150 dpi (1500x900 px), ``vram_vs_seqlen.png`` only when more than
one sequence length is present, dashed 100 % line and 0-110 y-range on the efficiency chart.

matplotlib (Agg) is used when importable; otherwise a small PIL line-chart renderer produces the
same charts, because the target image has no matplotlib.  Both are CPU rasterisers.
"""
from __future__ import annotations

import argparse
import math
from pathlib import Path
from typing import Callable, Dict, List, Optional, Tuple

import pandas as pd

W, H, DPI = 1500, 900, 150
PALETTE = [(31, 119, 180), (255, 127, 14), (44, 160, 44), (214, 39, 40), (148, 103, 189), (140, 86, 75)]
Series = Dict[str, Tuple[List[float], List[float]]]


def _nice_ticks(lo: float, hi: float, n: int = 6) -> List[float]:
    if hi <= lo:
        hi = lo + 1.0
    raw = (hi - lo) / max(1, n)
    mag = 10 ** math.floor(math.log10(raw))
    step = min((m for m in (1, 2, 2.5, 5, 10) if m * mag >= raw), default=10) * mag
    t0 = math.floor(lo / step) * step
    ticks, t = [], t0
    while t <= hi + step * 0.5:
        ticks.append(round(t, 10))
        t += step
    return ticks


def _fmt(v: float) -> str:
    if v == 0:
        return "0"
    a = abs(v)
    if a >= 1e6:
        return f"{v/1e6:g}M"
    if a >= 1e4:
        return f"{v/1e3:g}k"
    if a < 1e-3:
        return f"{v:.1e}"
    return f"{v:g}"


def _pil_chart(series: Series, title: str, xlabel: str, ylabel: str, path: Path,
               ylim: Optional[Tuple[float, float]] = None, hline: Optional[float] = None) -> None:
    from PIL import Image, ImageDraw, ImageFont

    img = Image.new("RGB", (W, H), "white")
    d = ImageDraw.Draw(img)
    try:
        font = ImageFont.load_default(size=22)
        small = ImageFont.load_default(size=18)
        big = ImageFont.load_default(size=28)
    except TypeError:
        font = small = big = ImageFont.load_default()
    L, R, T, B = 150, 60, 90, 110
    xs = sorted({x for sx, _ in series.values() for x in sx})
    ys = [y for _, sy in series.values() for y in sy]
    if hline is not None:
        ys.append(hline)
    ylo, yhi = (min(ys + [0.0]), max(ys) * 1.08) if ylim is None else ylim
    yt = _nice_ticks(ylo, yhi)
    if ylim is None:
        ylo, yhi = yt[0], yt[-1]
    xlo, xhi = (xs[0], xs[-1]) if len(xs) > 1 else (xs[0] - 1, xs[0] + 1)
    pad = (xhi - xlo) * 0.05
    xlo, xhi = xlo - pad, xhi + pad

    def X(v: float) -> float:
        return L + (v - xlo) / (xhi - xlo) * (W - L - R)

    def Y(v: float) -> float:
        return H - B - (v - ylo) / (yhi - ylo) * (H - T - B)

    for t in yt:
        if ylo <= t <= yhi:
            d.line([(L, Y(t)), (W - R, Y(t))], fill=(225, 225, 225), width=1)
            d.text((L - 12, Y(t)), _fmt(t), fill="black", font=small, anchor="rm")
    for x in xs:
        d.line([(X(x), T), (X(x), H - B)], fill=(235, 235, 235), width=1)
        d.text((X(x), H - B + 14), _fmt(x), fill="black", font=small, anchor="mt")
    d.rectangle([L, T, W - R, H - B], outline="black", width=2)
    if hline is not None:
        x = L
        while x < W - R:                                  # dashed "ideal" line
            d.line([(x, Y(hline)), (min(x + 14, W - R), Y(hline))], fill=(120, 120, 120), width=2)
            x += 24
    for i, (name, (sx, sy)) in enumerate(series.items()):
        col = PALETTE[i % len(PALETTE)]
        pts = [(X(a), Y(b)) for a, b in zip(sx, sy)]
        if len(pts) > 1:
            d.line(pts, fill=col, width=4)
        for px, py in pts:
            d.ellipse([px - 7, py - 7, px + 7, py + 7], fill=col)
        ly = T + 14 + i * 30
        d.line([(W - R - 230, ly + 10), (W - R - 190, ly + 10)], fill=col, width=4)
        d.text((W - R - 180, ly), name, fill="black", font=small)
    d.text((W // 2, 36), title, fill="black", font=big, anchor="mm")
    d.text((W // 2, H - 36), xlabel, fill="black", font=font, anchor="mm")
    lab = Image.new("RGBA", (H, 40), (255, 255, 255, 0))
    ImageDraw.Draw(lab).text((H // 2, 20), ylabel, fill="black", font=font, anchor="mm")
    img.paste(lab.rotate(90, expand=True), (18, 0), lab.rotate(90, expand=True))
    img.save(path, dpi=(DPI, DPI))


def _mpl_chart(series: Series, title: str, xlabel: str, ylabel: str, path: Path,
               ylim: Optional[Tuple[float, float]] = None, hline: Optional[float] = None) -> None:
    import matplotlib

    matplotlib.use("Agg")
    import matplotlib.pyplot as plt

    fig, ax = plt.subplots(figsize=(W / DPI, H / DPI))
    for name, (sx, sy) in series.items():
        ax.plot(sx, sy, marker="o", linewidth=2, label=name)
    if hline is not None:
        ax.axhline(hline, linestyle="--", color="gray", label="Ideal")
    if ylim is not None:
        ax.set_ylim(*ylim)
    ax.set_xlabel(xlabel)
    ax.set_ylabel(ylabel)
    ax.set_title(title)
    ax.grid(True, alpha=0.3)
    ax.legend()
    fig.tight_layout()
    fig.savefig(path, dpi=DPI)
    plt.close(fig)


def _renderer() -> Callable[..., None]:
    try:
        import matplotlib  # noqa: F401

        return _mpl_chart
    except Exception:
        return _pil_chart


def _by_strategy(df: pd.DataFrame, x: str, y: str) -> Series:
    out: Series = {}
    for strat, sub in df.groupby("strategy", sort=True):
        sub = sub.sort_values(x)
        out[str(strat).upper()] = (sub[x].astype(float).tolist(), sub[y].astype(float).tolist())
    return out


def plot_metrics(csv_path: str, output_dir: str) -> List[Path]:
    df = pd.read_csv(csv_path)
    out = Path(output_dir)
    out.mkdir(parents=True, exist_ok=True)
    draw = _renderer()
    made: List[Path] = []
    # one curve per strategy at the most common sequence length for the per-GPU charts
    main_seq = df["seq_len"].mode().iloc[0]
    per_gpu = df[df["seq_len"] == main_seq]

    def emit(name: str, series: Series, title: str, xl: str, yl: str, **kw) -> None:
        p = out / name
        draw(series, title, xl, yl, p, **kw)
        made.append(p)
        print(f"Saved: {p}")

    emit("tokens_per_sec_vs_gpu.png", _by_strategy(per_gpu, "world_size", "tokens_per_sec"),
         "Throughput vs Number of GPUs", "Number of GPUs", "Tokens/sec")
    emit("step_time_vs_gpu.png", _by_strategy(per_gpu, "world_size", "mean_step_time_sec"),
         "Step Time vs Number of GPUs", "Number of GPUs", "Mean Step Time (s)")
    if df["seq_len"].nunique() > 1:
        widest = df[df["world_size"] == df["world_size"].max()]
        emit("vram_vs_seqlen.png", _by_strategy(widest, "seq_len", "peak_vram_gb"),
             "Peak VRAM vs Sequence Length", "Sequence Length", "Peak VRAM (GB)")
    emit("scaling_efficiency.png", _by_strategy(per_gpu, "world_size", "scaling_efficiency_pct"),
         "Scaling Efficiency vs Number of GPUs", "Number of GPUs", "Scaling Efficiency (%)",
         ylim=(0.0, 110.0), hline=100.0)
    emit("gbps_vs_gpu.png", _by_strategy(per_gpu, "world_size", "h2d_gbps_per_gpu"),
         "H2D Transfer Rate vs Number of GPUs", "Number of GPUs", "H2D GB/s per GPU")
    return made


def main(argv=None) -> None:
    ap = argparse.ArgumentParser(description="Generate plots from metrics CSV")
    ap.add_argument("--results", required=True, help="Path to metrics.csv")
    ap.add_argument("--out", required=True, help="Output directory for plots")
    a = ap.parse_args(argv)
    plot_metrics(a.results, a.out)


if __name__ == "__main__":
    main()
