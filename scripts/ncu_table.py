#!/usr/bin/env python3
"""One markdown row per kernel launch in an .ncu-rep (captured with `ncu --set full`): duration, grid, registers,
SM / tensor-pipe / issue / DRAM utilisation, L2 traffic and the two most frequent warp states.
Usage: scripts/ncu_table.py gpurun_out/ncu/step.ncu-rep > profiles/round2/ncu_step_table.md"""
import csv
import io
import re
import subprocess
import sys

COLS = [("gpu__time_duration.sum", "us", 1e-3), ("launch__grid_size", "grid", 1), ("launch__block_size", "block", 1),
        ("launch__registers_per_thread", "regs", 1),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM %", 1),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor %", 1),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %", 1),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %", 1),
        ("lts__t_bytes.sum", "L2 MB", 1e-6)]
STALLS = "smsp__pcsamp_warps_issue_stalled_"


def num(v, unit, scale):
    try:
        x = float(v.replace(",", ""))
    except ValueError:
        return v
    # ncu prints durations in ns / us / ms and bytes in byte / Kbyte / Mbyte depending on magnitude
    mult = {"nsecond": 1e-3 / 1e-3, "usecond": 1e3 * 1e-3 / 1e-3, "msecond": 1e6, "second": 1e9,
            "byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
    x = x * mult * scale
    return f"{x:.1f}" if abs(x) < 1000 and x != int(x) else f"{x:.0f}"


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    print(f"# ncu --set full, one row per captured launch: {rep.split('/')[-1]}\n")
    print("| # | kernel | " + " | ".join(c[1] for c in COLS) + " | top warp states |")
    print("|---|---|" + "---|" * (len(COLS) + 1))
    for n, r in enumerate(rows[2:]):
        name = re.sub(r"\(.*", "", r[ix["Kernel Name"]]) if "Kernel Name" in ix else "?"
        name = re.sub(r"^(void )?b200::", "", name)
        cells = [num(r[ix[k]], units[ix[k]], sc) if k in ix else "-" for k, _, sc in COLS]
        st = sorted(((float(r[i] or 0), h[len(STALLS):]) for h, i in ix.items()
                     if h.startswith(STALLS) and not h.endswith("_not_issued")), reverse=True)
        tot = sum(s for s, _ in st) or 1
        top = ", ".join(f"{nm} {100 * s / tot:.0f}%" for s, nm in st[:2])
        print(f"| {n} | `{name}` | " + " | ".join(cells) + f" | {top} |")


if __name__ == "__main__":
    main()
