#!/usr/bin/env python3
"""One markdown row per kernel launch in an .ncu-rep (captured with `ncu --set full`): duration, grid, registers,
SM / tensor-pipe / issue / DRAM utilisation, L2 traffic and the two most frequent warp states.
Usage: scripts/ncu_table.py gpurun_out/ncu/step.ncu-rep > profiles/round2/ncu_step_table.md"""
import csv
import io
import re
import subprocess
import sys

COLS = [("gpu__time_duration.sum", "us", "time"), ("launch__grid_size", "grid", ""), ("launch__block_size", "block", ""),
        ("launch__registers_per_thread", "regs", ""),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM %", ""),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor %", ""),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %", ""),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %", ""),
        ("lts__t_bytes.sum", "L2 MB", "bytes")]
STALLS = "smsp__pcsamp_warps_issue_stalled_"
# ncu prints a value in whatever unit fits its magnitude
TO_US = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "s": 1e6, "second": 1e6}
TO_MB = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}


def num(v, unit, kind):
    try:
        x = float(v.replace(",", ""))
    except ValueError:
        return v
    if kind == "time":
        x *= TO_US.get(unit, 1.0)
    elif kind == "bytes":
        x *= TO_MB.get(unit, 1e-6)
    return f"{x:.0f}" if x == int(x) or abs(x) >= 1000 else f"{x:.1f}"


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
