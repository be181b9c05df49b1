#!/usr/bin/env python3
"""Summarise an .ncu-rep (captured with `ncu --set full --import-source on`) into markdown: headline
metrics from the raw page, opcode mix and the hottest source lines from the source page.
Usage: scripts/ncu_summary.py gpurun_out/prof_x.ncu-rep [launch index in the report, default 0] > profiles/x_ncu.md"""
import collections
import csv
import io
import re
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.max", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed.sum"]
STALLS = "smsp__pcsamp_warps_issue_stalled_"


SKIP = int(sys.argv[2]) if len(sys.argv) > 2 else 0


def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--launch-skip", str(SKIP), "--launch-count", "1", "--page", name, "--csv"],
                         capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    raw = page(rep, "raw")
    hdr, units, vals = raw[0], raw[1], raw[2]
    d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
    print(f"# ncu summary: {rep.split('/')[-1]}" + (f", launch {SKIP}" if SKIP else "") + "\n")
    print(f"Kernel: `{d.get('Kernel Name', ('?',))[0]}`\n")
    print("| metric | value | unit |\n|---|---|---|")
    for k in WANT:
        if k in d:
            print(f"| {k} | {d[k][0]} | {d[k][1]} |")
    stalls = sorted(((int(float(v[0])), k[len(STALLS):]) for k, v in d.items()
                     if k.startswith(STALLS) and not k.endswith("_not_issued")), reverse=True)
    tot = sum(s for s, _ in stalls) or 1
    print("\nWarp-state samples: " + ", ".join(f"{n} {100*s/tot:.0f}%" for s, n in stalls[:8]))
    src = page(rep, "source")
    h = src[1]
    ia, isamp, iex = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
    ops = collections.Counter()
    samp = collections.Counter()
    rows = [r for r in src[2:] if len(r) > max(ia, isamp, iex)]
    for r in rows:
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[ia])
        o = m.group(2).split(".")[0] if m else "?"
        try:
            ops[o] += int(r[iex] or 0)
            samp[o] += int(r[isamp] or 0)
        except ValueError:        # a repeated header row (several source views in one page)
            continue
    ti, ts = sum(ops.values()) or 1, sum(samp.values()) or 1
    print("\nOpcode mix (executed warp instructions / stall samples):\n")
    print("| opcode | inst % | samples % |\n|---|---|---|")
    for o, n in ops.most_common(14):
        print(f"| {o} | {100*n/ti:.1f} | {100*samp[o]/ts:.1f} |")
    rows = [r for r in rows if (r[isamp] or "0").isdigit()]
    hot = sorted(rows, key=lambda r: -int(r[isamp] or 0))[:10]
    print("\nHottest SASS lines by samples:\n")
    for r in hot:
        print(f"* {int(r[isamp] or 0):5d}  `{r[ia].strip()[:100]}`")


if __name__ == "__main__":
    main()
