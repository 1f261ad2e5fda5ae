#!/usr/bin/env python
"""Pipe utilisation and warp-stall breakdown of the launches in an .ncu-rep (run here, no GPU needed).

  python tools/ncu_stalls.py gpurun_out/prof.ncu-rep [profiles/NAME.md]
"""
import csv
import io
import subprocess
import sys

KEEP = ("gpu__time_duration.sum", "smsp__issue_active.avg.pct", "smsp__inst_executed.sum", "sm__inst_executed_pipe_",
        "sm__pipe_", "smsp__average_warps_issue_stalled", "smsp__average_warp_latency_issue_stalled", "dram__throughput.avg.pct",
        "gpu__dram_throughput", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__warps_active.avg",
        "sm__warps_active.avg.pct", "launch__registers_per_thread", "launch__grid_size", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "smsp__inst_issued.avg.per_cycle_active", "sm__throughput.avg.pct", "l1tex__throughput.avg.pct", "lts__throughput.avg.pct")


def main():
    src = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    out = [f"# ncu --set full, pipes and stalls: {src}\n"]
    for n, r in enumerate(rows[2:]):
        out.append(f"\n## launch {n}: `{r[hdr.index('Kernel Name')][:100]}`\n\n| metric | value | unit |\n|---|---:|---|")
        items = []
        for i, k in enumerate(hdr):
            if not any(t in k for t in KEEP):
                continue
            try:
                v = float(r[i].replace(",", ""))
            except ValueError:
                continue
            if v == 0:
                continue
            items.append((k, v, units[i]))
        for k, v, u in items:
            if "issue_stalled" in k and "ratio" not in k and "pct" not in k:
                continue
            out.append(f"| {k} | {v:.4g} | {u} |")
    text = "\n".join(out) + "\n"
    if len(sys.argv) > 2:
        open(sys.argv[2], "w").write(text)
    print(text)


if __name__ == "__main__":
    main()
