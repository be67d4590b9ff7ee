#!/usr/bin/env python
"""Summarises an .ncu-rep (read here, no GPU needed) into the handful of numbers the roofline
discussion uses.  python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/xyz.md"""
import csv
import io
import os
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("sm__cycles_elapsed.avg.per_second", "SM clock"),
    ("sm__cycles_active.avg", "SM active cycles"),
    ("sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg", "tensor (HMMA) pipe active cycles, sum over 4 sub-pipes"),
    ("sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "tensor pipe active % (realtime)"),
    ("sm__inst_executed_pipe_tensor_subpipe_hmma.sum", "UTCHMMA instructions"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput % of peak"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate"),
    ("l1tex__m_xbar2l1tex_read_bytes.sum", "L2->SM bytes"),
    ("l1tex__m_xbar2l1tex_read_bytes.sum.pct_of_peak_sustained_elapsed", "L2->SM % of peak"),
    ("launch__registers_per_thread", "registers/thread"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem/block"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    print(f"# ncu summary of `{os.path.basename(rep)}` (`ncu --set full --clock-control none --import-source on`)\n")
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")]
        print(f"## {name[:110]}\n")
        print("| metric | value |\n|---|---|")
        for key, label in KEYS:
            for i, h in enumerate(hdr):
                if h == key or h.endswith("." + key):
                    print(f"| {label} (`{key}`) | {r[i]} {units[i]} |")
                    break
        print()
    # top stall lines from the source page
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    heads = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
    if heads:
        h = rows[heads[0]]
        end = heads[1] - 1 if len(heads) > 1 else len(rows)
        data = [r for r in rows[heads[0] + 1:end] if len(r) > h.index("# Samples")]
        si, so = h.index("# Samples"), h.index("Source")
        tot = sum(int(r[si] or 0) for r in data)
        print(f"## warp-sample hot spots (first kernel, {tot} samples; waiting warps dominate by design)\n")
        print("| samples | SASS |\n|---|---|")
        for r in sorted(data, key=lambda r: -int(r[si] or 0))[:12]:
            print(f"| {r[si]} | `{r[so][:100]}` |")


if __name__ == "__main__":
    main()
