#!/usr/bin/env python
"""One table row per kernel launch of an .ncu-rep (read here, no GPU needed): duration, DRAM traffic and
rate as a fraction of the measured copy bandwidth, tensor-pipe activity, registers, shared memory.
    python tools/ncu_table.py gpurun_out/prof_step.ncu-rep > profiles/r2/ncu_step_kernels.md"""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def col(hdr, key):
    for i, h in enumerate(hdr):
        if h == key or h.endswith("." + key):
            return i
    return None


def to_bytes(v, unit):
    v = float(v.replace(",", "")) if v else 0.0
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(unit.split("/")[0], 1)


def to_us(v, unit):
    v = float(v.replace(",", "")) if v else 0.0
    return v * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6, "usecond": 1, "msecond": 1e3, "nsecond": 1e-3, "second": 1e6}.get(unit, 1)


def main():
    rep = sys.argv[1]
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = peaks.get("hbm_gbs", 6650.0)
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    c = {k: col(hdr, k) for k in (
        "Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_tensor_subpipe_hmma.sum", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.avg.per_second",
        "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed.sum", "sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg",
        "sm__cycles_elapsed.avg")}
    print(f"# Per-kernel ncu digest of `{os.path.basename(rep)}`\n")
    print("`ncu --set full --clock-control none --import-source on --profile-from-start off` around "
          "`tools/ncu_workload.py` (one wide-MLP step with the serial schedule, one emulated 4-rank fused sync of a "
          f"64 MB bucket, the 2-3-1 step, one TF32 GEMM).  DRAM fraction is of the measured copy bandwidth "
          f"({hbm:.0f} GB/s, MEASURED_PEAKS.json).\n")
    print("`tensor busy` = `sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg` / (4 sub-pipes x "
          "`sm__cycles_elapsed.avg`): the share of the kernel's cycles in which the tensor pipe was issuing.\n")
    print("| # | kernel | us | DRAM rd MB | DRAM wr MB | DRAM GB/s | of HBM peak | tensor busy | L2 hit % | regs | grid x block | dyn smem KB | SM GHz |")
    print("|---|---|---|---|---|---|---|---|---|---|---|---|---|")
    for n, r in enumerate(rows[2:]):
        name = r[c["Kernel Name"]]
        name = re.sub(r"void |unnamed>::|b200dp::|\(anonymous namespace\)::", "", name)
        name = re.sub(r"\(.*$", "", name)[:52]
        us = to_us(r[c["gpu__time_duration.sum"]], units[c["gpu__time_duration.sum"]])
        rd = to_bytes(r[c["dram__bytes_read.sum"]], units[c["dram__bytes_read.sum"]])
        wr = to_bytes(r[c["dram__bytes_write.sum"]], units[c["dram__bytes_write.sum"]])
        gbs = (rd + wr) / (us * 1e-6) / 1e9 if us else 0
        hm = r[c["sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg"]] if c["sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg"] is not None else ""
        el = r[c["sm__cycles_elapsed.avg"]] if c["sm__cycles_elapsed.avg"] is not None else ""
        try:
            tp = 100.0 * float(hm.replace(",", "")) / (4.0 * float(el.replace(",", "")))
        except Exception:
            tp = 0.0
        l2 = r[c["lts__t_sector_hit_rate.pct"]] if c["lts__t_sector_hit_rate.pct"] is not None else ""
        smem = to_bytes(r[c["launch__shared_mem_per_block_dynamic"]], units[c["launch__shared_mem_per_block_dynamic"]]) / 1024
        ghz = r[c["sm__cycles_elapsed.avg.per_second"]]
        print(f"| {n} | `{name}` | {us:.1f} | {rd/1e6:.1f} | {wr/1e6:.1f} | {gbs:.0f} | {gbs/hbm:.2f} | "
              f"{tp:.1f} % | {float(l2 or 0):.0f} | {r[c['launch__registers_per_thread']]} | "
              f"{r[c['launch__grid_size']]} x {r[c['launch__block_size']]} | {smem:.1f} | {float(ghz or 0):.2f} |")


if __name__ == "__main__":
    main()
