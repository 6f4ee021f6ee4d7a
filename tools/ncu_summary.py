#!/usr/bin/env python
"""Condense an ``ncu --set full`` report (read here, no GPU needed) into the few lines the roofline
claims rest on.  Usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep [kernel-regex] > profiles/x.txt"""
import csv
import io
import re
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_tensor.sum",
    "sm__inst_executed_pipe_tmem.avg.pct_of_peak_sustained_active",
    "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_uniform.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "launch__grid_size",
    "launch__block_size",
    "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic",
    "launch__occupancy_limit_shared_mem",
    "smsp__cycles_active.avg",
    "sm__cycles_elapsed.max",
]


def main():
    rep = sys.argv[1]
    pat = re.compile(sys.argv[2]) if len(sys.argv) > 2 else None
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, body = rows[0], rows[1], rows[2:]
    ki = hdr.index("Kernel Name")
    body = [r for r in body if pat is None or pat.search(r[ki])]
    print("report: %s   launches captured: %d" % (rep, len(body)))
    print("kernel: %s" % sorted({r[ki][:120] for r in body}))
    cols = {}
    for i, name in enumerate(hdr):
        short = name.split(".", 2)[-1] if name.startswith(("SM_", "TPC.", "GPC.", "FBPA.", "LTS.")) else name
        cols.setdefault(short, i)
        cols.setdefault(name, i)
    for w in WANT:
        if w in cols:
            i = cols[w]
            print("%-75s %-14s %s" % (w, units[i], [r[i] for r in body]))
    extra = [n for n in hdr if ("tensor" in n or "utc" in n.lower()) and "pct" in n and n not in WANT]
    for n in extra:
        i = hdr.index(n)
        vals = [r[i] for r in body]
        if any(v not in ("0", "", "n/a") for v in vals):
            print("%-75s %-14s %s" % (n[-75:], units[i], vals))


if __name__ == "__main__":
    main()
