#!/usr/bin/env python
"""Reduce `ncu -i X.ncu-rep --page raw --csv` to the table profiles/rNN_ncu_full_summary.md holds.
  python tools/ncu_full_table.py gpurun_out/r02g/ncu_full_raw.csv "title line" > profiles/r02_ncu_full_summary.md"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
cols = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "time us"), ("dram__bytes_read.sum", "dram read MB"),
        ("dram__bytes_write.sum", "dram write MB"),
        ("sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe cycles active %"),
        ("sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active", "tensor inst %"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"), ("lts__t_sector_hit_rate.pct", "L2 hit %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"), ("sm__inst_executed.avg.per_cycle_active", "IPC"),
        ("launch__registers_per_thread", "regs"), ("launch__shared_mem_per_block_dynamic", "dyn smem KB")]
ix = {h: i for i, h in enumerate(hdr)}


def conv(name, v, u):
    v = float(v.replace(",", ""))
    if name == "gpu__time_duration.sum":
        return "%.1f" % (v * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(u, 1))
    if "dram__bytes" in name:
        return "%.1f" % (v * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1, "Gbyte": 1e3}.get(u, 1))
    if "shared_mem" in name:
        return "%.1f" % (v * {"byte": 1e-3, "Kbyte": 1}.get(u, 1))
    return "%.4g" % v


print("# " + (sys.argv[2] if len(sys.argv) > 2 else "ncu --set full") + "\n")
print("| " + " | ".join(c[1] for c in cols) + " |")
print("|---|" + "---:|" * (len(cols) - 1))
for r in rows[2:]:
    out = []
    for name, lab in cols:
        v = r[ix[name]]
        out.append("`%s`" % v.split("(")[0] if name == "Kernel Name" else conv(name, v, units[ix[name]]))
    print("| " + " | ".join(out) + " |")
