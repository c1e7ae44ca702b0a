#!/usr/bin/env python
"""Key metrics of every kernel in an .ncu-rep (reads `ncu -i FILE --page raw --csv` on stdin or runs it)."""
import csv
import subprocess
import sys

WANT = [
    ("time_us", "gpu__time_duration.sum"),
    ("dram_rd_MB", "dram__bytes_read.sum"),
    ("dram_wr_MB", "dram__bytes_write.sum"),
    ("dram_pct", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    ("tensor_pct", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
    ("warps_active_pct", "sm__warps_active.avg.pct_of_peak_sustained_active"),
    ("issue_active_pct", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
    ("regs", "launch__registers_per_thread"),
    ("grid", "launch__grid_size"),
    ("block", "launch__block_size"),
    ("smem_dyn_KB", "launch__shared_mem_per_block_dynamic"),
    ("inst_M", "smsp__inst_executed.sum"),
]


def main():
    path = sys.argv[1]
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    print("| kernel | " + " | ".join(k for k, _ in WANT) + " |")
    print("|---|" + "---|" * len(WANT))
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")]
        name = name.split("(")[0].replace("void kdl::", "").replace("kdl::", "")
        vals = []
        for k, m in WANT:
            if m not in hdr:
                vals.append("-")
                continue
            i = hdr.index(m)
            v, u = r[i], units[i]
            try:
                f = float(v.replace(",", ""))
                if u in ("byte",): f /= 1e6
                if u in ("Kbyte",): f /= 1e3
                if u in ("Gbyte",): f *= 1e3
                if u == "ns": f /= 1e3
                if u == "ms": f *= 1e3
                if k == "inst_M": f /= 1e6
                if k == "smem_dyn_KB" and u == "byte": f = f * 1e6 / 1024
                if k == "smem_dyn_KB" and u == "Kbyte": f = f * 1e3 / 1.024 / 1e3
                vals.append(f"{f:.1f}" if f < 1e5 else f"{f:.0f}")
            except ValueError:
                vals.append(v)
        print(f"| {name[:60]} | " + " | ".join(vals) + " |")


if __name__ == "__main__":
    main()
