#!/usr/bin/env python
"""Per-kernel summary of an `ncu --set full` report (read here, no GPU needed).

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/rNN_ncu_<what>.md
"""
import csv
import io
import subprocess
import sys

WANT = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe % (active)"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor pipe % (elapsed)"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex %"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2 %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
    ("launch__registers_per_thread", "regs/thread"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("sm__cycles_elapsed.max", "cycles"),
]


def main(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    cols = [(k, n) for k, n in WANT if k in idx]
    print("| # | kernel | " + " | ".join(n for _, n in cols) + " |")
    print("|---|---|" + "---|" * len(cols))
    for j, r in enumerate(rows[2:]):
        name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "")
        cells = []
        for k, _ in cols:
            v, u = r[idx[k]], units[idx[k]]
            try:
                v = "%.4g" % float(v.replace(",", ""))
            except ValueError:
                pass
            cells.append((v + " " + u).strip() if u not in ("%", "") else v)
        print("| %d | `%s` | " % (j, name) + " | ".join(cells) + " |")


if __name__ == "__main__":
    main(sys.argv[1])
