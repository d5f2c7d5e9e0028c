#!/usr/bin/env python
"""DRAM and L2 -> SM bytes per launch of the per-layer kernels, from an `ncu --set full` capture of the kernels INSIDE
a step (tools/profile_step.py: consecutive launches of one encoder layer in the second denoise step, so every kernel
finds L2 in the state its predecessor left it in -- unlike an isolated launch).  Read here, no GPU needed:

    python tools/ncu_traffic.py gpurun_out/r02k/prof_layer.ncu-rep --rows 25216 > profiles/r02k_ncu_layer.md
    (also rewrites profiles/traffic.json, which bench.py reports as roofline.traffic / traffic_l2_to_sm)

Capture (one GPU, after the same command exited 0 without ncu):
    ncu --set full --clock-control none --import-source on -k regex:'gemm_h16_tcgen05|attention_tc|gemm_ln_tcgen05|ffn_ln_tcgen05' \
        -s 47 -c 4 -o gpurun_out/r02k/prof_layer python tools/profile_step.py --workload cfg2a --denoise-steps 2
(34 matching launches per denoise step: 1 input projection + 8 x (in_proj, attention, out_proj+LN, feed-forward) + 1
output projection; skipping 34 + 1 + 3 x 4 lands on layer 3 of the second step.)
"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DRAM = ("dram__bytes_read.sum", "dram__bytes_write.sum")
# bytes that crossed the L2 -> SM crossbar: L1TEX-side read returns (generic loads) + everything the L2 served to this
# SM's TMA / bulk-copy units shows up in lts__t_bytes; both are recorded when the report has them
L2_SM = ("l1tex__m_xbar2l1tex_read_bytes.sum", "lts__t_bytes.sum", "lts__t_sectors_op_read.sum",
         "lts__t_sectors_srcunit_tex_op_read.sum", "l1tex__m_l1tex2xbar_write_bytes.sum")
OTHER = ("gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
         "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
         "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
         "lts__t_sector_hit_rate.pct", "launch__registers_per_thread", "sm__cycles_elapsed.max")
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "sector": 32.0, "usecond": 1.0, "us": 1.0, "msecond": 1e3,
        "ms": 1e3, "nsecond": 1e-3, "ns": 1e-3}


def algorithmic(name, M, d=512, F=1024):
    """Algorithmic HBM bytes per launch at the cfg2a shapes: activations in / out, weights L2-resident."""
    if "ffn_ln" in name:
        return M * d * 8.0, "ffn_ln_tcgen05_kernel M=%d d=%d F=%d" % (M, d, F)
    if "gemm_ln" in name:
        return M * (2.0 * d + 8.0 * d), "gemm_ln_tcgen05_kernel M=%d N=%d K=%d" % (M, d, d)
    if "attention_tc" in name:
        return 4.0 * 2 * M * d, "attention_tc_kernel M=%d S=197 H=4" % M
    if "gemm_h16" in name:
        return M * d * 2.0 + M * 3 * d * 2.0, "gemm_h16_tcgen05_kernel M=%d N=%d K=%d" % (M, 3 * d, d)
    return None, name


def main():
    path = sys.argv[1]
    M = int(sys.argv[sys.argv.index("--rows") + 1]) if "--rows" in sys.argv else 25216
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}

    def val(r, k):
        if k not in idx or r[idx[k]] in ("", "n/a"):
            return None
        v = float(r[idx[k]].replace(",", ""))
        return v * UNIT.get(units[idx[k]], 1.0)
    traffic = {"_about": "per launch, from an ncu --set full capture of consecutive launches of one encoder layer inside "
                         "the second denoise step of tools/profile_step.py (cfg2a); regenerate with tools/ncu_traffic.py",
               "_source": os.path.relpath(path, ROOT)}
    print("| kernel | duration us | dram read MB | dram write MB | dram total / algorithmic | L2->SM MB (l1tex__m_xbar2l1tex_read_bytes) | "
          "L2->SM / algorithmic | L2 hit % | tensor pipe % (active) |")
    print("|---|---|---|---|---|---|---|---|---|")
    for r in rows[2:]:
        name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "").replace("gd::", "")
        alg, label = algorithmic(name, M)
        rd, wr = val(r, DRAM[0]) or 0.0, val(r, DRAM[1]) or 0.0
        xbar = val(r, "l1tex__m_xbar2l1tex_read_bytes.sum")       # bytes the crossbar delivered into the SMs (TMA loads,
        l2 = xbar                                                   # generic loads, DSMEM traffic): the L2 -> SM figure
        ent = {"dram_bytes_per_launch": rd + wr, "read": rd, "write": wr, "algorithmic": alg,
               "l2_to_sm_bytes_per_launch": l2,
               "tma_load_bytes": val(r, "l1tex__m_xbar2l1tex_read_bytes_mem_global_op_tma_ld.sum"),
               "dsmem_bytes": val(r, "l1tex__m_xbar2l1tex_read_bytes_mem_dshared.sum"),
               "sm_to_l2_write_bytes": val(r, "l1tex__m_l1tex2xbar_write_bytes.sum"),
               "duration_us": val(r, "gpu__time_duration.sum"),
               "tensor_pipe_pct_active": val(r, OTHER[1]), "l2_hit_pct": val(r, "lts__t_sector_hit_rate.pct")}
        traffic[label] = ent
        print("| `%s` | %.1f | %.1f | %.1f | %s | %s | %s | %s | %s |"
              % (name, ent["duration_us"] or 0, rd / 1e6, wr / 1e6, ("%.2f" % ((rd + wr) / alg)) if alg else "",
                 ("%.0f" % (l2 / 1e6)) if l2 else "", ("%.1f" % (l2 / alg)) if (l2 and alg) else "",
                 ("%.0f" % ent["l2_hit_pct"]) if ent["l2_hit_pct"] is not None else "",
                 ("%.1f" % ent["tensor_pipe_pct_active"]) if ent["tensor_pipe_pct_active"] is not None else ""))
    if "--no-write" not in sys.argv:
        with open(os.path.join(ROOT, "profiles", "traffic.json"), "w") as fh:
            json.dump(traffic, fh, indent=1)


if __name__ == "__main__":
    main()
