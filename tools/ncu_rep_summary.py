#!/usr/bin/env python
"""Key `ncu --set full` metrics of every kernel in one or more .ncu-rep files (read with `ncu -i ... --page raw --csv`).

    python tools/ncu_rep_summary.py gpurun_out/prof_r1b/*.ncu-rep
"""
import csv
import io
import subprocess
import sys

KEYS = [("gpu__time_duration.sum", "time"), ("sm__pipe_tensor_subpipe_hmma_cycles_active", "tensor%"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
        ("sm__inst_executed_pipe_uniform", None),
        ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("launch__registers_per_thread", "regs"), ("lts__t_sector_hit_rate.pct", "l2hit%"),
        ("sm__cycles_active.avg", "sm_cycles")]
for path in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rd = list(csv.reader(io.StringIO(out)))
    hdr, units = rd[0], rd[1]
    for row in rd[2:]:
        d = dict(zip(hdr, row))
        u = dict(zip(hdr, units))
        name = d.get("Kernel Name", "?").split("(")[0][-40:]
        parts = [f"{path.split('/')[-1]:28s} {name:28s} grid {d.get('Grid Size', '')} block {d.get('Block Size', '')}"]
        for k, lab in KEYS:
            for h in hdr:
                if h.startswith(k) and lab:
                    parts.append(f"{lab}={d[h]}{u[h] if u[h] not in ('%', '') else ''}")
                    break
        print("  ".join(parts))
