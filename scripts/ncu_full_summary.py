#!/usr/bin/env python3
"""Condense one `ncu --set full` report (.ncu-rep) into the handful of numbers DESIGN.md / bench.py quote.
Usage: ncu_full_summary.py REPORT.ncu-rep [algorithmic_flop] [algorithmic_bytes]   (reads it with `ncu -i ... --page raw --csv`)"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
flop = float(sys.argv[2]) if len(sys.argv) > 2 else 0.0
abytes = float(sys.argv[3]) if len(sys.argv) > 3 else 0.0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
WANT = [
    ("Kernel Name", "kernel"), ("launch__grid_size", "grid"), ("launch__block_size", "block"), ("launch__cluster_size", "cluster"),
    ("launch__registers_per_thread", "registers/thread"), ("launch__shared_mem_per_block_dynamic", "dynamic smem/block"),
    ("gpu__time_duration.sum", "duration"), ("sm__cycles_elapsed.max", "SM cycles elapsed (max)"),
    ("sm__cycles_active.avg", "SM cycles active (avg over all SMs)"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput % of peak"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe active % (of active cycles)"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor pipe active % (of elapsed)"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "XU (SFU) pipe % of peak (active)"),
    ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "FMA pipe active %"),
    ("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "ALU pipe active %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("smsp__sass_inst_executed_op_utcmma.sum", "tcgen05.mma instructions (UTCxMMA)"),
    ("smsp__sass_inst_executed_op_tmem_ldt.sum", "tcgen05.ld instructions"),
    ("smsp__sass_inst_executed_op_tmem_stt.sum", "tcgen05.st instructions"),
    ("dram__bytes_read.sum", "DRAM bytes read"), ("dram__bytes_write.sum", "DRAM bytes written"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput % of peak"),
    ("lts__t_bytes.sum", "L2 bytes"), ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "shared-memory wavefronts (LSU)"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "shared-memory bank conflicts"),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long_scoreboard / issue"),
    ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stall barrier / issue"),
    ("smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "stall mio_throttle / issue"),
    ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "stall math_pipe_throttle / issue"),
    ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall short_scoreboard / issue"),
    ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall wait / issue"),
]
for d in rows[2:]:
    vals = {h: (d[i], units[i]) for i, h in enumerate(hdr)}
    print(f"# {rep}  (ncu --set full --clock-control none: cold caches, serialised launch)")
    dur_us = None
    for key, label in WANT:
        if key in vals:
            v, u = vals[key]
            print(f"{label:48s} {v} {u}")
            if key == "gpu__time_duration.sum":
                f = float(v.replace(",", ""))
                dur_us = f * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u, 1.0)
    if dur_us and flop:
        print(f"{'algorithmic TFLOP/s at this duration':48s} {flop / dur_us / 1e6:.1f}")
    if dur_us and abytes:
        print(f"{'algorithmic GB/s at this duration':48s} {abytes / dur_us / 1e3:.1f}")
    print()
