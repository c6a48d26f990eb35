#!/usr/bin/env python3
"""Turn an .ncu-rep (one `ncu --set full` capture) into the text summary kept under profiles/.
usage: python tools/ncu_summary.py gpurun_out/<name>.ncu-rep "<command that was profiled>" > profiles/<name>_ncu_full.txt"""
import csv, io, subprocess, sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum ', 'dram__bytes_write.sum ', 'launch__registers_per_thread ',
        'launch__grid_size', 'launch__block_size', 'launch__cluster_size', 'launch__shared_mem_per_block_dynamic',
        'launch__occupancy_limit', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct',
        'smsp__inst_executed.sum ', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed', 'sm__inst_executed_pipe_fma.avg.pct',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum ', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum ',
        'smsp__average_warps_issue_stalled', 'sm__icc_request_hit_rate', 'sm__cycles_elapsed.max ',
        'lts__t_sector_hit_rate.pct', 'sm__inst_executed.avg.per_cycle_elapsed']
rep, cmd = sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else ""
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
print(f"# ncu --set full --clock-control none --import-source on   ({rep.split('/')[-1]}, read with --page raw --csv)")
print(f"# command: {cmd}")
for vals in rows[2:]:
    for h, u, v in zip(hdr, units, vals):
        if h == 'Kernel Name' or any(k in h + ' ' for k in KEYS):
            print(f"{h} [{u}] = {v}")
