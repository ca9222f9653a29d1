#!/bin/bash
# End-of-round-2 ncu evidence for the sweep (run under gpurun): --set full of the k_sweep7 launches of one 20 000-pair
# batch, pipe utilisation and issue / occupancy figures extracted as text.
set -x
python profiles/exp/exp_s7_variants.py 20000 0 > gpurun_out/r02ax_plain.log 2>&1 || exit 1
ncu --set full --clock-control none -k regex:"k_sweep7" -c 14 -f -o /tmp/r02ax_full \
    python profiles/exp/exp_s7_variants.py 20000 0 > gpurun_out/r02ax_ncu_full.log 2>&1
ncu -i /tmp/r02ax_full.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv, sys
rows = list(csv.reader(sys.stdin))
hdr = rows[0]
want = ['Kernel Name', 'gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__block_size', 'launch__grid_size',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed']
for w in want:
    if w in hdr:
        i = hdr.index(w)
        print(w, [r[i][:34] for r in rows[1:]])
for i, h in enumerate(hdr):
    if 'warp_issue_stalled' in h and 'per_warp_active' in h:
        print(h, [r[i] for r in rows[1:]])
" > gpurun_out/r02ax_pipes.txt 2>&1
ls -la /tmp/r02ax_full.ncu-rep gpurun_out | tail -6
