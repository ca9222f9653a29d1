#!/bin/bash
# End-of-round-2 ncu evidence (run under gpurun): launch list of the bench command with the packed DP kernel, and one
# --set full capture of k_dp16 on a 20 000-pair batch; text summaries go to gpurun_out/.
set -x
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --pairs-per-step 200000 > gpurun_out/r02ah_bench_small.json 2> gpurun_out/r02ah.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02ah_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --pairs-per-step 200000 > gpurun_out/r02ah_ncu_launch.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:"k_dp16" -c 2 -f -o /tmp/r02ah_full \
    python profiles/exp/exp_s7_variants.py 20000 0 > gpurun_out/r02ah_ncu_full.log 2>&1
python profiles/ncu_summary.py /tmp/r02ah_full.ncu-rep --lines 40 > gpurun_out/r02ah_summary.txt 2>&1
ncu -i /tmp/r02ah_full.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv, sys
rows = list(csv.reader(sys.stdin))
hdr = rows[0]
want = ['Kernel Name', 'gpu__time_duration.sum', 'launch__registers_per_thread', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum']
for w in want:
    if w in hdr:
        i = hdr.index(w)
        print(w, [r[i] for r in rows[1:]])
cols = [i for i, h in enumerate(hdr) if 'pipe' in h]
for i in cols:
    print(hdr[i], [r[i] for r in rows[1:]])
" > gpurun_out/r02ah_pipes.txt 2>&1
sz=$(stat -c %s /tmp/r02ah_full.ncu-rep); echo "report bytes $sz"
if [ "$sz" -lt 30000000 ]; then cp /tmp/r02ah_full.ncu-rep gpurun_out/; fi
ls -la gpurun_out | tail -8
