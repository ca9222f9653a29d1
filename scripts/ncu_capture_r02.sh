#!/bin/bash
# Round-2 ncu evidence (run under gpurun): launch list of the bench command, then one --set full capture of the sweep
# and DP kernels on a 20 000-pair batch; text summaries go to gpurun_out/, the report itself only when it is small.
set -x
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --pairs-per-step 200000 > gpurun_out/r02w_bench_small.json 2> gpurun_out/r02w.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02w_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --pairs-per-step 200000 > gpurun_out/r02w_ncu_launch.log 2>&1
python profiles/exp/exp_s7_variants.py 20000 0 > gpurun_out/r02w_plain.log 2>&1 || exit 1
ncu --set full --import-source on --clock-control none -k regex:"k_sweep7|k_dp" -c 18 -f -o /tmp/r02w_full \
    python profiles/exp/exp_s7_variants.py 20000 0 > gpurun_out/r02w_ncu_full.log 2>&1
ncu -i /tmp/r02w_full.ncu-rep --page raw --csv > gpurun_out/r02w_raw.csv 2>/dev/null
python profiles/ncu_summary.py /tmp/r02w_full.ncu-rep --lines 25 > gpurun_out/r02w_summary.txt 2>&1
for k in "k_sweep7<0,12,9,256,2>" "k_sweep7<2,12,9,256,2>" "k_sweep7<0,12,9,320,2>" "k_sweep7<2,12,9,192,2>" "k_dp<4,false>"; do
  python profiles/ncu_lines.py /tmp/r02w_full.ncu-rep "$k" 30 phases >> gpurun_out/r02w_lines.txt 2>&1
done
sz=$(stat -c %s /tmp/r02w_full.ncu-rep); echo "report bytes $sz"
if [ "$sz" -lt 45000000 ]; then cp /tmp/r02w_full.ncu-rep gpurun_out/; fi
ls -la gpurun_out | tail -12
