#!/bin/bash
# Round check on the GPU box: parity tests, the bench line, the ncu launch list of the same bench
# command and one full ncu capture of the sweep / DP kernels (summaries land in gpurun_out/).
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r01.json 2> gpurun_out/bench.err
tail -2 gpurun_out/bench.err; cat gpurun_out/bench_r01.json
timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --pairs-per-step 20000 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_r01.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline --pairs-per-step 20000 > gpurun_out/ncu1.log 2>&1
python profiles/exp/prof_target.py > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"k_sweep6|k_sweep5|k_dp" -c 6 -o /tmp/prof_g6 \
    python profiles/exp/prof_target.py > gpurun_out/ncu2.log 2>&1
python profiles/ncu_summary.py /tmp/prof_g6.ncu-rep --lines 30 > gpurun_out/prof_g6_summary.txt 2>&1
ncu -i /tmp/prof_g6.ncu-rep --page raw --csv > gpurun_out/prof_g6_raw.csv 2>/dev/null
[ $(stat -c %s /tmp/prof_g6.ncu-rep) -lt 40000000 ] && cp /tmp/prof_g6.ncu-rep gpurun_out/
