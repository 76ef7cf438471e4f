#!/bin/bash
# ncu evidence (1 GPU): launch list of one bench run, then --set full of the two dominant kernels
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu"
$CMD > gpurun_out/r2_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_ -c 400 --csv --log-file gpurun_out/r2_launches.csv $CMD > gpurun_out/r2_ncu_list.log 2>&1
tail -3 gpurun_out/r2_ncu_list.log
ncu --set full --clock-control none --import-source on -k regex:k_correlate -s 2 -c 1 -o gpurun_out/r2_prof_correlate -f $CMD > gpurun_out/r2_ncu_corr.log 2>&1
tail -3 gpurun_out/r2_ncu_corr.log
ncu --set full --clock-control none --import-source on -k regex:k_apply -s 3 -c 1 -o gpurun_out/r2_prof_apply -f python bench.py --steps 2 --warmup 3 --no-cpu --mode track > gpurun_out/r2_ncu_apply.log 2>&1
tail -3 gpurun_out/r2_ncu_apply.log
ls -la gpurun_out/*.ncu-rep
