#!/bin/bash
# One-GPU measurement pass behind profiles/rNN_*: run under gpurun, e.g.
#   gpurun --timeout 1500 -- 'bash tools/capture_profiles.sh r1d'
# Every number that is reported comes from a run WITHOUT the profiler; the ncu passes come after.
tag=${1:-rXX}; out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | grep -v "^asynch\|exited$" | tail -3 > $out/${tag}_pytest.txt; cat $out/${tag}_pytest.txt
timeout 600 python bench.py --steps 20 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo bench=$?
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_reference_arm.json 2> $out/${tag}_ref.err; echo ref=$?
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu > $out/ncu_list.log 2>&1; echo ncu1=$?
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_apply -s 2 -c 1 -o $out/k_apply_${tag} \
    python bench.py --steps 2 --warmup 1 --no-cpu --mode track > $out/ncu_apply.log 2>&1; echo ncu2=$?
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_correlate -s 1 -c 1 -o $out/k_correlate_${tag} \
    python bench.py --steps 2 --warmup 1 --no-cpu --mode sync > $out/ncu_corr.log 2>&1; echo ncu3=$?
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks_throttle_reasons.active --format=csv > $out/${tag}_smi.csv
