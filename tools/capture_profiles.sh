#!/bin/bash
# Evidence capture on the GPU box, one sub-command per gpurun call (run from the repository root):
#   gpurun -- bash tools/capture_profiles.sh tests [TAG]         smoke, GPU tests, both bench arms
#   gpurun -- bash tools/capture_profiles.sh ncu [TAG]           launch list + ncu --set full of k_correlate / k_apply
#   gpurun -- bash tools/capture_profiles.sh big [TAG]           launch list + ncu --set full of the large-block kernels
#   gpurun --gpus N -- bash tools/capture_profiles.sh multi N [TAG]   multi-GPU tests + bench with the product gather
# Outputs land in gpurun_out/TAG_* (default TAG r3); summarise a report with
#   python tools/ncu_summary.py gpurun_out/TAG_prof_correlate.ncu-rep > profiles/TAG_ncu_k_correlate.txt
# tools/build_variant.sh NAME "-DFLAG" builds a library variant for A/B runs (COHCUDA_LIB=... selects it).
# A number printed by a run under ncu is never a bench value: every ncu pass follows a plain run that exited 0.
set -u
what=${1:-tests}
mkdir -p gpurun_out
case "$what" in
tests)
    tag=${2:-r3}
    python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; tail -1 gpurun_out/${tag}_smoke.log
    python -m pytest tests -x -q -m gpu > gpurun_out/${tag}_pytest_gpu.log 2>&1; tail -3 gpurun_out/${tag}_pytest_gpu.log
    python bench.py --impl reference > gpurun_out/${tag}_bench_reference.json 2> gpurun_out/${tag}_bench_reference.err
    python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -c 400 gpurun_out/${tag}_bench.json
    ;;
ncu)
    tag=${2:-r3}
    CMD="python bench.py --steps 2 --warmup 3 --no-cpu --sustained 0"
    $CMD > gpurun_out/${tag}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${tag}_plain.log; exit 1; }
    ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_ -c 400 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu_list.log 2>&1
    ncu --set full --clock-control none --import-source on -k regex:k_correlate -s 2 -c 1 -o gpurun_out/${tag}_prof_correlate -f $CMD > gpurun_out/${tag}_ncu_corr.log 2>&1
    ncu --set full --clock-control none --import-source on -k regex:k_apply -s 3 -c 1 -o gpurun_out/${tag}_prof_apply -f $CMD --mode track > gpurun_out/${tag}_ncu_apply.log 2>&1
    ls -la gpurun_out/*.ncu-rep
    ;;
big)
    tag=${2:-r3}
    CMD="python bench.py --config big65536 --steps 2 --warmup 3 --no-cpu --sustained 0"
    $CMD > gpurun_out/${tag}_big_plain.json 2> gpurun_out/${tag}_big_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${tag}_big_plain.err; exit 1; }
    ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_ -c 400 --csv --log-file gpurun_out/${tag}_big_launches.csv $CMD > gpurun_out/${tag}_big_ncu_list.log 2>&1
    for k in k_big_split k_big_sub k_big_merge; do
        ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -o gpurun_out/${tag}_prof_$k -f $CMD > gpurun_out/${tag}_ncu_$k.log 2>&1
    done
    ls -la gpurun_out/*.ncu-rep
    ;;
multi)
    n=${2:-2}; tag=${3:-r3}
    python -m pytest tests/test_multigpu.py -x -q -m gpu > gpurun_out/${tag}_multigpu_pytest_${n}gpu.log 2>&1; tail -3 gpurun_out/${tag}_multigpu_pytest_${n}gpu.log
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 \
        bench.py --gpus $n > gpurun_out/${tag}_bench_${n}gpu.json 2> gpurun_out/${tag}_bench_${n}gpu.err
    tail -c 600 gpurun_out/${tag}_bench_${n}gpu.json
    ;;
*)
    echo "usage: $0 tests|ncu|big|multi [N] [TAG]"; exit 2 ;;
esac
