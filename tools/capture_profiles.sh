#!/bin/bash
# The command sequences behind profiles/r2_* (each line is one gpurun call from the repository root):
#   gpurun -- bash tools/gpu_a.sh            smoke, GPU tests, bench (optionally kernel variants via COHCUDA_LIB)
#   gpurun -- bash tools/gpu_c.sh            ncu: launch list + --set full of k_correlate_persist and k_apply
#   gpurun -- bash tools/gpu_misc.sh         sustained run, configs 1 / 3 / 5, both bench arms
#   gpurun --gpus 2 -- bash tools/gpu_b.sh 2 multi-GPU tests + bench with the product gather on 2 GPUs
#   gpurun --gpus 8 -- bash tools/gpu_8.sh   config 4 parity tests, host-link probe with 8 ranks, bench at N = 8
# then: python tools/ncu_summary.py gpurun_out/r2_prof_correlate.ncu-rep > profiles/r2_ncu_k_correlate.txt  (same for k_apply)
# tools/build_variant.sh NAME "-DFLAG" builds a library variant for A/B runs (tools/gpu_a3.sh / gpu_a4.sh).
echo "see the comments in this file"
