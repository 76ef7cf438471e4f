#!/bin/bash
mkdir -p gpurun_out
for v in "" $VARIANTS; do
  COHCUDA_LIB=$PWD/coherent-rtlsdr_b200/lib/libcoherent_cuda$v.so timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu --mode track > gpurun_out/r2a4_bench$v.json 2> gpurun_out/r2a4_bench$v.err
  tail -c 300 gpurun_out/r2a4_bench$v.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2a4_bench$v.json"))
    print("variant[$v]", "track %.4g" % d["value"], "k_apply ms", round(d["kernel_ms_per_step"]["track"]["k_apply"],4), "frac", round(d["roofline"]["frac"],4), "sync %.4g" % d["modes"]["sync"]["value"])
except Exception as e:
    print("variant[$v] failed", e)
PY
done
