#!/bin/bash
mkdir -p gpurun_out
for v in "" $VARIANTS; do
  COHCUDA_LIB=$PWD/coherent-rtlsdr_b200/lib/libcoherent_cuda$v.so timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r2a3_bench$v.json 2> gpurun_out/r2a3_bench$v.err
  tail -c 300 gpurun_out/r2a3_bench$v.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2a3_bench$v.json"))
    print("variant[$v]", d["value"], d["kernel_ms_per_step"]["sync"], "track", d["modes"]["track"]["value"], "lag_ok", d["e2e"]["lag_known_answer_ok"])
except Exception as e:
    print("variant[$v] failed", e)
PY
done
