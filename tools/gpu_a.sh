#!/bin/bash
# single-GPU check of the round-2 tree: smoke, GPU tests, bench with the correlate variants (A/B)
mkdir -p gpurun_out
timeout 120 python __graft_entry__.py smoke > gpurun_out/r2a_smoke.txt 2>&1; tail -3 gpurun_out/r2a_smoke.txt
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2a_pytest.txt 2>&1; tail -30 gpurun_out/r2a_pytest.txt
for v in ""; do
  COHCUDA_LIB=$PWD/coherent-rtlsdr_b200/lib/libcoherent_cuda$v.so timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r2a_bench1$v.json 2> gpurun_out/r2a_bench1$v.err
  tail -c 300 gpurun_out/r2a_bench1$v.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2a_bench1$v.json"))
    print("variant[$v]", d["value"], d["kernel_ms_per_step"], "track", d["modes"]["track"]["value"], "e2e", d["e2e"]["value"])
except Exception as e:
    print("variant[$v] failed", e)
PY
done
