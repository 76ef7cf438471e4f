#!/bin/bash
# multi-GPU check (2 or more GPUs): product gather tests, then the bench line with with_gather
N=${1:-2}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_multigpu.py -q > gpurun_out/r2_multigpu_pytest_${N}gpu.txt 2>&1; tail -25 gpurun_out/r2_multigpu_pytest_${N}gpu.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu > gpurun_out/r2_bench_${N}gpu.json 2> gpurun_out/r2_bench_${N}gpu.err
tail -c 1500 gpurun_out/r2_bench_${N}gpu.err
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2_bench_${N}gpu.json"))
    print("value", d["value"], "e2e", d["e2e"]["value"])
    print(json.dumps(d.get("with_gather"), indent=1)[:3000])
except Exception as e:
    print("bench failed", e)
PY
