#!/bin/bash
# the 8-GPU evidence of the round in one call: config 4 parity tests, host-link probe with 8 concurrent ranks,
# bench at N = 8 (with_gather through the product path), multi-GPU log kept under gpurun_out/
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2_topo_8gpu.txt 2>&1
timeout 900 python -m pytest tests/test_multigpu.py -q -k "config4 or delay_step or (blocks and nccl)" > gpurun_out/r2_multigpu_pytest_8gpu.txt 2>&1; tail -15 gpurun_out/r2_multigpu_pytest_8gpu.txt
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 tools/pcie_probe_ranks.py > gpurun_out/r2_pcie_probe_8ranks.json 2> gpurun_out/r2_pcie_probe_8ranks.err; cat gpurun_out/r2_pcie_probe_8ranks.json
timeout 120 python tools/pcie_probe_ranks.py > gpurun_out/r2_pcie_probe_1rank.json 2>/dev/null; cat gpurun_out/r2_pcie_probe_1rank.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu > gpurun_out/r2_bench_8gpu.json 2> gpurun_out/r2_bench_8gpu.err
tail -c 1200 gpurun_out/r2_bench_8gpu.err
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2_bench_8gpu.json"))
    print("value", d["value"], "e2e", d["e2e"]["value"])
    for k,v in d["with_gather"]["transports"].items():
        print(k, {m: (round(v[m]["value"]/1e9,1), round(v[m]["root_ingest_gbs"],1)) for m in ("sync","track") if m in v}, v.get("bytes_identical_to_single_gpu"), v.get("error"))
except Exception as e:
    print("bench failed", e)
PY
