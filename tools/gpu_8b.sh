#!/bin/bash
# NCCL transport tuning at N GPUs (gather only)
N=${1:-8}
mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 tools/gather_bench.py --gather-transports nccl --steps 5 $EXTRA > gpurun_out/r2_gather_$tag.json 2> gpurun_out/r2_gather_$tag.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2_gather_$tag.json"))
    v=d["transports"]["nccl"]
    print("$tag", {m:(round(v[m]["value"]/1e9,1), round(v[m]["root_ingest_gbs"],1)) for m in ("sync","track") if m in v}, v.get("bytes_identical_to_single_gpu"), v.get("error"))
except Exception as e:
    print("$tag failed", e); print(open("gpurun_out/r2_gather_$tag.err").read()[-800:])
PY
}
EXTRA="" run res16 COHCUDA_GATHER_RESERVE=16
EXTRA="" run res32 COHCUDA_GATHER_RESERVE=32
EXTRA="" run res32_ch8 COHCUDA_GATHER_RESERVE=32 NCCL_MIN_P2P_NCHANNELS=8
EXTRA="--gather-chunk 256" run res32_chunk256 COHCUDA_GATHER_RESERVE=32
