"""Host<->device copy bandwidth with ALL ranks copying at once (torchrun, one process per GPU): the box limit that
bench.py's e2e number at N GPUs has to be judged against.  Each rank moves 369 MB each way (one URA21 step) from /
to its own pinned buffers, both directions concurrently, started together behind a barrier; run once with the
process left where the launcher put it and once bound to the CPUs next to its GPU (NVML's affinity), with the
pinned buffers allocated after the binding.

    python -m torch.distributed.run --nproc-per-node 8 tools/pcie_probe_ranks.py
"""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def bind(index):
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = [i for i in range(ncpu) if (words[i // 64] >> (i % 64)) & 1]
        cpus = [c for c in cpus if c in os.sched_getaffinity(0)]
        if cpus:
            os.sched_setaffinity(0, cpus)
        return len(cpus)
    except Exception:
        return -1


def measure(dev, n, reps=6):
    h_in = torch.empty(n, dtype=torch.uint8, pin_memory=True); h_in.fill_(1)
    h_out = torch.empty(n, dtype=torch.uint8, pin_memory=True); h_out.fill_(2)
    d_in = torch.empty(n, dtype=torch.uint8, device=dev); d_out = torch.zeros(n, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def both():
        with torch.cuda.stream(s1):
            d_in.copy_(h_in, non_blocking=True)
        with torch.cuda.stream(s2):
            h_out.copy_(d_out, non_blocking=True)
    both(); torch.cuda.synchronize()
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        both()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    return n / dt / 1e9


def main():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29533")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = 369098752
    out = {"ranks": world, "bytes_each_way_per_rank": n}
    for label in ("unbound", "bound_to_gpu_numa_node"):
        ncpu = bind(local) if label != "unbound" else len(os.sched_getaffinity(0))
        g = measure(dev, n)
        all_g = [None] * world
        dist.all_gather_object(all_g, (g, ncpu))
        out[label] = {"gbs_per_direction_per_rank": [round(x[0], 1) for x in all_g],
                      "aggregate_gbs_per_direction": round(sum(x[0] for x in all_g), 1), "cpus_allowed": [x[1] for x in all_g]}
    if rank == 0:
        print(json.dumps(out))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
