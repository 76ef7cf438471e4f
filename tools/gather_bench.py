"""Only the with_gather part of bench.py (torchrun, one process per GPU): config 4's channel-group gather through the
library, per transport -- for tuning runs on 8 GPUs, where every second of box time is charged eight times.
    python -m torch.distributed.run --nproc-per-node 8 tools/gather_bench.py --gather-transports nccl --steps 5"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--blocks", type=int, default=1024)
    ap.add_argument("--gather-transports", default="nccl,peer_copy,peer_store")
    ap.add_argument("--gather-chunk", type=int, default=0)
    ap.add_argument("--store-ctas", type=int, default=0)
    ap.add_argument("--gather-verify-all", action="store_true")
    ap.add_argument("--alpha", type=float, default=0.5)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    bench._reserve_stdout()
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    import cohpkg
    pkg = cohpkg.load()
    stream = torch.cuda.Stream(device=dev)
    C, B, nb = bench.C_SIG, bench.BLOCKSIZE, args.blocks
    eng = pkg.CoherentCuda(C, B, apply_shift=True, device=local, max_batch_blocks=nb, stream=stream.cuda_stream, alpha=args.alpha)
    iq, _, _ = bench.gen_device_iq(torch, C, B, nb, 0xC0DE + 1 + 1000 * rank, dev, ref_seed=0xC0DE + 1)

    def barrier():
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        eng.flush()
        e1.record(stream)
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    out = bench.run_with_gather(torch, dist, pkg, eng, iq, args, rank, world, local, dev, stream, timed, barrier, nb, C + 1, B, B // 2)
    if rank == 0:
        out["env"] = {k: v for k, v in os.environ.items() if k.startswith("NCCL_") or k.startswith("COHCUDA_")}
        bench.emit(out)
    eng.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
