"""Worker for tests/test_sharding.py (CPU, gloo) and tests/test_multigpu.py (GPU, nccl): one
process per rank runs its share through coherent_rtlsdr_b200.sharding and rank 0 gathers,
assembles and saves the result.  With --engine oracle the per-rank engine is a stand-in built
on the CPU oracle (TEST ONLY: it lets the host-side partitioning, halo and gather logic run
without a GPU); with --engine cuda it is the product's CoherentCuda handle."""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


class OracleEngine:
    """Same ``batch`` signature as CoherentCuda, arithmetic by oracle/ (test infrastructure)."""

    def __init__(self, n_signal, blocksize, alpha=0.5, noheader=False, apply_shift=False, max_lag_slots=0, **_):
        from oracle import oracle as O
        self.O = O
        self.cfg = O.make_config(n_signal, blocksize, alpha=alpha, noheader=int(noheader),
                                 apply_shift=int(apply_shift), int_dot=1, max_lag_slots=max_lag_slots)
        self.state = O.new_state(n_signal + 1)
        self.noheader = noheader

    def batch(self, iq, lag_request=None, lag_mode=1, refnoise=None, readcnt0=0, globalseqn0=0, **_):
        r = self.O.process_batch(self.cfg, iq, state=self.state, lag_request=lag_request, lag_mode=lag_mode,
                                 refnoise=refnoise, readcnt0=readcnt0)
        if not self.noheader:
            gs = (globalseqn0 + np.arange(iq.shape[0])).astype(np.uint32)
            r["packets"][:, :4] = gs.view(np.int8).reshape(-1, 4)
        return r

    def close(self):
        pass


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", required=True)
    ap.add_argument("--engine", default="oracle")
    ap.add_argument("--case", required=True)        # npz with iq, lag_request, refnoise, meta
    ap.add_argument("--out", required=True)
    ap.add_argument("--transport", default="host")  # host | plan (CPU: cohcuda_gather_plan placement) | nccl | peer_copy | peer_store
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    kw = {}
    if a.engine == "cuda":
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        kw = {"device": local}
        factory = None
    else:
        dist.init_process_group("gloo")
        factory = OracleEngine
    import cohpkg
    cohpkg.load()
    from coherent_rtlsdr_b200 import sharding as S
    d = np.load(a.case)
    C, B, noheader, shift, slots = (int(v) for v in d["meta"])
    device_gather = a.transport in ("nccl", "peer_copy", "peer_store")
    sh = S.ShardedCoherent(C, B, rank, world, mode=a.mode, noheader=bool(noheader), apply_shift=bool(shift),
                           max_lag_slots=slots, engine_factory=factory, transport=a.transport if device_gather else None, **kw)
    if device_gather:
        # the product path: packets are gathered and assembled on the devices by libcoherent_cuda
        sh.setup_gather(d["iq"].shape[0], chunk_blocks=int(os.environ.get("COH_TEST_CHUNK", "0")))
        full = sh.run_gathered(d["iq"], lag_request=d["lag_request"], refnoise=d["refnoise"], readcnt0=8, globalseqn0=100)
        if rank == 0:
            np.savez(a.out, **full)
        sh.close()
        dist.barrier()
        dist.destroy_process_group()
        return
    local_out = sh.run(d["iq"], lag_request=d["lag_request"], refnoise=d["refnoise"], readcnt0=8, globalseqn0=100)
    parts = S.gather_packets(local_out, dst=0)
    if a.transport == "plan":
        # the library's placement arithmetic (cohcuda_gather_plan, no CUDA call) against the numpy assembly
        pk = cohpkg.load()
        import ctypes
        lib = pk.load_library()
        nb = d["iq"].shape[0]
        if a.mode == "channels":
            counts = [c1 - c0 for c0, c1 in S.channel_groups(C, world)]
            shard = pk.SHARD_CHANNELS
        else:
            counts = [b1 - b0 for b0, b1 in S.block_ranges(nb, world)]
            shard = pk.SHARD_BLOCKS
        cnt = (ctypes.c_uint32 * world)(*counts)
        cfg = pk.GatherConfig(rank, world, 0, shard, pk.GATHER_NCCL, cnt, nb, 0, 0)
        plans = []
        for r in range(world):
            g = pk.GatherGeometry()
            assert lib.cohcuda_gather_plan(ctypes.byref(cfg), r, B, noheader, C, ctypes.byref(g)) == 0
            plans.append(g)
        head = None
        if a.mode == "channels":
            rows = local_out["columns"].reshape(nb, -1)
            if rank == 0:
                hdr = 0 if noheader else 16 + 4 * (C + 1)
                head = np.zeros((nb, hdr + B), np.int8)
                if not noheader:
                    h32 = np.zeros((nb, 4 + C + 1), np.uint32)
                    h32[:, 0] = 100 + np.arange(nb); h32[:, 1] = C + 1; h32[:, 2] = B // 2
                    h32[:, 4:] = (8 + np.arange(nb, dtype=np.uint32))[:, None]
                    head[:, :hdr] = h32.view(np.int8).reshape(nb, hdr)
                head[:, hdr:] = local_out["ref"]
        else:
            rows = local_out["packets"]
        planned = S.gather_by_plan(rows, rank, plans, nb, root_head=head)
    if rank == 0:
        if a.mode == "channels":
            full = S.assemble_channel_packets(parts, C, B, bool(noheader), readcnt0=8, globalseqn0=100)
        else:
            full = S.assemble_block_packets(parts)
        if a.transport == "plan":
            assert (planned == full["packets"]).all(), "cohcuda_gather_plan placement differs from the numpy assembly"
            assert plans[0].packet_len == full["packets"].shape[1]
        np.savez(a.out, **full, pre_calls=np.array([p.get("pre_calls", 0) for p in parts]),
                 halo=np.array([p.get("halo", 0) for p in parts]))
    sh.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
