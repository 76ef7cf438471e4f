"""Multi-GPU partitioning of the block sync + correction path (one process per GPU).

The path shards with no exchange step inside the arithmetic: given the reference
channel's block, every (block, channel) item is independent; the only cross-block state is per
channel -- the phase IIR (src/csdrdevice.cc:58-69 of the reference) and the lag held since the
last measurement (include/csdrdevice.h:138-151).  Two partitionings:

* ``channels`` (BASELINE config 4, 168 channels): rank g owns a contiguous group of signal
  channels for ALL blocks; the reference channel is replicated to every rank and its spectrum
  recomputed locally.  Per-channel state never crosses ranks, so the result is identical to the
  single-engine run by construction.  The reference's cap of nfft-1 lag measurements per block
  (src/ccoherent.cc:124) is a cross-channel rule, so it is resolved here on the host
  (``effective_requests``) and the engines run uncapped.
* ``blocks`` (BASELINE config 5, sustained stream): rank r owns a contiguous range of blocks for
  all channels.  State at the start of the range is reconstructed locally, without waiting for
  the previous rank: the held lag by re-measuring the last block before the range in which each
  channel was measured (a handful of single-block calls), the phase IIR by warming up on a halo
  of ``lookback`` noise-enabled blocks (the truncated tail weighs (1-alpha)^lookback <= 2^-48,
  below fp32 resolution -- the same truncation the CUDA path's decoupled look-back applies
  inside one engine).

The only collective is the gather of the packed int8 packets to rank 0, which feeds the
reference's unchanged ZMQ publisher.  On GPUs it happens INSIDE libcoherent_cuda
(``cohcuda_gather_*``, csrc/api_gather.inc): every rank's k_apply writes its columns / rows so that
they travel chunk by chunk (NCCL send/recv, a copy-engine peer copy, or k_apply's own NVLink stores)
to their final place in the root's device buffer while the next chunk is computed; this module only
plans (who owns what) and carries the 136-byte attach blob and the per-block scalars
(``ShardedCoherent.setup_gather`` / ``run_gathered``).  ``gather_packets`` + ``assemble_*`` are the
host-side restatement of the same placement, used by the CPU tests (gloo) as the checker.

No arithmetic of the path happens in this file: engines are ``CoherentCuda`` handles (or, in the
CPU tests only, an injected stand-in with the same ``batch`` signature).
"""
from __future__ import annotations

import math

import numpy as np


# ------------------------------------------------------------------------------ planning
def effective_requests(lag_request, max_lag_slots):
    """Apply the reference's per-block slot cap (src/ccoherent.cc:124: slot 0 is the reference,
    at most nfft-1 signal channels are queued, in channel order) to a [n_blocks][C+1] request
    matrix.  Returns the uint8 matrix of requests that are actually measured."""
    req = np.ascontiguousarray(lag_request, dtype=np.uint8).copy()
    req[:, 0] = 0
    req = (req != 0).astype(np.uint8)
    if max_lag_slots and max_lag_slots < req.shape[1] - 1:
        order = np.cumsum(req, axis=1)
        req[order > max_lag_slots] = 0
    return req


def requests_for_mode(n_blocks, nch, lag_mode):
    req = np.zeros((n_blocks, nch), dtype=np.uint8)
    if lag_mode:
        req[:, 1:] = 1
    return req


def block_ranges(n_blocks, world):
    """Contiguous, balanced block ranges [b0, b1) per rank (empty ranges allowed)."""
    return [(r * n_blocks // world, (r + 1) * n_blocks // world) for r in range(world)]


def channel_groups(n_signal, world):
    """Contiguous groups of SIGNAL channels per rank as (c0, c1) with 1 <= c0 <= c1 <= n_signal+1
    (channel 0, the reference, is replicated everywhere)."""
    return [(1 + r * n_signal // world, 1 + (r + 1) * n_signal // world) for r in range(world)]


def lookback_depth(alpha):
    """Blocks after which the phase IIR has forgotten its initial state to below fp32
    resolution: (1-alpha)^depth <= 2^-48 (same rule as libcoherent_cuda, api.cu)."""
    if alpha >= 1.0:
        return 1
    return int(min(max(math.ceil(48.0 * math.log(2.0) / -math.log(1.0 - alpha)), 1), 2048))


def plan_block_shard(req_eff, refnoise, b0, lookback):
    """What rank with range start b0 must run before its own range.

    Returns (pre, h0): ``pre`` = ordered list of (block, flags[C+1]) single-block correlation
    calls that restore each channel's held lag as of block h0, and ``h0`` <= b0 the first block
    of the warm-up halo ([h0, b0) is processed with the original flags and discarded)."""
    nb, nch = req_eff.shape
    if b0 <= 0:
        return [], 0
    en = np.ones(nb, dtype=np.uint8) if refnoise is None else (np.asarray(refnoise) != 0).astype(np.uint8)
    # halo: the newest `lookback`+1 noise-enabled blocks before b0 (or everything back to 0)
    need, h0 = lookback + 1, b0
    while h0 > 0 and need > 0:
        h0 -= 1
        need -= int(en[h0])
    if need > 0:
        h0 = 0
    if h0 == 0:
        return [], 0
    pre = {}
    for c in range(1, nch):
        hit = np.flatnonzero(req_eff[:h0, c])
        if hit.size:
            pre.setdefault(int(hit[-1]), np.zeros(nch, dtype=np.uint8))[c] = 1
    return sorted(pre.items()), h0


# ------------------------------------------------------------------------------ engines
def _default_factory(n_signal, blocksize, **kw):
    from . import CoherentCuda          # the CUDA engine; raises if the library / GPU is missing
    return CoherentCuda(n_signal, blocksize, **kw)


class ShardedCoherent:
    """One rank's share of a batch.  ``mode`` = "blocks" or "channels"."""

    def __init__(self, n_signal, blocksize, rank, world, mode="blocks", alpha=0.5, noheader=False,
                 apply_shift=False, max_lag_slots=0, engine_factory=None, transport=None, **engine_kw):
        assert mode in ("blocks", "channels")
        self.transport = transport          # None: host-side gather (checker); else the library's device gather
        self.C, self.B, self.nch = n_signal, blocksize, n_signal + 1
        self.rank, self.world, self.mode = rank, world, mode
        self.alpha, self.noheader, self.max_lag_slots = alpha, bool(noheader), max_lag_slots
        self.lookback = lookback_depth(alpha)
        self.hdr_len = 0 if noheader else 16 + 4 * self.nch
        self.packet_len = self.hdr_len + self.nch * blocksize
        factory = engine_factory or _default_factory
        if mode == "channels":
            self.c0, self.c1 = channel_groups(n_signal, world)[rank]
            n_local = self.c1 - self.c0
            # host gather: columns only, rank 0 assembles header + readcnt + all columns afterwards;
            # device gather: the root's k_apply writes the gathered header itself
            self.engine = factory(n_local, blocksize, alpha=alpha, noheader=True if transport is None else noheader,
                                  apply_shift=apply_shift,
                                  max_lag_slots=0, **engine_kw) if n_local > 0 else None
        else:
            self.engine = factory(n_signal, blocksize, alpha=alpha, noheader=noheader, apply_shift=apply_shift,
                                  max_lag_slots=0, **engine_kw)

    def close(self):
        if self.engine is not None and hasattr(self.engine, "close"):
            self.engine.close()

    # -------------------------------------------------------------------------------------
    def run(self, iq, lag_request=None, lag_mode=1, refnoise=None, readcnt0=0, globalseqn0=0):
        """iq: the whole batch, uint8 [n_blocks][C+1][B] (every rank sees the same array or a
        memmap of it; only this rank's share is touched).  Returns this rank's outputs:
        blocks  -> dict(b0, b1, packets[b1-b0][packet_len], lag, mag, phase [b1-b0][C+1])
        channels-> dict(c0, c1, columns[n_blocks][c1-c0][B] (+ ref column on rank 0 as 'ref'),
                        lag, mag, phase [n_blocks][c1-c0])"""
        nb = iq.shape[0]
        assert iq.shape[1:] == (self.nch, self.B), iq.shape
        req = requests_for_mode(nb, self.nch, lag_mode) if lag_request is None else lag_request
        req = effective_requests(req, self.max_lag_slots)
        rn = None if refnoise is None else np.ascontiguousarray(refnoise, dtype=np.uint8)
        if self.mode == "channels":
            return self._run_channels(iq, req, rn, nb)
        return self._run_blocks(iq, req, rn, nb, readcnt0, globalseqn0)

    def _run_channels(self, iq, req, rn, nb):
        c0, c1 = self.c0, self.c1
        out = {"c0": c0, "c1": c1, "n_blocks": nb}
        if self.rank == 0:
            out["ref"] = (np.ascontiguousarray(iq[:, 0, :]) ^ np.uint8(0x80)).view(np.int8)   # cpacketizer.cc:151
        if c1 == c0:
            out.update(columns=np.zeros((nb, 0, self.B), np.int8), lag=np.zeros((nb, 0), np.float32),
                       mag=np.zeros((nb, 0), np.float32), phase=np.zeros((nb, 0), np.complex64))
            return out
        sel = [0] + list(range(c0, c1))
        r = self.engine.batch(np.ascontiguousarray(iq[:, sel, :]), lag_request=np.ascontiguousarray(req[:, sel]),
                              refnoise=rn)
        cols = r["packets"].reshape(nb, len(sel), self.B)[:, 1:, :]
        out.update(columns=np.ascontiguousarray(cols), lag=r["lag"][:, 1:], mag=r["mag"][:, 1:], phase=r["phase"][:, 1:])
        return out

    def _run_blocks(self, iq, req, rn, nb, readcnt0, globalseqn0):
        b0, b1 = block_ranges(nb, self.world)[self.rank]
        out = {"b0": b0, "b1": b1, "n_blocks": nb}
        if b1 == b0:
            out.update(packets=np.zeros((0, self.packet_len), np.int8), lag=np.zeros((0, self.nch), np.float32),
                       mag=np.zeros((0, self.nch), np.float32), phase=np.zeros((0, self.nch), np.complex64))
            return out
        pre, h0 = plan_block_shard(req, rn, b0, self.lookback)
        for blk, flags in pre:                       # restore the held lags as of block h0
            self.engine.batch(np.ascontiguousarray(iq[blk:blk + 1]), lag_request=flags[None, :],
                              refnoise=np.zeros(1, np.uint8), readcnt0=readcnt0 + blk, globalseqn0=globalseqn0 + blk)
        r = self.engine.batch(np.ascontiguousarray(iq[h0:b1]), lag_request=np.ascontiguousarray(req[h0:b1]),
                              refnoise=None if rn is None else rn[h0:b1],
                              readcnt0=readcnt0 + h0, globalseqn0=globalseqn0 + h0)
        k = b0 - h0
        out.update(packets=r["packets"][k:], lag=r["lag"][k:], mag=r["mag"][k:], phase=r["phase"][k:], halo=k,
                   pre_calls=len(pre))
        return out


    # ------------------------------------------------------------------------------------- device gather
    def setup_gather(self, n_blocks, group=None, chunk_blocks=0, store_ctas=0):
        """Collective: gather_init on every rank, export on rank 0, the blob through
        torch.distributed, attach everywhere.  n_blocks = blocks of the whole batch."""
        import torch.distributed as dist
        from . import SHARD_BLOCKS, SHARD_CHANNELS, TRANSPORTS
        assert self.transport in TRANSPORTS, self.transport
        if self.mode == "channels":
            counts = [c1 - c0 for c0, c1 in channel_groups(self.C, self.world)]
            assert min(counts) >= 1, "every rank needs at least one signal channel"
            shard = SHARD_CHANNELS
        else:
            counts = [b1 - b0 for b0, b1 in block_ranges(n_blocks, self.world)]
            shard = SHARD_BLOCKS
        self.gather_blocks = n_blocks
        self.geo = self.engine.gather_init(self.rank, self.world, counts, shard, TRANSPORTS[self.transport],
                                           max_blocks=max(n_blocks, 1), chunk_blocks=chunk_blocks, store_ctas=store_ctas)
        blob = [self.engine.gather_export() if self.rank == 0 else None]
        dist.broadcast_object_list(blob, src=0, group=group)
        self.engine.gather_attach(blob[0])
        dist.barrier(group=group)

    def run_gathered(self, iq, lag_request=None, lag_mode=1, refnoise=None, readcnt0=0, globalseqn0=0, group=None,
                     fetch=True):
        """Like run() + gather + assembly, but the packets never leave the devices: returns, on rank 0,
        dict(packets (if fetch), lag, mag, phase) of the whole batch and None elsewhere."""
        import torch.distributed as dist
        nb = iq.shape[0]
        assert nb == self.gather_blocks, "setup_gather() was sized for another batch"
        req = requests_for_mode(nb, self.nch, lag_mode) if lag_request is None else lag_request
        req = effective_requests(req, self.max_lag_slots)
        rn = None if refnoise is None else np.ascontiguousarray(refnoise, dtype=np.uint8)
        if self.mode == "channels":
            sel = [0] + list(range(self.c0, self.c1))
            r = self.engine.batch(np.ascontiguousarray(iq[:, sel, :]), lag_request=np.ascontiguousarray(req[:, sel]),
                                  refnoise=rn, readcnt0=readcnt0, globalseqn0=globalseqn0, gather=True)
            small = {"c0": self.c0, "c1": self.c1, "lag": r["lag"][:, 1:], "mag": r["mag"][:, 1:], "phase": r["phase"][:, 1:]}
        else:
            b0, b1 = block_ranges(nb, self.world)[self.rank]
            small = {"b0": b0, "b1": b1}
            if b1 > b0:
                pre, h0 = plan_block_shard(req, rn, b0, self.lookback)
                for blk, flags in pre:                       # restore the held lags as of block h0
                    self.engine.batch(np.ascontiguousarray(iq[blk:blk + 1]), lag_request=flags[None, :],
                                      refnoise=np.zeros(1, np.uint8), readcnt0=readcnt0 + blk, globalseqn0=globalseqn0 + blk)
                r = self.engine.batch(np.ascontiguousarray(iq[h0:b1]), lag_request=np.ascontiguousarray(req[h0:b1]),
                                      refnoise=None if rn is None else rn[h0:b1], readcnt0=readcnt0 + h0,
                                      globalseqn0=globalseqn0 + h0, gather=True, gather_skip_blocks=b0 - h0)
                k = b0 - h0
                small.update(lag=r["lag"][k:], mag=r["mag"][k:], phase=r["phase"][k:], halo=k, pre_calls=len(pre))
        # every rank's transfers are complete (batch() synchronised its streams); the barrier makes that
        # known to the root before it reads the gathered packets
        dist.barrier(group=group)
        parts = [None] * self.world if self.rank == 0 else None
        dist.gather_object(small, parts, dst=0, group=group)
        if self.rank != 0:
            return None
        if self.mode == "channels":
            lag = np.zeros((nb, self.nch), np.float32); mag = np.zeros((nb, self.nch), np.float32)
            phase = np.zeros((nb, self.nch), np.complex64); phase[:, 0] = 1.0
            for p in parts:
                lag[:, p["c0"]:p["c1"]], mag[:, p["c0"]:p["c1"]], phase[:, p["c0"]:p["c1"]] = p["lag"], p["mag"], p["phase"]
            out = {"lag": lag, "mag": mag, "phase": phase}
        else:
            out = {k: np.concatenate([p[k] for p in parts if "lag" in p], axis=0) for k in ("lag", "mag", "phase")}
            out["halo"] = np.array([p.get("halo", 0) for p in parts]); out["pre_calls"] = np.array([p.get("pre_calls", 0) for p in parts])
        if fetch:
            out["packets"] = self.engine.gather_fetch(0, nb)
        return out


# ------------------------------------------------------------------------------ collective
def assemble_channel_packets(parts, n_signal, blocksize, noheader, readcnt0=0, globalseqn0=0):
    """Rank-0 assembly for the ``channels`` partitioning: parts = per-rank dicts (rank order).
    Builds [hdr0 16][readcnt 4(C+1)][(C+1) columns x B] per block (include/cpacketizer.h:32-37,
    src/cpacketizer.cc:109-118,137-172 of the reference)."""
    nb, nch, B = parts[0]["n_blocks"], n_signal + 1, blocksize
    hdr = 0 if noheader else 16 + 4 * nch
    pk = np.zeros((nb, hdr + nch * B), dtype=np.int8)
    if not noheader:
        h32 = np.zeros((nb, 4 + nch), dtype=np.uint32)
        h32[:, 0] = globalseqn0 + np.arange(nb, dtype=np.uint32)
        h32[:, 1] = nch
        h32[:, 2] = B // 2
        h32[:, 4:] = (readcnt0 + np.arange(nb, dtype=np.uint32))[:, None]
        pk[:, :hdr] = h32.view(np.int8).reshape(nb, hdr)
    cols = pk[:, hdr:].reshape(nb, nch, B)
    cols[:, 0, :] = parts[0]["ref"]
    lag = np.zeros((nb, nch), np.float32); mag = np.zeros((nb, nch), np.float32)
    phase = np.zeros((nb, nch), np.complex64); phase[:, 0] = 1.0
    for p in parts:
        c0, c1 = p["c0"], p["c1"]
        cols[:, c0:c1, :] = p["columns"]
        lag[:, c0:c1], mag[:, c0:c1], phase[:, c0:c1] = p["lag"], p["mag"], p["phase"]
    return {"packets": pk, "lag": lag, "mag": mag, "phase": phase}


def assemble_block_packets(parts):
    """Rank-0 assembly for the ``blocks`` partitioning: concatenation in rank order."""
    return {k: np.concatenate([p[k] for p in parts], axis=0) for k in ("packets", "lag", "mag", "phase")}


def gather_packets(local, dst=0, group=None):
    """Gather every rank's outputs to rank ``dst`` over torch.distributed (NCCL on GPUs -- the
    int8 payload then travels over NVLink --, gloo on CPU).  Returns the per-rank list on ``dst``
    and None elsewhere.  Payload tensors may be CUDA or CPU; small metadata rides along as a
    python object."""
    import torch
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    backend = dist.get_backend(group)
    meta = {k: v for k, v in local.items() if not isinstance(v, np.ndarray)}
    shapes = {k: (v.shape, str(v.dtype)) for k, v in local.items() if isinstance(v, np.ndarray)}
    metas = [None] * world
    dist.all_gather_object(metas, (meta, shapes), group=group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    parts = [dict(m) for m, _ in metas] if rank == dst else None
    keys = sorted(set().union(*[s.keys() for _, s in metas]))        # a key may exist on some ranks only
    for k in keys:
        sizes = [int(np.prod(s[k][0])) * np.dtype(s[k][1]).itemsize if k in s else 0 for _, s in metas]
        mine = None
        if sizes[rank]:
            mine = torch.from_numpy(np.ascontiguousarray(local[k]).view(np.uint8).reshape(-1)).to(dev)
        if rank == dst:
            bufs = [torch.empty(n, dtype=torch.uint8, device=dev) if n else None for n in sizes]
            ops = [dist.P2POp(dist.irecv, bufs[r], r, group) for r in range(world) if r != dst and sizes[r]]
            bufs[dst] = mine
        else:
            ops = [dist.P2POp(dist.isend, mine, dst, group)] if mine is not None else []
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        if rank == dst:
            for r in range(world):
                if k not in metas[r][1]:
                    continue
                shp, dt = metas[r][1][k]
                raw = bufs[r].cpu().numpy() if bufs[r] is not None else np.zeros(0, np.uint8)
                parts[r][k] = raw.view(np.dtype(dt)).reshape(shp)
    return parts


def gather_by_plan(local_rows, plan_rank, plans, n_blocks, dst=0, group=None, root_head=None):
    """CPU checker of the library's placement arithmetic: ``plans`` = cohcuda_gather_plan() of every
    rank (GatherGeometry), ``local_rows`` = this rank's contribution as uint8/int8 [rows][run_bytes]
    (its columns of every block, or its whole packets).  Rows travel over torch.distributed and are
    placed at ``run_offset`` of gathered rows ``first_block + i`` -- exactly what api_gather.inc does
    on the device.  Returns the gathered [total_blocks][packet_len] int8 matrix on ``dst``."""
    import torch
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    me = plans[plan_rank]
    local_rows = np.ascontiguousarray(local_rows).view(np.int8)
    assert local_rows.shape[1] == me.run_bytes, (local_rows.shape, me.run_bytes)
    shapes = [None] * world
    dist.all_gather_object(shapes, local_rows.shape, group=group)
    if rank != dst:
        if local_rows.size:
            dist.send(torch.from_numpy(local_rows.reshape(-1).copy()), dst, group=group)
        return None
    G = np.zeros((n_blocks, plans[dst].packet_len), np.int8)
    if root_head is not None:                      # channel shards: the root also writes header, readcnt and column 0
        G[:, :root_head.shape[1]] = np.ascontiguousarray(root_head).view(np.int8)
    for r in range(world):
        rows = local_rows if r == dst else None
        if r != dst:
            buf = torch.empty(int(np.prod(shapes[r])), dtype=torch.int8)
            if buf.numel():
                dist.recv(buf, r, group=group)
            rows = buf.numpy().reshape(shapes[r])
        p = plans[r]
        G[p.first_block:p.first_block + rows.shape[0], p.run_offset:p.run_offset + p.run_bytes] = rows
    return G
