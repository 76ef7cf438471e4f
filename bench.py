#!/usr/bin/env python
"""bench.py -- coherent samples/s of the block sync + correction path on synthetic IQ.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--mode sync|track] [--impl reference]

One "step" = one pass of the hot path over one batch: URA21 geometry (21 signal + 1 reference
channel, 8192-sample blocks, 1024 blocks = BASELINE.json configs[1]).  `value` = coherent samples/s
((C+1)*L*blocks / time, every column of the emitted matrix counts) with the input already in HBM;
`e2e` = the same through cohcuda_batch with pinned HOST buffers (H2D + kernels + D2H timed).
N > 1 (torchrun): one process per GPU, each rank runs its own 21 signal channels against its own copy
of the ONE shared reference channel (weak scaling, no data-path collective in `value`).  `with_gather`
is BASELINE.json configs[3] when N = 8 (168 = 8 x 21 channels, channel-group sharded): the same step
through the library's device gather (cohcuda_gather_*, three transports), the gathered
[hdr][readcnt][1 + 21 N columns] matrix verified byte for byte against a single-GPU engine that
processes all channels, with the bytes the root received per second.
`--impl reference` times the reference's CPU path (oracle/_ref when built, else the oracle port)
on the host cores for the same metric.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

C_SIG, BLOCKSIZE, N_BLOCKS = 21, 16384, 1024
CONFIG = "ura21"
# BASELINE.json configs other than the bench line (configs[1] = ura21), for BASELINE.md's table:
#   --config four      configs[0]: four.cfg geometry, 3+1 channels x 8192 samples, 64 blocks
#   --config big65536  configs[2]: 21+1 channels x 65536-sample blocks, resident chunk of 256 blocks
#   --config stream    configs[4]: URA21 ring of 1024 blocks replayed; 1 % of the (block, channel) items
#                      re-request a lag (drift re-sync), everything else tracks: use --mode mixed
# configs[3] (168 channels, channel groups of 21 per GPU) is the ura21 line at --gpus 8 with_gather.
PRESETS = {"ura21": (21, 16384, 1024), "four": (3, 16384, 64), "big65536": (21, 131072, 256),
           "stream": (21, 16384, 1024)}
CONFIG_INDEX = {"four": 0, "ura21": 1, "big65536": 2, "stream": 4}   # index into BASELINE.json "configs" (3 = ura21 at --gpus 8 with_gather)
REQ = None                        # host [n_blocks][C+1] lag-request flags for --mode mixed
STEPS = None                      # [(block, channel, new delay)] delay steps of --mode mixed
BYTES_PER_SAMPLE = 4.0            # 2 B u8 IQ in + 2 B int8 IQ out (packet mode)
FLOP_PER_SAMPLE_SYNC = 311.0      # SURVEY 8d algorithmic flops per coherent sample, SYNC mode


def ncu_traffic(kernel):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture (profiles/)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)[kernel]
        return t["bytes_per_launch"], t["source"]
    except Exception:
        return None, None


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.draw")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
            # NVML start-up takes driver locks for a while on an 8-GPU box: do not let it
            # overlap the timed region -- wait for the first sample before returning
            t0 = time.time()
            while not self.rows and time.time() - t0 < 15.0 and self.proc.poll() is None:
                time.sleep(0.05)
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=3)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i] == "Active" for r in self.rows)]
        pw = []
        for r in self.rows:
            try:
                pw.append(float(r[6]))
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "sm_mhz_min": min(sm) if sm else None,
                "power_w_max": max(pw) if pw else None}


def gen_device_iq(torch, n_signal, blocksize, n_blocks, seed, device, max_delay=64, ref_seed=None, delay_steps=None):
    """Synthetic IQ of SURVEY 8d generated on the device (same model as synth.make_blocks).
    ref_seed: seed of the reference-noise stream when it is shared by several channel groups
    (config 4: every rank's 21 channels see the same reference); default = seed.
    delay_steps: [(block, channel 1.., new_delay)] -- the channel's delay changes from that block on (drift /
    re-sync scenario of config 5); returns the per-block delay matrix as 4th value when given."""
    L = blocksize // 2
    gr = torch.Generator(device=device); gr.manual_seed(seed if ref_seed is None else ref_seed)
    g = torch.Generator(device=device); g.manual_seed(seed ^ 0x5EED)
    rng = np.random.Generator(np.random.PCG64(seed))
    gains = rng.uniform(0.7, 1.0, n_signal); phases = rng.uniform(-np.pi, np.pi, n_signal)
    delays = rng.integers(-max_delay, max_delay + 1, n_signal)
    pad = max_delay + 1
    T = n_blocks * L + 2 * pad
    r = torch.randn(T, 2, generator=gr, device=device) * 40.0
    iq = torch.empty(n_blocks, n_signal + 1, blocksize, dtype=torch.uint8, device=device)

    def quant(v):
        return torch.clamp(torch.round(v + 127.5), 0, 255).to(torch.uint8)

    iq[:, 0, :] = quant(r[pad:pad + n_blocks * L]).reshape(n_blocks, blocksize)
    dmat = np.tile(np.asarray(delays, dtype=np.int64), (n_blocks, 1))
    for (b0, ch, nd) in (delay_steps or []):
        assert abs(nd) <= max_delay
        dmat[b0:, ch - 1] = nd
    ar = torch.arange(L, device=device)
    for k in range(n_signal):
        if delay_steps:
            start = torch.as_tensor(pad + np.arange(n_blocks) * L - dmat[:, k], device=device)
            seg = r[(start[:, None] + ar[None, :]).reshape(-1)]
        else:
            s = pad - int(delays[k])
            seg = r[s:s + n_blocks * L]
        c, sn = float(gains[k] * np.cos(phases[k])), float(gains[k] * np.sin(phases[k]))
        w = torch.randn(n_blocks * L, 2, generator=g, device=device) * 4.0
        x = torch.stack((seg[:, 0] * c - seg[:, 1] * sn, seg[:, 0] * sn + seg[:, 1] * c), dim=1) + w
        iq[:, k + 1, :] = quant(x).reshape(n_blocks, blocksize)
    if delay_steps:
        return iq, delays, phases, dmat
    return iq, delays, phases


# ------------------------------------------------------------------ CPU baseline / reference arm
# kind "reference": the reference's OWN translation units (oracle/_ref/libref_cdsp.so: cdsp.cc,
# ccoherent.cc, csdrdevice.cc, crtlsdr.cc, cpacketizer.cc compiled unmodified against shims) fed
# block by block through csdrdevice::swapbuffer -- its single DSP thread per process, as designed
# (src/ccoherent.cc:245).  nfft = C+1 lets it correlate every channel of a block in SYNC mode (the
# stock value 8 would measure only 7 channels per block).  kind "port": the C restatement
# (oracle/liboracle.so), used only where the prebuilt reference library is absent.
def _ref_worker(q_in, q_out, n_signal, B, mode, reps):
    from oracle import oracle as O
    iq = q_in.get()
    nb = iq.shape[0]
    req = np.zeros((nb, n_signal + 1), np.uint8)
    if mode == "sync":
        req[:, 1:] = 1
    with O.RefCdsp(n_signal, B, nfft=n_signal + 1) as R:
        R.run_batch(iq[:1], lag_request=req[:1])               # warm the instance up
        q_out.put("ready"); q_in.get()                         # start line
        t0 = time.perf_counter()
        for _ in range(reps):
            R.run_batch(iq, lag_request=req)
        q_out.put(time.perf_counter() - t0)


def cpu_pass(O, iq, mode, threads, reps=1):
    """`reps` passes of the CPU path over iq[nb][C+1][B] on `threads` host cores (contiguous block
    ranges, one DSP thread / state per range); returns (seconds, kind)."""
    nb, nch, B = iq.shape
    ranges = [(i * nb // threads, (i + 1) * nb // threads) for i in range(threads)]
    ranges = [r for r in ranges if r[1] > r[0]]
    if O.ref_available():
        import multiprocessing as mp
        ctx = mp.get_context("spawn")                          # one process per DSP thread (cpacketize is static)
        procs = []
        for r in ranges:
            qi, qo = ctx.Queue(), ctx.Queue()
            pr = ctx.Process(target=_ref_worker, args=(qi, qo, nch - 1, B, mode, reps), daemon=True)
            pr.start(); qi.put(np.ascontiguousarray(iq[r[0]:r[1]]))
            procs.append((pr, qi, qo))
        for _, _, qo in procs:
            assert qo.get(timeout=300) == "ready"
        t0 = time.perf_counter()
        for _, qi, _ in procs:
            qi.put("go")
        times = [qo.get(timeout=3600) for _, _, qo in procs]
        wall = time.perf_counter() - t0
        for pr, _, _ in procs:
            pr.join(timeout=30)
        return max(max(times), wall if threads > 1 else 0.0), "reference"
    cfg = O.make_config(nch - 1, B)
    lag_mode = 1 if mode == "sync" else 0

    def work(r):
        for _ in range(reps):
            O.process_batch(cfg, iq[r[0]:r[1]], lag_mode=lag_mode)

    t0 = time.perf_counter()
    if len(ranges) == 1:
        work(ranges[0])
    else:
        ts = [threading.Thread(target=work, args=(r,)) for r in ranges]
        [t.start() for t in ts]
        [t.join() for t in ts]
    return time.perf_counter() - t0, "port"


def workload_name(L, nb, mode):
    """config.workload, the same string on both arms"""
    return (f"{CONFIG}: {C_SIG}+1 channels x {L}-sample blocks x {nb} blocks per GPU "
            f"(BASELINE.json configs[{CONFIG_INDEX[CONFIG]}]), {mode.upper()} mode, software shift on")


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import oracle as O
    cohpkg = __import__("cohpkg")
    synth = cohpkg.synth()
    cores = os.cpu_count() or 1
    mode = args.mode
    # one step = the GPU arm's whole step (N_BLOCKS blocks of the preset: 0.4 s of SYNC work on 16 cores);
    # --cpu-blocks bounds it to a sample on small hosts, and the line then says so
    sample_blocks = args.cpu_blocks or N_BLOCKS
    iq, _ = synth.make_blocks(C_SIG, BLOCKSIZE, sample_blocks, seed=0xC0DE + 1)
    if args.warmup:
        cpu_pass(O, iq[:max(cores, 1)], mode, cores, reps=args.warmup)
    t, kind = cpu_pass(O, iq, mode, cores, reps=args.steps)
    samples = (C_SIG + 1) * (BLOCKSIZE // 2) * sample_blocks * args.steps
    v = samples / t
    b3 = cpu_b3(O, iq[:min(sample_blocks, 12)], mode, cores)
    line = {
        "impl": "reference", "metric": "coherent samples/sec", "value": v, "unit": "samples/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(BLOCKSIZE // 2, sample_blocks, mode),
                   "mode": mode, "blocks_per_gpu": sample_blocks, "channels": C_SIG + 1, "L": BLOCKSIZE // 2,
                   "host_cores": cores,
                   "sample": "whole step" if sample_blocks == N_BLOCKS else f"{sample_blocks} of {N_BLOCKS} blocks per step"},
        "cpu_baseline": {"value": v, "unit": "samples/s", "cores": cores, "kind": kind,
                         "sample": f"{sample_blocks} blocks x {C_SIG + 1} channels per step, {args.steps} steps, "
                                   f"{cores} DSP threads (one process each) over contiguous block ranges"},
        "e2e": {"value": v, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "cpu_b3_numpy_scipy": b3,
    }
    emit(line)
    return 0


def cpu_b3(O, iq, mode, cores):
    """the numpy / scipy.fft (pocketfft, complex64) restatement of the same path, timed
    beside the reference's own translation units, and one 16384-point transform through the shim FFT that
    stands in for fftw3f against pocketfft -- so that the CPU arm's FFT speed is on the record."""
    try:
        nb, nch, B = iq.shape
        lag_mode = 1 if mode == "sync" else 0
        O.np_process_batch_c64(iq[:1], lag_mode)
        t0 = time.perf_counter(); O.np_process_batch_c64(iq, lag_mode, workers=1); t1 = time.perf_counter() - t0
        t0 = time.perf_counter(); O.np_process_batch_c64(iq, lag_mode, workers=cores); tn = time.perf_counter() - t0
        n = nch * (B // 2) * nb
        return {"value_1_thread": n / t1, "value_fft_workers": n / tn, "fft_workers": cores, "unit": "samples/s",
                "sample": f"{nb} blocks x {nch} channels, {mode} mode", "fft_us_per_16384": O.fft_us(16384)}
    except Exception as e:
        return {"error": f"{type(e).__name__}: {e}"[:200]}


def bind_to_gpu_numa_node(index):
    """Run this rank on the CPU cores next to its GPU (NVML's ideal affinity) so that the pinned
    host buffers of the e2e path are first-touched on that NUMA node; harmless if unavailable."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = [i for i in range(ncpu) if (words[i // 64] >> (i % 64)) & 1]
        allowed = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in allowed]
        if cpus:
            os.sched_setaffinity(0, cpus)
    except Exception:
        pass


# ------------------------------------------------------------------------ gather (configs[3])
class _DevBytes:
    """a raw device pointer as something torch.as_tensor accepts (no copy)"""
    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|i1", "data": (int(ptr), False), "version": 2}


def run_with_gather(torch, dist, pkg, eng, iq, args, rank, world, local, dev, stream, timed, barrier, nb, nch, B, L):
    """The sharded product path: every rank's engine runs its channel group with gather = 1; the
    library moves the packed columns into the root's [hdr][readcnt][1 + 21 N columns] packets while
    the next chunk is computed.  Verified against ONE engine that processes all channels."""
    import ctypes
    N = 1 + world * C_SIG
    counts = [C_SIG] * world
    total_samples = N * L * nb                       # one shared reference column + every rank's 21 columns
    out = {"config": f"{world} x {C_SIG} = {world * C_SIG} signal channels + 1 shared reference, channel-group sharded, "
                     f"{nb} blocks of {L} samples; gathered packet = {N} columns",
           "transports": {}}
    init_state = eng.get_state()                     # carried state at this point (restored before each verification)

    def reset_state():
        st = (pkg.ChanState * nch)()
        for c in range(nch):
            st[c].phasecorr[0] = st[c].phasecorrprev[0] = 1.0
        eng.set_state(st)

    def step(mode):
        eng.batch_raw(iq=iq.data_ptr(), iq_on_device=1, n_blocks=nb, lag_mode=pkg.LAG_ALL if mode == "sync" else pkg.LAG_NONE,
                      refnoise_enabled=1, gather=1)

    single = None
    for name in args.gather_transports.split(","):
        tr = pkg.TRANSPORTS[name]
        res = {}
        err = None
        try:
            geo = eng.gather_init(rank, world, counts, pkg.SHARD_CHANNELS, tr, max_blocks=nb,
                                  chunk_blocks=args.gather_chunk, store_ctas=args.store_ctas)
        except Exception as e:
            err = f"{type(e).__name__}: {e}"[:300]
        blob = [None]
        if rank == 0 and err is None:
            try:
                blob = [eng.gather_export()]
            except Exception as e:
                err = f"{type(e).__name__}: {e}"[:300]
        dist.broadcast_object_list(blob, src=0)
        if err is None and blob[0] is not None:
            try:
                eng.gather_attach(blob[0])
            except Exception as e:
                err = f"{type(e).__name__}: {e}"[:300]
        elif err is None:
            err = "root could not export"
        # a transport the box cannot do (no peer access, no libnccl) is reported, not fatal -- on every rank alike
        flag = torch.tensor([0 if err is None else 1], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MAX)
        if int(flag.item()):
            errs = [None] * world
            dist.all_gather_object(errs, err)
            out["transports"][name] = {"error": next((e for e in errs if e), "failed on another rank")}
            continue
        if True:
            barrier()
            # ---- verification: fresh state everywhere, one SYNC step, compare with a single engine ------
            reset_state()
            step("sync"); eng.sync(); barrier()
            if rank == 0:
                if single is None:
                    big = torch.empty((nb, N, B), dtype=torch.uint8, device=dev)
                    for g in range(world):
                        part, _, _ = gen_device_iq(torch, C_SIG, B, nb, 0xC0DE + 1 + 1000 * g, dev, ref_seed=0xC0DE + 1)
                        if g == 0:
                            big[:, 0, :] = part[:, 0, :]
                        big[:, 1 + g * C_SIG:1 + (g + 1) * C_SIG, :] = part[:, 1:, :]
                        del part
                    e1 = pkg.CoherentCuda(world * C_SIG, B, apply_shift=True, device=local, max_batch_blocks=nb,
                                          stream=stream.cuda_stream, alpha=args.alpha)
                    assert (e1.packet_len, e1.dev_pitch, e1.dev_lead) == (geo.packet_len, geo.dev_pitch, geo.dev_lead)
                    pk1 = torch.zeros(nb * e1.dev_pitch + 16, dtype=torch.int8, device=dev)
                    e1.batch_raw(iq=big.data_ptr(), iq_on_device=1, n_blocks=nb, lag_mode=pkg.LAG_ALL, refnoise_enabled=1,
                                 packets=pk1.data_ptr(), packets_on_device=1)
                    e1.sync(); e1.close()
                    single = pk1[:nb * e1.dev_pitch].view(nb, e1.dev_pitch)[:, geo.dev_lead:geo.dev_lead + geo.packet_len]
                    # the checker: the first two blocks through the CPU oracle (fresh state, integer-dot specification)
                    oracle_rows = None
                    if not args.no_cpu:
                        try:
                            from oracle import oracle as O
                            ocfg = O.make_config(world * C_SIG, B, apply_shift=1, int_dot=1)
                            oracle_rows = O.process_batch(ocfg, big[:2].cpu().numpy(), lag_mode=1)["packets"]
                        except Exception as e:
                            oracle_rows = f"{type(e).__name__}: {e}"[:200]
                    del big
                # every byte of every gathered packet, compared on the device (the gathered buffer has e1's layout)
                gp = torch.as_tensor(_DevBytes(eng.gather_packets_ptr(), nb * geo.dev_pitch), device=dev)
                gp = gp.view(nb, geo.dev_pitch)[:, geo.dev_lead:geo.dev_lead + geo.packet_len]
                res["verified_blocks"] = nb
                res["bytes_identical_to_single_gpu"] = bool(torch.equal(gp, single))
                if isinstance(oracle_rows, np.ndarray):
                    res["oracle_blocks"] = int(oracle_rows.shape[0])
                    res["bytes_identical_to_int_dot_oracle"] = bool((eng.gather_fetch(0, oracle_rows.shape[0]) == oracle_rows).all())
                elif oracle_rows is not None:
                    res["oracle_check_error"] = oracle_rows
                del gp
            # ---- timing: SYNC and TRACK ---------------------------------------------------------------
            for mode in ("sync", "track"):
                b0 = eng.gather_bytes()
                ms = timed(lambda: step(mode), args.steps, 3)
                moved = (eng.gather_bytes() - b0) / (args.steps + 3)       # payload per step at this rank
                t = torch.tensor([float(moved)], device=dev, dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                root_bytes = (world - 1) * C_SIG * B * nb                  # what crosses NVLink into the root per step
                res[mode] = {"value": total_samples * args.steps / (ms * 1e-3), "unit": "samples/s",
                             "ms_per_step": ms / args.steps,
                             "root_ingest_gbs": root_bytes * args.steps / (ms * 1e-3) / 1e9,
                             "root_ingest_bytes_per_step": root_bytes, "payload_bytes_counted_by_library": int(t.item())}
        out["transports"][name] = res
        barrier()
    eng.set_state(init_state)
    best = max((r["sync"]["value"] for r in out["transports"].values() if "sync" in r), default=None)
    out["value"] = best; out["unit"] = "samples/s"
    out["collective"] = ("gather of the packed int8 columns to rank 0 inside libcoherent_cuda; limit = NVLink ingest of the "
                         "root (measured peer copy 770 GB/s per direction, B200_PROFILING.md)")
    return out


# ------------------------------------------------------------------------------------ GPU arm
def run_gpu_arm(args):
    import torch
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL's version / debug lines go to stdout by default: keep stdout for the ONE JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    bind_to_gpu_numa_node(local)
    import cohpkg
    pkg = cohpkg.load()
    stream = torch.cuda.Stream(device=dev)
    nb, nch, B, L = args.blocks, C_SIG + 1, BLOCKSIZE, BLOCKSIZE // 2
    eng = pkg.CoherentCuda(C_SIG, B, apply_shift=True, device=local, max_batch_blocks=nb,
                           stream=stream.cuda_stream, alpha=args.alpha)
    # each rank holds its own 21 signal channels and its own copy of the ONE shared reference channel
    dmat = None
    if args.mode == "mixed":
        # config 5's drift re-sync: delays STEP at known blocks, the channel asks for its lag again there
        # (src/console.cc:285-289 -> include/csdrdevice.h:129-151), every fifth request three blocks late
        iq, delays, _, dmat = gen_device_iq(torch, C_SIG, B, nb, 0xC0DE + 1 + 1000 * rank, dev, ref_seed=0xC0DE + 1,
                                            delay_steps=STEPS)
    else:
        iq, delays, _ = gen_device_iq(torch, C_SIG, B, nb, 0xC0DE + 1 + 1000 * rank, dev, ref_seed=0xC0DE + 1)
    pk_dev = torch.zeros(nb * eng.dev_pitch + 16, dtype=torch.int8, device=dev)
    torch.cuda.synchronize()
    samples_per_step = nch * L * nb

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if dist is None:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def step_dev(mode):
        eng.batch_raw(iq=iq.data_ptr(), iq_on_device=1, n_blocks=nb,
                      lag_request=REQ if mode == "mixed" else None,
                      lag_mode=pkg.LAG_ALL if mode == "sync" else pkg.LAG_NONE,
                      refnoise_enabled=0 if args.no_refnoise else 1,
                      packets=pk_dev.data_ptr(), packets_on_device=1)

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        eng.flush()          # host-buffer steps: the last packet read-back belongs to the timed region
        e1.record(stream)
        barrier()
        return max_over_ranks(e0.elapsed_time(e1))

    hbm_peak, peak_src = measured_peaks()
    sampler = ClockSampler(local)
    # the acquisition pass leaves the measured lags held, so TRACK runs with real shifts applied
    step_dev("sync"); eng.sync()
    launches0 = eng.launch_count()
    if rank == 0:
        sampler.start()
    ms_main = timed(lambda: step_dev(args.mode), args.steps, args.warmup)
    launches = (eng.launch_count() - launches0) * args.steps // (args.steps + args.warmup)
    clocks = sampler.stop() if rank == 0 else None
    other = "sync" if args.mode == "track" else "track"
    ms_other = timed(lambda: step_dev(other), args.steps, args.warmup)

    # per-kernel durations (CUDA events on the launching stream, one launch at a time)
    def profile(mode):
        eng.profile(True)
        for _ in range(2):
            step_dev(mode)
        eng.sync()
        p = eng.profile_read()
        eng.profile(False)
        return p
    prof = {m: profile(m) for m in {"sync", "track", args.mode}}
    correlate_dominates = prof[args.mode]["k_correlate"]["ms"] > prof[args.mode]["k_apply"]["ms"]

    def roofline(mode):
        if mode != "track" and correlate_dominates:
            k = prof[mode]["k_correlate"]; units = 2 * C_SIG * nb * L      # two profiled steps
            name = "k_correlate"
        else:
            k = prof[mode]["k_apply"]; units = 2 * nch * nb * L
            name = "k_apply"
        gbs = units * BYTES_PER_SAMPLE / (k["ms"] * 1e-3) / 1e9
        traffic, traffic_src = ncu_traffic(name) if (CONFIG == "ura21" and nb == 1024 and mode != "mixed") else (None, None)
        r = {"kernel": name, "bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s",
             "frac": gbs / hbm_peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
             "launches_timed": k["launches"], "avg_launch_ms": k["ms"] / max(k["launches"], 1),
             "algorithmic_bytes_per_launch": units * BYTES_PER_SAMPLE / max(k["launches"], 1)}
        if mode == "sync" and L == 8192:
            # the kernel's real ceiling is the FP32 pipe (76 FLOP/B against a machine balance of ~11):
            # 148 SMs x 128 lanes x 2 flop x 1.965 GHz; FFT butterflies are add-heavy, so even a
            # permanently busy FMA pipe delivers ~55-60 % of that figure
            r["fp32_tflops_algorithmic"] = units * FLOP_PER_SAMPLE_SYNC / (k["ms"] * 1e-3) / 1e12
            r["fp32_peak_tflops"] = 74.4
            r["fp32_frac"] = r["fp32_tflops_algorithmic"] / 74.4
            r["note"] = "FP32-pipe bound, not HBM bound; frac is against HBM as the contract asks, fp32_frac against the FP32 peak"
        return r

    # ---- end to end through the C ABI with pinned host buffers ------------------------------
    h_iq = torch.empty((nb, nch, B), dtype=torch.uint8, pin_memory=True); h_iq.copy_(iq)
    h_pk = torch.empty((nb, eng.packet_len), dtype=torch.int8, pin_memory=True)
    h_lag = torch.empty((nb, nch), dtype=torch.float32, pin_memory=True)
    h_mag = torch.empty((nb, nch), dtype=torch.float32, pin_memory=True)
    h_ph = torch.empty((nb, nch, 2), dtype=torch.float32, pin_memory=True)

    def step_e2e():
        eng.batch_raw(iq=h_iq.data_ptr(), iq_on_device=0, n_blocks=nb,
                      lag_request=REQ if args.mode == "mixed" else None,
                      lag_mode=pkg.LAG_ALL if args.mode == "sync" else pkg.LAG_NONE, refnoise_enabled=1,
                      packets=h_pk.data_ptr(), packets_on_device=0, lag=h_lag.data_ptr(),
                      mag=h_mag.data_ptr(), phasecorr=h_ph.data_ptr())
    ms_e2e = timed(step_e2e, max(args.steps // 2, 2), 3)
    e2e_steps = max(args.steps // 2, 2)
    lag_ok = bool((h_lag[:, 1:].numpy() == delays[None, :]).all()) if args.mode == "sync" else None
    if args.mode == "mixed":
        # every served request must return the delay in force AT THAT BLOCK, and the lag held afterwards is that one
        served = REQ[:, 1:].astype(bool)
        lag_ok = bool((h_lag[:, 1:].numpy()[served] == dmat[served]).all())

    # ---- sustained run (BASELINE configs[4] is "sustained throughput"): seconds, not a 50 ms burst ----------
    sustained = None
    if args.sustained > 0:
        sustained = {}
        for m, mult in ((args.mode, 1), (other, 1)):
            n_steps = int(args.sustained * (12 if m == "track" else 1))
            smp = ClockSampler(local)
            if rank == 0:
                smp.start()
            ms = timed(lambda: step_dev(m), n_steps, 3)
            ck = smp.stop() if rank == 0 else None
            sustained[m] = {"steps": n_steps, "seconds": ms * 1e-3, "value": world * samples_per_step * n_steps / (ms * 1e-3),
                            "unit": "samples/s", "clocks": ck}

    with_gather = None
    if dist is not None:
        with_gather = run_with_gather(torch, dist, pkg, eng, iq, args, rank, world, local, dev, stream,
                                      timed, barrier, nb, nch, B, L)

    if rank == 0:
        val = lambda ms, steps: world * samples_per_step * steps / (ms * 1e-3)
        line = {
            "metric": "coherent samples/sec", "value": val(ms_main, args.steps), "unit": "samples/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_main / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(L, nb, args.mode),
                       "mode": args.mode, "blocks_per_gpu": nb, "channels": nch, "L": L,
                       "parallelism": f"{world} x {C_SIG} signal channels, one shared reference channel replicated to every rank" if world > 1 else "single GPU",
                       "l2": f"input {nb * nch * B >> 20} MiB + output {nb * nch * B >> 20} MiB per step exceed the 126 MB L2 (no flush needed)"},
            "roofline": roofline(args.mode),
            "modes": {args.mode: {"value": val(ms_main, args.steps), "ms_per_step": ms_main / args.steps},
                      other: {"value": val(ms_other, args.steps), "ms_per_step": ms_other / args.steps,
                              "roofline": roofline(other)}},
            "kernel_ms_per_step": {m: {k: v["ms"] / 2 for k, v in prof[m].items()} for m in prof},
            "e2e": {"value": val(ms_e2e, e2e_steps), "unit": "samples/s",
                    "h2d_bytes_per_step": world * int(h_iq.numel()),
                    "d2h_bytes_per_step": world * int(h_pk.numel() + 4 * h_lag.numel() + 4 * h_mag.numel() + 4 * h_ph.numel()),
                    "ms_per_step": ms_e2e / e2e_steps, "lag_known_answer_ok": lag_ok},
            "gpu_launches": int(launches),
            "clocks": clocks,
        }
        if with_gather:
            # compact scalars first (a truncated line still carries them), the full record last
            trs = {k: v for k, v in with_gather["transports"].items() if "sync" in v}
            if trs:
                best = max(trs, key=lambda k: trs[k]["sync"]["value"])
                summary = {"gather_best": best, "gather_value": trs[best]["sync"]["value"],
                           "gather_ms_per_step": trs[best]["sync"]["ms_per_step"],
                           "gather_efficiency": (ms_main / args.steps) / trs[best]["sync"]["ms_per_step"] if args.mode == "sync" else None,
                           "root_ingest_gbs": trs[best]["sync"]["root_ingest_gbs"],
                           "gather_verified": all(v.get("bytes_identical_to_single_gpu") for v in trs.values()),
                           "gather_limit": "root NVLink ingest" if trs[best]["sync"]["root_ingest_gbs"] > 600 else "chunked schedule (link not saturated)"}
                line = {**{k: line[k] for k in ("metric", "value", "unit", "n_gpus")}, **summary,
                        **{k: v for k, v in line.items() if k not in ("metric", "value", "unit", "n_gpus")}}
            line["with_gather"] = with_gather
        if sustained:
            line["sustained"] = sustained
        if world == 1 and not args.no_cpu:
            from oracle import oracle as O
            synth = cohpkg.synth()
            probe, _ = synth.make_blocks(C_SIG, B, 4, seed=0xC0DE + 1)
            t_probe, _ = cpu_pass(O, probe, args.mode, 1)
            n_cpu = int(min(max(4, 4 * args.cpu_seconds / max(t_probe, 1e-6)), 2048))
            cpu_iq, _ = synth.make_blocks(C_SIG, B, n_cpu, seed=0xC0DE + 1)
            t_cpu, kind = cpu_pass(O, cpu_iq, args.mode, 1)
            line["cpu_b3_numpy_scipy"] = cpu_b3(O, cpu_iq[:8], "track" if args.mode == "mixed" else args.mode, os.cpu_count() or 1)
            line["cpu_baseline"] = {"value": nch * L * n_cpu / t_cpu, "unit": "samples/s", "cores": 1,
                                    "kind": kind,
                                    "sample": f"{n_cpu} blocks x {nch} channels of the same workload, "
                                              f"{'track' if args.mode == 'mixed' else args.mode} mode, "
                                              f"1 thread (the reference's single DSP thread), {t_cpu:.1f} s"}
        emit(line)
    eng.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


_JSON_FD = None


def _reserve_stdout():
    """stdout carries exactly ONE line, the JSON result.  Libraries write there too (NCCL prints its
    version banner on fd 1 whatever NCCL_DEBUG_FILE says), so fd 1 is pointed at stderr for the whole
    run and the JSON line goes out through a private duplicate of the original stdout."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    os.write(_JSON_FD if _JSON_FD is not None else 1, data)


def main():
    _reserve_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mode", default="sync", choices=["sync", "track", "mixed"])
    ap.add_argument("--config", default="ura21", choices=list(PRESETS))
    ap.add_argument("--blocks", type=int, default=1024)
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    ap.add_argument("--cpu-blocks", type=int, default=0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--alpha", type=float, default=0.5, help="phase IIR weight (reference: 0.5)")
    ap.add_argument("--no-refnoise", action="store_true", help="diagnostic: noise source off (no phase estimate)")
    ap.add_argument("--sustained", type=int, default=500, help="also time this many back-to-back SYNC steps (12x as many TRACK steps) with the clock / power record")
    ap.add_argument("--gather-transports", default="nccl,peer_copy,peer_store")
    ap.add_argument("--gather-chunk", type=int, default=0, help="blocks per gather transfer (0 = library default)")
    ap.add_argument("--store-ctas", type=int, default=0, help="CTAs of the NVLink-storing k_apply (0 = library default)")
    ap.add_argument("--gather-verify-all", action="store_true", help="(kept for old command lines: every gathered block is compared by default)")
    args = ap.parse_args()
    global C_SIG, BLOCKSIZE, N_BLOCKS, CONFIG, REQ, STEPS
    CONFIG = args.config
    C_SIG, BLOCKSIZE, N_BLOCKS = PRESETS[CONFIG]
    if args.blocks == 1024:
        args.blocks = N_BLOCKS
    if args.mode == "mixed":
        rng = np.random.default_rng(5)
        # about 1 % of the (block, channel) items re-request a lag; every one of them follows a delay step
        n_steps = max(1, int(0.01 * args.blocks * C_SIG))
        REQ = np.zeros((args.blocks, C_SIG + 1), np.uint8); REQ[0, 1:] = 1
        STEPS = []
        for i in range(n_steps):
            b0 = int(rng.integers(1, max(args.blocks - 4, 2))); ch = int(rng.integers(1, C_SIG + 1))
            STEPS.append((b0, ch, int(rng.integers(-64, 65))))
        STEPS.sort()
        for i, (b0, ch, _) in enumerate(STEPS):
            REQ[min(b0 + (3 if i % 5 == 0 else 0), args.blocks - 1), ch] = 1
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
