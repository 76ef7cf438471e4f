"""Host-side multi-GPU logic on CPU: partition planning, halo/state reconstruction and the gather
to rank 0 (torch.distributed, gloo, world_size 2).  The per-rank engine is an oracle-backed
stand-in (tests/_shard_worker.py) so that what is tested here is exactly the code in
coherent-rtlsdr_b200/sharding.py that also drives the CUDA engines; the gathered matrix must equal
the single-engine result byte for byte (SURVEY section 4, item v)."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def S(pkg):
    import importlib
    return importlib.import_module("coherent_rtlsdr_b200.sharding")


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def make_case(synth, path, C, B, nb, seed, noheader=0, shift=1, slots=0, gated=True, req_p=0.08):
    rng = np.random.default_rng(seed)
    delays = rng.integers(-B // 16, B // 16 + 1, C)
    iq, _ = synth.make_blocks(C, B, nb, seed=seed, delays=delays)
    req = np.zeros((nb, C + 1), np.uint8)
    req[0, 1:] = 1                                   # acquisition
    req[1:, 1:] = rng.random((nb - 1, C)) < req_p    # sporadic re-sync requests
    noise = np.ones(nb, np.uint8)
    if gated:
        noise[nb // 3: nb // 3 + 7] = 0
    np.savez(path, iq=iq, lag_request=req, refnoise=noise, meta=np.array([C, B, noheader, shift, slots]))
    return iq, req, noise, delays


def run_world(mode, case, out, world=2, engine="oracle", transport="host", env=None):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", "_shard_worker.py"), "--mode", mode, "--engine", engine,
           "--case", case, "--out", out, "--transport", transport]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT, env=dict(os.environ, **(env or {})))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]


def single(oracle, iq, req, noise, C, B, noheader, shift, slots):
    cfg = oracle.make_config(C, B, noheader=noheader, apply_shift=shift, int_dot=1, max_lag_slots=slots)
    r = oracle.process_batch(cfg, iq, lag_request=req, refnoise=noise, readcnt0=8)
    if not noheader:
        gs = (100 + np.arange(iq.shape[0])).astype(np.uint32)
        r["packets"][:, :4] = gs.view(np.int8).reshape(-1, 4)
    return r


def test_effective_requests_is_the_reference_slot_cap(S, synth, oracle):
    C, B, nb = 9, 512, 12
    iq, _ = synth.make_blocks(C, B, nb, seed=3, delays=np.arange(C) - 4)
    rng = np.random.default_rng(1)
    req = (rng.random((nb, C + 1)) < 0.8).astype(np.uint8)
    capped = oracle.process_batch(oracle.make_config(C, B, max_lag_slots=7), iq, lag_request=req)
    eff = S.effective_requests(req, 7)
    assert eff[:, 0].sum() == 0 and eff.sum(axis=1).max() <= 7
    free = oracle.process_batch(oracle.make_config(C, B, max_lag_slots=0), iq, lag_request=eff)
    np.testing.assert_array_equal(capped["packets"], free["packets"])
    np.testing.assert_array_equal(capped["lag"], free["lag"])


def test_partitions_cover_everything_once(S):
    for n, w in [(1024, 8), (10, 4), (3, 8), (0, 2), (1000003, 8)]:
        r = S.block_ranges(n, w)
        assert r[0][0] == 0 and r[-1][1] == n and all(a[1] == b[0] for a, b in zip(r, r[1:]))
        assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1
    for C, w in [(168, 8), (21, 2), (3, 8), (1, 1)]:
        g = S.channel_groups(C, w)
        assert g[0][0] == 1 and g[-1][1] == C + 1 and all(a[1] == b[0] for a, b in zip(g, g[1:]))
    assert S.lookback_depth(0.5) == 48 and S.lookback_depth(1.0) == 1


def test_block_shard_plan(S):
    nb, nch = 300, 4
    req = np.zeros((nb, nch), np.uint8); req[0, 1:] = 1; req[120, 2] = 1; req[250, 3] = 1
    noise = np.ones(nb, np.uint8); noise[180:190] = 0
    pre, h0 = S.plan_block_shard(req, noise, 200, 48)
    assert h0 == 200 - 49 - 10                        # 49 enabled blocks + the 10 gated ones
    assert [b for b, _ in pre] == [0, 120]            # ch1,ch3 last measured in block 0, ch2 in block 120
    assert pre[0][1].tolist() == [0, 1, 0, 1] and pre[1][1].tolist() == [0, 0, 1, 0]
    assert S.plan_block_shard(req, noise, 30, 48) == ([], 0)      # halo reaches the start: exact state
    assert S.plan_block_shard(req, noise, 0, 48) == ([], 0)


@pytest.mark.parametrize("mode,noheader,slots", [("blocks", 0, 0), ("channels", 0, 3), ("channels", 1, 0)])
def test_world2_gloo_gather_equals_single_engine(S, synth, oracle, tmp_path, mode, noheader, slots):
    C, B, nb = 5, 512, 230
    case, out = str(tmp_path / "case.npz"), str(tmp_path / "out.npz")
    iq, req, noise, _ = make_case(synth, case, C, B, nb, seed=11, noheader=noheader, shift=1, slots=slots)
    # transport "plan": rank 0 also assembles through cohcuda_gather_plan's geometry (the arithmetic the
    # device gather uses) and the worker asserts it equals the numpy assembly
    run_world(mode, case, out, world=2, transport="plan")
    got = np.load(out)
    want = single(oracle, iq, req, noise, C, B, noheader, 1, slots)
    np.testing.assert_array_equal(got["lag"], want["lag"])
    np.testing.assert_array_equal(got["phase"].view(np.uint32), want["phase"].view(np.uint32))
    np.testing.assert_array_equal(got["packets"], want["packets"])
    if mode == "blocks":
        assert got["halo"][1] > 0                     # rank 1 really started from a halo, not from block 0


def test_world3_block_shards_with_empty_and_short_ranges(S, synth, oracle, tmp_path):
    """Ragged case: fewer blocks than the halo, so every rank's halo reaches block 0."""
    C, B, nb = 2, 256, 7
    case, out = str(tmp_path / "case.npz"), str(tmp_path / "out.npz")
    iq, req, noise, _ = make_case(synth, case, C, B, nb, seed=5, gated=False)
    run_world("blocks", case, out, world=3)
    got = np.load(out)
    want = single(oracle, iq, req, noise, C, B, 0, 1, 0)
    np.testing.assert_array_equal(got["packets"], want["packets"])


def test_block_shards_across_a_delay_step(S, synth, oracle, tmp_path):
    """Drift re-sync over a shard boundary (BASELINE configs[4], src/console.cc:285-289): a delay steps four
    blocks BEFORE rank 1's range and is re-measured two blocks later (inside rank 1's warm-up halo); another
    steps inside rank 1's range.  The held lag each rank restores must be the one in force at its halo start."""
    C, B, nb = 5, 512, 240
    edge = nb // 2
    steps = [(edge - 4, 2, 37), (edge + 9, 4, -18), (nb - 30, 1, 3)]
    iq, _ = synth.make_blocks(C, B, nb, seed=77, delays=np.array([0, 5, -5, 9, 60]), delay_steps=steps)
    req = np.zeros((nb, C + 1), np.uint8); req[0, 1:] = 1
    req[edge - 2, 2] = 1; req[edge + 9, 4] = 1; req[nb - 25, 1] = 1
    noise = np.ones(nb, np.uint8); noise[edge - 20: edge - 12] = 0
    case, out = str(tmp_path / "case.npz"), str(tmp_path / "out.npz")
    np.savez(case, iq=iq, lag_request=req, refnoise=noise, meta=np.array([C, B, 0, 1, 0]))
    run_world("blocks", case, out, world=2, transport="plan")
    got = np.load(out)
    want = single(oracle, iq, req, noise, C, B, 0, 1, 0)
    assert want["lag"][edge, 2] == 37 and want["lag"][edge - 3, 2] == 5     # the step really is in force across the edge
    np.testing.assert_array_equal(got["lag"], want["lag"])
    np.testing.assert_array_equal(got["phase"].view(np.uint32), want["phase"].view(np.uint32))
    np.testing.assert_array_equal(got["packets"], want["packets"])
    assert got["halo"][1] > 0 and got["pre_calls"][1] >= 1
