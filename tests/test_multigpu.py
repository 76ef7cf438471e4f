"""Multi-GPU parity (needs >= 2 GPUs; skipped on a single-GPU box): one process per GPU, block-sharded
and channel-group-sharded, gathered to rank 0 THROUGH THE PRODUCT PATH -- libcoherent_cuda's device
gather (cohcuda_gather_*, csrc/api_gather.inc) with each of its transports.  The gathered matrix must
equal the single-GPU engine's output byte for byte (SURVEY section 4, item v) and the oracle's.
The packet layout being matched: include/cpacketizer.h:32-37, src/cpacketizer.cc:91-96,137-172."""
import os

import numpy as np
import pytest

from test_sharding import make_case, run_world

pytestmark = pytest.mark.gpu

TRANSPORTS = ["nccl", "peer_copy", "peer_store"]


def _ngpu():
    import torch
    return torch.cuda.device_count()


def _check(pkg, oracle, got, iq, req, noise, C, B, slots, shift=1):
    eng = pkg.CoherentCuda(C, B, apply_shift=bool(shift), max_lag_slots=slots)
    one = eng.batch(iq, lag_request=req, refnoise=noise, readcnt0=8, globalseqn0=100)
    eng.close()
    np.testing.assert_array_equal(got["lag"], one["lag"])
    np.testing.assert_array_equal(got["phase"].view(np.uint32), one["phase"].view(np.uint32))
    np.testing.assert_array_equal(got["packets"], one["packets"])
    cfg = oracle.make_config(C, B, apply_shift=shift, int_dot=1, max_lag_slots=slots)
    ref = oracle.process_batch(cfg, iq, lag_request=req, refnoise=noise, readcnt0=8)
    np.testing.assert_array_equal(got["packets"][:, 4:], ref["packets"][:, 4:])     # all but globalseqn (offset 100)
    np.testing.assert_array_equal(got["lag"], ref["lag"])


@pytest.mark.parametrize("mode,C,B,nb,slots,transport", [
    ("channels", 21, 16384, 40, 7, "nccl"), ("channels", 21, 16384, 40, 7, "peer_copy"), ("channels", 21, 16384, 40, 7, "peer_store"),
    ("blocks", 21, 16384, 140, 0, "nccl"), ("blocks", 5, 512, 230, 0, "peer_copy"), ("blocks", 5, 512, 230, 0, "peer_store"),
    ("channels", 6, 2048, 64, 0, "nccl"),
    # the large-block path (fft_big.cu) under a gather, uneven channel groups (5 = 3 + 2 on two GPUs)
    ("channels", 5, 32768, 20, 0, "nccl"), ("channels", 5, 32768, 20, 0, "peer_copy"), ("blocks", 3, 65536, 24, 0, "peer_store")])
def test_gathered_matrix_equals_single_gpu(pkg, synth, oracle, tmp_path, mode, C, B, nb, slots, transport):
    world = min(_ngpu(), 8, C if mode == "channels" else 8)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    case, out = str(tmp_path / "case.npz"), str(tmp_path / "out.npz")
    iq, req, noise, _ = make_case(synth, case, C, B, nb, seed=21 + C, noheader=0, shift=1, slots=slots)
    # 16-block transfers: several chunks per rank, the last one ragged
    run_world(mode, case, out, world=world, engine="cuda", transport=transport, env={"COH_TEST_CHUNK": "16"})
    _check(pkg, oracle, np.load(out), iq, req, noise, C, B, slots)


@pytest.mark.parametrize("transport", TRANSPORTS)
def test_config4_168_channels_over_8_gpus(pkg, synth, oracle, tmp_path, transport):
    """BASELINE.json configs[3]: 168 = 8 x 21 signal channels + ONE shared reference, channel groups of 21
    per GPU, one 169-column packet per block (packetlength(N = 169), src/cpacketizer.cc:91-96)."""
    if _ngpu() < 8:
        pytest.skip("needs 8 GPUs")
    C, B, nb = 168, 16384, 20
    case, out = str(tmp_path / "case.npz"), str(tmp_path / "out.npz")
    iq, req, noise, _ = make_case(synth, case, C, B, nb, seed=168, noheader=0, shift=1, slots=0, req_p=0.3)
    run_world("channels", case, out, world=8, engine="cuda", transport=transport, env={"COH_TEST_CHUNK": "8"})
    got = np.load(out)
    assert got["packets"].shape == (nb, 16 + 4 * 169 + 169 * B)
    h = got["packets"][:, :16].copy().view(np.uint32)
    assert (h[:, 1] == 169).all() and (h[:, 2] == B // 2).all()
    _check(pkg, oracle, got, iq, req, noise, C, B, 0)


def test_block_shards_across_a_delay_step(pkg, synth, oracle, tmp_path):
    """Drift re-sync over shard boundaries (BASELINE configs[4]): a delay steps a few blocks BEFORE a shard
    starts and is re-measured late, inside the next shard's warm-up halo; another steps inside a shard.
    plan_block_shard must restore the right held lag on every rank."""
    world = min(_ngpu(), 4)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    C, B, nb = 5, 2048, 240
    edge = nb // world                                # first block of rank 1
    steps = [(edge - 4, 2, 37), (edge + 9, 4, -18), (nb - 30, 1, 3)]
    iq, _ = synth.make_blocks(C, B, nb, seed=77, delays=np.array([0, 5, -5, 9, 60]), delay_steps=steps)
    req = np.zeros((nb, C + 1), np.uint8); req[0, 1:] = 1
    req[edge - 2, 2] = 1; req[edge + 9, 4] = 1; req[nb - 25, 1] = 1
    noise = np.ones(nb, np.uint8); noise[edge - 20: edge - 12] = 0
    case, out = str(tmp_path / "case.npz"), str(tmp_path / "out.npz")
    np.savez(case, iq=iq, lag_request=req, refnoise=noise, meta=np.array([C, B, 0, 1, 0]))
    run_world("blocks", case, out, world=world, engine="cuda", transport="nccl", env={"COH_TEST_CHUNK": "16"})
    _check(pkg, oracle, np.load(out), iq, req, noise, C, B, 0)
