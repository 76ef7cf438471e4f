"""Multi-GPU parity (needs >= 2 GPUs; skipped on a single-GPU box): one process per GPU over
NCCL, block-sharded and channel-group-sharded, gathered to rank 0 -- the gathered matrix must
equal the single-GPU engine's output byte for byte (SURVEY section 4, item v) and the oracle's."""
import os

import numpy as np
import pytest

from test_sharding import make_case, run_world

pytestmark = pytest.mark.gpu


def _ngpu():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("mode,C,B,nb,slots", [("blocks", 21, 16384, 140, 0), ("channels", 21, 16384, 40, 7),
                                               ("blocks", 5, 512, 230, 0), ("channels", 6, 2048, 64, 0)])
def test_gathered_matrix_equals_single_gpu(pkg, synth, oracle, tmp_path, mode, C, B, nb, slots):
    world = min(_ngpu(), 8)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    case, out = str(tmp_path / "case.npz"), str(tmp_path / "out.npz")
    iq, req, noise, _ = make_case(synth, case, C, B, nb, seed=21 + C, noheader=0, shift=1, slots=slots)
    run_world(mode, case, out, world=world, engine="cuda")
    got = np.load(out)
    eng = pkg.CoherentCuda(C, B, apply_shift=True, max_lag_slots=slots)
    one = eng.batch(iq, lag_request=req, refnoise=noise, readcnt0=8, globalseqn0=100)
    eng.close()
    np.testing.assert_array_equal(got["lag"], one["lag"])
    np.testing.assert_array_equal(got["phase"].view(np.uint32), one["phase"].view(np.uint32))
    np.testing.assert_array_equal(got["packets"], one["packets"])
    cfg = oracle.make_config(C, B, apply_shift=1, int_dot=1, max_lag_slots=slots)
    ref = oracle.process_batch(cfg, iq, lag_request=req, refnoise=noise, readcnt0=8)
    hdr = 16 + 4 * (C + 1)
    np.testing.assert_array_equal(got["packets"][:, 4:], ref["packets"][:, 4:])     # all but globalseqn (offset 100)
    np.testing.assert_array_equal(got["lag"], ref["lag"])
