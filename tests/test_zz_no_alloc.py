"""cohcuda_lags_static -- the lag measurement for callers that cannot allocate GPU memory.

Some users of libcoherent_cuda must not reach cudaMalloc: a receiver thread that has to hand over a block
every few milliseconds (the allocator serialises against the whole device), a plug-in inside a program
whose own pool already owns the GPU, a process with a fixed memory map.  For them the library carries its
working memory as a fixed array in its own device image and offers one handle-free entry point over it
(include/coherent_cuda.h).  This file checks that entry point against the ordinary one, cohcuda_batch with
lag_mode = COHCUDA_LAG_ALL, which allocates its buffers: same input, same lag and mag.

The fixed array is shared by all calls and keeps the previous call's contents, so every case here makes
at least two calls in a row -- with different inputs and different geometries -- and each must still match.

The file name sorts last on purpose: `pytest -x` stops at the first failure, and a failure here should
not hide the parity tests.  It passes alone in a fresh process and after the rest of the suite.
"""
import numpy as np
import pytest


def _batch_lags(pkg, iq, C, B):
    """lag / mag from the existing, allocating entry point"""
    eng = pkg.CoherentCuda(C, B)
    out = eng.batch(iq, lag_mode=pkg.LAG_ALL)
    eng.close()
    return out["lag"], out["mag"]


def _static_lags(pkg, iq, C, B):
    """lag / mag from cohcuda_lags_static; the caller (this test, through torch) owns input and outputs"""
    import torch
    nb = iq.shape[0]
    d_iq = torch.from_numpy(iq).cuda()
    d_lag = torch.full((nb, C + 1), float("nan"), dtype=torch.float32, device="cuda")
    d_mag = torch.full((nb, C + 1), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    pkg.lags_static(d_iq.data_ptr(), C, B, nb, d_lag.data_ptr(), d_mag.data_ptr())   # default stream
    torch.cuda.synchronize()
    return d_lag.cpu().numpy(), d_mag.cpu().numpy()


def _case(synth, C, B, nb, seed):
    L = B // 2
    rng = np.random.default_rng(seed)
    maxd = min(250, L // 4)
    delays = rng.integers(-maxd, maxd + 1, C)
    iq, _ = synth.make_blocks(C, B, nb, seed=seed, delays=delays)
    return iq, delays


def _check(pkg, iq, C, B, delays):
    lag_s, mag_s = _static_lags(pkg, iq, C, B)
    lag_b, mag_b = _batch_lags(pkg, iq, C, B)
    np.testing.assert_array_equal(lag_s, lag_b)                    # integer lag: bit-exact, as everywhere in tests/
    if B >= 512:                                                   # and it is the injected delay (tiny blocks: too few samples to insist)
        assert (lag_s[:, 1:] == delays[None, :]).all()
    np.testing.assert_allclose(mag_s, mag_b, rtol=1e-4)            # the suite's tolerance for mag ...
    np.testing.assert_array_equal(mag_s.view(np.uint32), mag_b.view(np.uint32))   # ... and, same kernels, the same bits


@pytest.mark.gpu
def test_two_calls_in_a_row_match_the_allocating_entry_point(pkg, synth):
    # URA21 geometry, the most blocks the fixed array takes at this block size (persistent kernel + ticket)
    C, B, nb = 21, 16384, pkg.STATIC_MAX_SPECTRUM_BYTES // (8 * 16384)
    iq1, d1 = _case(synth, C, B, nb, seed=0xA110C)
    iq2, d2 = _case(synth, C, B, nb, seed=0xA110C + 1)
    _check(pkg, iq1, C, B, d1)
    _check(pkg, iq2, C, B, d2)                                     # second call: stale spectra / peaks of the first must not show
    lag_again, _ = _static_lags(pkg, iq1, C, B)
    assert (lag_again[:, 1:] == d1[None, :]).all()


@pytest.mark.gpu
def test_geometry_changes_between_calls(pkg, synth):
    # other block sizes use other twiddle tables and the one-item-per-CTA kernels; going back and forth
    # between them and the 16384-byte case exercises everything the array holds
    for C, B, nb, seed in [(5, 2048, 40, 11), (3, 16384, 6, 12), (1, 64, 5, 13), (7, 8192, 9, 14), (3, 16384, 6, 15)]:
        iq, d = _case(synth, C, B, nb, seed)
        _check(pkg, iq, C, B, d)


@pytest.mark.gpu
def test_signed_input_and_side_stream(pkg, synth):
    """input_is_signed = 1 (the in-place XOR of cbuffer::setbufferptr already ran) on a caller-owned stream"""
    import torch
    C, B, nb = 3, 16384, 4
    iq, d = _case(synth, C, B, nb, seed=77)
    lag_b, mag_b = _batch_lags(pkg, iq, C, B)
    d_iq = torch.from_numpy(iq ^ np.uint8(0x80)).cuda()
    d_lag = torch.zeros((nb, C + 1), dtype=torch.float32, device="cuda")
    d_mag = torch.zeros((nb, C + 1), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    for _ in range(2):
        pkg.lags_static(d_iq.data_ptr(), C, B, nb, d_lag.data_ptr(), d_mag.data_ptr(), input_is_signed=True,
                        stream=side.cuda_stream)
    side.synchronize()
    np.testing.assert_array_equal(d_lag.cpu().numpy(), lag_b)
    np.testing.assert_allclose(d_mag.cpu().numpy(), mag_b, rtol=1e-4)


def test_sizes_beyond_the_fixed_array_are_rejected(pkg):
    """The limits are checked before any CUDA call, so this runs without a GPU.  0x1000 stands in for
    device pointers that are never dereferenced."""
    p = 0x1000
    with pytest.raises(pkg.CohCudaError, match="COHCUDA_STATIC_MAX_BLOCKSIZE"):
        pkg.lags_static(p, 3, 32768, 1, p, p)
    with pytest.raises(pkg.CohCudaError, match="COHCUDA_STATIC_MAX_SPECTRUM_BYTES"):
        pkg.lags_static(p, 3, 16384, pkg.STATIC_MAX_SPECTRUM_BYTES // (8 * 16384) + 1, p, p)
    with pytest.raises(pkg.CohCudaError, match="COHCUDA_STATIC_MAX_ITEMS"):
        pkg.lags_static(p, 200, 1024, 32, p, p)
    with pytest.raises(pkg.CohCudaError, match="power of two"):
        pkg.lags_static(p, 3, 3000, 1, p, p)
    with pytest.raises(pkg.CohCudaError, match="non-null"):
        pkg.lags_static(None, 3, 16384, 1, p, p)
    with pytest.raises(pkg.CohCudaError, match="16-byte aligned"):
        pkg.lags_static(p + 4, 3, 16384, 1, p, p)
