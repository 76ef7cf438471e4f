"""GPU parity: libcoherent_cuda (through the C ABI) vs the CPU oracle on identical seeded input.

Tolerances (BASELINE.json north_star): u8->cf32 conversion and integer lag bit-exact; phase
within 1e-4 rad; corrected samples within 1e-4 relative L2.  Against the oracle's integer-dot
variant (the GPU path's own specification) packets must be byte-exact.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _cols(packets, eng):
    nb = packets.shape[0]
    return packets[:, eng.hdr_len:].reshape(nb, eng.nch, eng.B)


def _rel_l2(a, b):
    a = a.astype(np.float64); b = b.astype(np.float64)
    return np.sqrt(((a - b) ** 2).sum() / max((b ** 2).sum(), 1e-30))


@pytest.mark.parametrize("B,C,nb,shift", [(16384, 3, 6, False), (16384, 3, 6, True),
                                          (2048, 5, 40, True), (512, 2, 70, False), (64, 1, 5, True),
                                          # blocks above 8192 samples: the split transform of fft_big.cu
                                          # (131072 bytes = BASELINE config 3's 65536-sample blocks)
                                          (32768, 3, 4, False), (65536, 2, 3, True), (131072, 3, 3, True),
                                          # BASELINE config 3 at its full width: 21 channels x 65536 samples
                                          (131072, 21, 2, True)])
def test_sync_mode_matches_oracle(pkg, synth, oracle, B, C, nb, shift):
    L = B // 2
    maxd = min(250, L // 4)
    rng = np.random.default_rng(B + C)
    delays = rng.integers(-maxd, maxd + 1, C)
    delays[0] = 0
    iq, info = synth.make_blocks(C, B, nb, seed=0xC0DE + B, delays=delays)
    eng = pkg.CoherentCuda(C, B, apply_shift=shift)
    out = eng.batch(iq, lag_mode=pkg.LAG_ALL, readcnt0=8)
    # specification oracle: integer dot -> byte-exact
    cfg = oracle.make_config(C, B, apply_shift=int(shift), int_dot=1)
    ref = oracle.process_batch(cfg, iq, lag_mode=1, readcnt0=8)
    np.testing.assert_array_equal(out["lag"], ref["lag"])                       # integer lag bit-exact
    assert (out["lag"][:, 1:] == delays[None, :]).all()                         # known answer
    np.testing.assert_allclose(out["mag"], ref["mag"], rtol=1e-4)
    np.testing.assert_array_equal(out["phase"].view(np.uint32), ref["phase"].view(np.uint32))
    np.testing.assert_array_equal(out["packets"], ref["packets"])
    # reference-faithful oracle (volk float dot): tolerances of the north star
    cfg_f = oracle.make_config(C, B, apply_shift=int(shift), int_dot=0)
    ref_f = oracle.process_batch(cfg_f, iq, lag_mode=1, readcnt0=8)
    dphi = np.angle(out["phase"] * np.conj(ref_f["phase"]))
    assert np.abs(dphi).max() < 1e-4
    assert np.abs(out["packets"].astype(int) - ref_f["packets"].astype(int)).max() <= 1
    assert (out["packets"] != ref_f["packets"]).mean() < 1e-3
    eng.close()


@pytest.mark.parametrize("B,C,nb", [(16384, 21, 64), (131072, 3, 9), (256, 4, 130)])
def test_track_mode_matches_oracle(pkg, synth, oracle, B, C, nb):
    iq, info = synth.make_blocks(C, B, nb, seed=7 + B, delays=np.zeros(C, dtype=int))
    eng = pkg.CoherentCuda(C, B)
    out = eng.batch(iq, lag_mode=pkg.LAG_NONE)
    cfg = oracle.make_config(C, B, int_dot=1)
    ref = oracle.process_batch(cfg, iq, lag_mode=0)
    np.testing.assert_array_equal(out["phase"].view(np.uint32), ref["phase"].view(np.uint32))
    np.testing.assert_array_equal(out["packets"], ref["packets"])
    np.testing.assert_array_equal(out["lag"], ref["lag"])
    # injected phase comes back (conj(p) ~ exp(j*phi)) once the IIR has converged
    if nb >= 40:
        est = np.angle(np.conj(out["phase"][-1, 1:]))
        err = np.angle(np.exp(1j * (est - info["phases"])))
        assert np.abs(err).max() < 5e-3 * max(1.0, (8192 / (B // 2)) ** 0.5)
    eng.close()


# ---- against the reference's own output (tests/golden/*.npz, from oracle/_ref) ----------------
import glob
import os

_GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
_CASES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(_GOLD, "*.npz")))


@pytest.mark.parametrize("name", _CASES)
def test_cuda_path_matches_reference_golden(pkg, oracle, name):
    """Reference configuration (no software shift, nfft-1 = 7 lag slots) on the reference's own
    inputs: integer lag bit-exact, phase within 1e-4 rad, corrected float samples within 1e-4
    relative L2 (BASELINE.json north_star tolerances); header, readcnt and the ref column
    byte-exact; int8 packet bytes differ from the reference's only by isolated +-1 LSB
    rounding flips (one flip in 1e4 bytes already is 1e-4 relative L2 of an int8 signal, so
    the L2 bound is applied to the float matrix, as the north star states it)."""
    import torch
    d = np.load(os.path.join(_GOLD, name + ".npz"))
    C, B, nb, noheader, nfft, _ = (int(x) for x in d["meta"])
    eng = pkg.CoherentCuda(C, B, noheader=bool(noheader), apply_shift=False, max_lag_slots=nfft - 1)
    cf = torch.zeros((nb, C + 1, B // 2, 2), dtype=torch.float32, device="cuda")
    out = eng.batch(d["iq"], lag_request=d["lag_request"], refnoise=d["refnoise"], readcnt0=8,
                    cf32_dev=cf.data_ptr())
    # the float matrix the reference packs: the oracle, which test_oracle.py pins bit-for-bit
    # to these very fixtures (packets, lag, mag, phase)
    cfg = oracle.make_config(C, B, noheader=noheader, apply_shift=0, int_dot=0, max_lag_slots=nfft - 1)
    ref = oracle.process_batch(cfg, d["iq"], lag_request=d["lag_request"], refnoise=d["refnoise"],
                               readcnt0=8, want_cf32=True)
    np.testing.assert_array_equal(ref["packets"], d["packets"])
    got_cf = cf.cpu().numpy().view(np.complex64)[..., 0]
    np.testing.assert_array_equal(got_cf[:, 0].view(np.uint32), ref["cf32"][:, 0].view(np.uint32))  # u8->cf32 bit-exact
    # a channel without energy has |corr| = 0 and a NaN phase factor from then on (src/csdrdevice.cc:63,
    # unguarded in the reference): the NaNs must sit in the same places, everything else obeys the tolerances
    nan_ph = np.isnan(d["phase"])
    np.testing.assert_array_equal(np.isnan(out["phase"]), nan_ph)
    np.testing.assert_array_equal(np.isnan(got_cf), np.isnan(ref["cf32"]))
    fin = ~np.isnan(ref["cf32"])
    num = np.linalg.norm((got_cf[fin] - ref["cf32"][fin]).astype(np.complex128))
    assert num / np.linalg.norm(ref["cf32"][fin].astype(np.complex128)) < 1e-4
    np.testing.assert_array_equal(out["lag"], d["lag"])
    np.testing.assert_allclose(out["mag"], d["mag"], rtol=1e-4)
    assert np.abs(np.angle(out["phase"][~nan_ph] * np.conj(d["phase"][~nan_ph]))).max() < 1e-4
    np.testing.assert_allclose(np.abs(out["phase"][~nan_ph]), np.abs(d["phase"][~nan_ph]), atol=1e-5)
    hdr = eng.hdr_len
    np.testing.assert_array_equal(out["packets"][:, :hdr + B], d["packets"][:, :hdr + B])
    assert (out["packets"] != d["packets"]).mean() < 1e-3          # isolated +-1 LSB rounding flips only
    assert np.abs(out["packets"].astype(int) - d["packets"].astype(int)).max() <= 1
    eng.close()


def test_streaming_block_calls_match_batch(pkg, synth):
    """cohcuda_submit_block / wait_block (the per-block call ccoherent::threadf makes) give the
    same bytes as one cohcuda_batch over the same blocks, state carried across calls."""
    C, B, nb = 3, 4096, 9
    iq, _ = synth.make_blocks(C, B, nb, seed=77, delays=np.array([3, -20, 64]))
    req = np.zeros((nb, C + 1), np.uint8); req[0, 1:] = 1; req[5, 2] = 1
    a = pkg.CoherentCuda(C, B, apply_shift=True)
    whole = a.batch(iq, lag_request=req, readcnt0=0)
    a.close()
    s = pkg.CoherentCuda(C, B, apply_shift=True)
    for b in range(nb):
        s.submit_block([iq[b, c] for c in range(C + 1)], readcnt=np.full(C + 1, b, np.uint32), lag_request=req[b])
        r = s.wait_block()
        np.testing.assert_array_equal(r["packet"], whole["packets"][b])
        np.testing.assert_array_equal(r["lag"], whole["lag"][b])
        np.testing.assert_array_equal(r["phase"].view(np.uint32), whole["phase"][b].view(np.uint32))
        assert r["globalseqn"] == b and (r["measured"][1:] == req[b, 1:]).all()
    st = s.get_state()
    assert [st[c].held_lag for c in range(1, C + 1)] == [3, -20, 64]
    s.close()


def test_cdsp_primitives_match_oracle(pkg, oracle):
    """The ten cdsp:: statics as single kernels through the C ABI (include/cdsp.h:36-71)."""
    import ctypes as C_
    eng = pkg.CoherentCuda(1, 64)
    L, lib, h = oracle.lib(), eng.lib, eng.h
    rng = np.random.default_rng(5)
    p = lambda a: a.ctypes.data
    u8 = rng.integers(0, 256, 8192, dtype=np.uint8); s8g = np.empty_like(u8); s8o = np.empty_like(u8)
    assert lib.cohcuda_cdsp_convtosigned(h, p(u8), p(s8g), 8192) == 0
    L.orc_convtosigned(p(u8), p(s8o), 8192); np.testing.assert_array_equal(s8g, s8o)
    fg, fo = np.empty(8192, np.float32), np.empty(8192, np.float32)
    assert lib.cohcuda_cdsp_convtofloat(h, p(fg), p(s8g), 8192) == 0
    L.orc_convtofloat(p(fo), p(s8o), 8192); np.testing.assert_array_equal(fg.view(np.uint32), fo.view(np.uint32))
    x = (rng.standard_normal(8192) * 0.7).astype(np.float32)
    qg, qo = np.empty(8192, np.int8), np.empty(8192, np.int8)
    assert lib.cohcuda_cdsp_convto8bit(h, p(qg), p(x), 4096) == 0
    L.orc_convto8bit(p(qo), p(x), 8192); np.testing.assert_array_equal(qg, qo)
    z = (rng.standard_normal(4096) + 1j * rng.standard_normal(4096)).astype(np.complex64)
    w = (rng.standard_normal(4096) + 1j * rng.standard_normal(4096)).astype(np.complex64)
    og, oo = np.empty_like(z), np.empty_like(z)
    assert lib.cohcuda_cdsp_scalarmul(h, p(og), p(z), C_.c_float(0.6), C_.c_float(-0.8), 4096) == 0
    L.orc_scalarmul(p(oo), p(z), oracle.Cf32(0.6, -0.8), 4096)
    np.testing.assert_array_equal(og.view(np.uint32), oo.view(np.uint32))
    assert lib.cohcuda_cdsp_conjugatemul(h, p(og), p(z), p(w), 4096) == 0
    L.orc_conjugatemul(p(oo), p(z), p(w), 4096)
    np.testing.assert_allclose(og, oo, rtol=1e-6, atol=1e-6)
    mg, mo = np.empty(4096, np.float32), np.empty(4096, np.float32)
    assert lib.cohcuda_cdsp_magsquared(h, p(mg), p(z), 4096) == 0
    L.orc_magsquared(p(mo), p(z), 4096); np.testing.assert_allclose(mg, mo, rtol=1e-6)
    mo[100] = mo[700] = mo.max() + 1
    idx = np.zeros(1, np.uint32)
    assert lib.cohcuda_cdsp_indexofmax(h, p(idx), p(mo), 4096) == 0 and idx[0] == 100
    res = np.empty(2, np.float32)
    assert lib.cohcuda_cdsp_conj_dotproduct(h, p(res), p(z), p(w), 4096) == 0
    want = np.vdot(w.astype(np.complex128), z.astype(np.complex128))      # sum z * conj(w)
    assert abs(complex(res[0], res[1]) - want) / abs(want) < 1e-5
    r = np.empty(1, np.float32)
    assert lib.cohcuda_cdsp_rms(h, p(r), p(x), 8192) == 0
    assert abs(r[0] - np.sqrt((x.astype(np.float64) ** 2).mean())) < 1e-5
    for n in (32, 256, 16384):
        xi = (rng.standard_normal((3, n)) + 1j * rng.standard_normal((3, n))).astype(np.complex64)
        for sign in (-1, 1):
            yo = np.empty_like(xi)
            assert lib.cohcuda_cdsp_fft(h, p(yo), p(xi), n, 3, sign) == 0
            want = np.fft.fft(xi.astype(np.complex128), axis=1) if sign < 0 else np.fft.ifft(xi.astype(np.complex128), axis=1) * n
            assert np.linalg.norm(yo - want) / np.linalg.norm(want) < 2e-6
    eng.close()


@pytest.mark.parametrize("alpha", [0.3, 0.9, 1.0])
def test_iir_weight_other_than_reference_default(pkg, synth, oracle, alpha):
    """alpha = 0.5 (src/csdrdevice.cc:60) makes every IIR product exact; other weights exercise
    the individually rounded multiply and add of csdrdevice.cc:67 (no fused multiply-add) and the
    look-back depth that follows from alpha."""
    C, B, nb = 3, 2048, 150
    iq, _ = synth.make_blocks(C, B, nb, seed=99, delays=np.array([0, 5, -9]))
    noise = np.ones(nb, np.uint8); noise[40:47] = 0
    eng = pkg.CoherentCuda(C, B, alpha=alpha, apply_shift=True)
    out = eng.batch(iq, lag_mode=pkg.LAG_ALL, refnoise=noise)
    cfg = oracle.make_config(C, B, alpha=alpha, apply_shift=1, int_dot=1)
    ref = oracle.process_batch(cfg, iq, lag_mode=1, refnoise=noise)
    np.testing.assert_array_equal(out["phase"].view(np.uint32), ref["phase"].view(np.uint32))
    np.testing.assert_array_equal(out["packets"], ref["packets"])
    eng.close()


def test_exact_tie_inside_the_fused_correlation_kernel(pkg, synth, oracle):
    """A signal channel with no energy (all-0x80 rows) makes every one of the 2L correlation values an
    exact 0.0 tie: the fused kernel's per-thread scan, warp shuffles and cross-warp step must return the
    FIRST index like volk_32f_index_max_32u (src/cdsp.cc:135-139) => lag = 0 - L, mag = 0; the channel's
    phase factor is NaN from then on (src/csdrdevice.cc:63) and its packet bytes 0."""
    C, B, nb = 3, 16384, 4
    iq, _ = synth.make_blocks(C, B, nb, seed=0xC0DE + 5, delays=np.array([2, -7, 30]))
    iq[1:3, 2, :] = 0x80
    eng = pkg.CoherentCuda(C, B, apply_shift=False)
    out = eng.batch(iq, lag_mode=pkg.LAG_ALL)
    eng.close()
    ref = oracle.process_batch(oracle.make_config(C, B, apply_shift=0, int_dot=1), iq, lag_mode=1)
    assert (out["lag"][1:3, 2] == -(B // 2)).all() and (out["mag"][1:3, 2] == 0).all()
    np.testing.assert_array_equal(out["lag"], ref["lag"])
    assert np.isnan(out["phase"][1:, 2]).all() and not np.isnan(out["phase"][:, [0, 1, 3]]).any()
    np.testing.assert_array_equal(np.isnan(out["phase"]), np.isnan(ref["phase"]))
    np.testing.assert_array_equal(out["packets"], ref["packets"])


def _drift_case(synth, C, B, nb, seed, steps, late):
    """BASELINE config 5's scenario: delays step at known blocks, the channel loses `synced` and requests
    its lag again (src/console.cc:285-289 -> include/csdrdevice.h:129-151,173-180), some of them late."""
    rng = np.random.default_rng(seed)
    delays = rng.integers(-40, 41, C)
    iq, info = synth.make_blocks(C, B, nb, seed=seed, delays=delays, delay_steps=steps)
    req = np.zeros((nb, C + 1), np.uint8); req[0, 1:] = 1
    for (b0, ch, _) in steps:
        req[min(b0 + late.get((b0, ch), 0), nb - 1), ch] = 1
    noise = np.ones(nb, np.uint8); noise[nb // 2: nb // 2 + 5] = 0
    return iq, info, req, noise


def test_drift_resync_delay_steps_match_oracle(pkg, synth, oracle):
    """Held lags are REPLACED mid-batch: after a step the old lag keeps being applied until the channel's
    request is served (a few blocks late for some), then the new one -- lag, phase and packet bytes
    equal the oracle's at every block, and the measured lags are the injected delays."""
    C, B, nb = 6, 4096, 150
    steps = [(20, 1, 17), (20, 4, -33), (57, 2, 5), (58, 2, -5), (101, 6, 120), (140, 1, 0)]
    late = {(20, 4): 3, (101, 6): 7}
    iq, info, req, noise = _drift_case(synth, C, B, nb, 0xD81F, steps, late)
    eng = pkg.CoherentCuda(C, B, apply_shift=True)
    out = eng.batch(iq, lag_request=req, refnoise=noise, readcnt0=8)
    eng.close()
    ref = oracle.process_batch(oracle.make_config(C, B, apply_shift=1, int_dot=1), iq, lag_request=req, refnoise=noise, readcnt0=8)
    np.testing.assert_array_equal(out["lag"], ref["lag"])
    np.testing.assert_array_equal(out["phase"].view(np.uint32), ref["phase"].view(np.uint32))
    np.testing.assert_array_equal(out["packets"], ref["packets"])
    meas = req[:, 1:].astype(bool)
    assert (out["lag"][:, 1:][meas] == info["delays"][meas]).all()       # every served request returns the CURRENT delay
    assert out["lag"][23, 4] == -33 and out["lag"][22, 4] != -33          # the late request: old lag held for 3 blocks


def test_sticky_requests_rotate_through_the_slot_cap(pkg, synth, oracle):
    """sticky_requests = 1: a request stays pending until it gets one of the nfft-1 correlation slots, as
    csdrdevice::lagrequested does (src/ccoherent.cc:124,266): 9 channels asking once are served 7 + 2 over
    two blocks.  Equal to the oracle fed with the requests as they end up being served."""
    C, B, nb, slots = 9, 1024, 6, 7
    iq, _ = synth.make_blocks(C, B, nb, seed=33, delays=np.arange(C) * 3 - 10)
    req = np.zeros((nb, C + 1), np.uint8); req[0, 1:] = 1; req[3, 2] = 1
    served = np.zeros_like(req); served[0, 1:8] = 1; served[1, 8:] = 1; served[3, 2] = 1
    eng = pkg.CoherentCuda(C, B, apply_shift=True, max_lag_slots=slots)
    out = eng.batch(iq, lag_request=req, sticky_requests=True)
    eng.close()
    ref = oracle.process_batch(oracle.make_config(C, B, apply_shift=1, int_dot=1, max_lag_slots=slots), iq, lag_request=served)
    np.testing.assert_array_equal(out["lag"], ref["lag"])
    np.testing.assert_array_equal(out["packets"], ref["packets"])


def test_failed_host_registration_does_not_poison_the_next_launch(pkg, synth):
    """cohcuda_register_host is best effort for the drop-in (librtlsdr's buffers may refuse pinning):
    a failed cudaHostRegister must not leave the runtime's last-error set for the next kernel launch."""
    C, B = 2, 2048
    iq, _ = synth.make_blocks(C, B, 2, seed=9, delays=np.array([1, -1]))
    eng = pkg.CoherentCuda(C, B)
    buf = np.ascontiguousarray(iq[0])
    eng.register_host(buf)
    with pytest.raises(pkg.CohCudaError):
        eng.register_host(buf)                       # already registered -> cudaErrorHostMemoryAlreadyRegistered
    eng.submit_block([buf[c] for c in range(C + 1)], lag_request=np.array([0, 1, 1], np.uint8))
    r = eng.wait_block()
    assert (r["lag"][1:] == [1, -1]).all()
    eng.unregister_host(buf)
    eng.close()


def test_back_to_back_host_batches_overlap_safely(pkg, synth):
    """Host-buffer batches alternate between two device buffer sets so that one batch's input copies
    overlap the previous batch's packet read-back.  Three batches issued without any wait in between
    (then one cohcuda_sync) must give the bytes of a single batch over the same blocks."""
    C, B, nb = 3, 16384, 60
    iq, _ = synth.make_blocks(C, B, nb, seed=303, delays=np.array([5, -40, 0]))
    ref_eng = pkg.CoherentCuda(C, B, apply_shift=True)
    whole = ref_eng.batch(iq, lag_mode=pkg.LAG_ALL, readcnt0=8)
    ref_eng.close()
    eng = pkg.CoherentCuda(C, B, apply_shift=True)
    outs = []
    for rep in range(2):                                  # second round reuses both buffer sets
        outs = []
        for a, b in ((0, 20), (20, 40), (40, 60)):
            o = {"packets": np.zeros((b - a, eng.packet_len), np.int8), "lag": np.zeros((b - a, C + 1), np.float32),
                 "part": np.ascontiguousarray(iq[a:b])}
            eng.batch_raw(iq=o["part"], iq_on_device=0, n_blocks=b - a, lag_mode=pkg.LAG_ALL, refnoise_enabled=1,
                          readcnt0=8 + a, globalseqn0=a, packets=o["packets"], packets_on_device=0, lag=o["lag"])
            outs.append(o)
        eng.sync()
        if rep == 0:
            eng.set_state(pkg.CoherentCuda(C, B).get_state())     # fresh channel state for the second round
    eng.close()
    got = np.concatenate([o["packets"] for o in outs])
    np.testing.assert_array_equal(got, whole["packets"])
    np.testing.assert_array_equal(np.concatenate([o["lag"] for o in outs]), whole["lag"])
