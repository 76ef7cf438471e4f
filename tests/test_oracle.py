"""CPU tests of the checker itself: the C restatement (oracle/liboracle.so) against
  (1) the committed golden vectors produced by the reference's own translation units
      (tests/golden/*.npz, made by tests/golden/make_golden.py from oracle/_ref),
  (2) the reference build itself when it is present (build container only),
  (3) an independent numpy/float64 restatement and known answers by construction.
None of this touches the GPU or the product library.
"""
import glob
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLD, "*.npz")))


def load_case(name):
    d = np.load(os.path.join(GOLD, name + ".npz"))
    C, B, nb, noheader, nfft, sent_len = (int(x) for x in d["meta"])
    return d, C, B, nb, noheader, nfft, sent_len


def test_golden_fixtures_present():
    assert {"four_cfg", "edge_delay", "mixed_slots", "raw_mode", "track_only", "dead_channel", "delay_steps",
            "long_block", "ura21_nfft8", "ura21_nfft22", "big65536"} <= set(CASES)


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_golden(oracle, name):
    """Bit-for-bit: packets (header, readcnt, every column), lag, mag, phase factors."""
    d, C, B, nb, noheader, nfft, _ = load_case(name)
    cfg = oracle.make_config(C, B, noheader=noheader, apply_shift=0, int_dot=0, max_lag_slots=nfft - 1)
    out = oracle.process_batch(cfg, d["iq"], lag_request=d["lag_request"], refnoise=d["refnoise"], readcnt0=8)
    np.testing.assert_array_equal(out["lag"], d["lag"])
    np.testing.assert_array_equal(out["mag"].view(np.uint32), d["mag"].view(np.uint32))
    np.testing.assert_array_equal(out["phase"].view(np.uint32), d["phase"].view(np.uint32))
    np.testing.assert_array_equal(out["packets"], d["packets"])


@pytest.mark.parametrize("name", CASES)
def test_golden_known_answers(name):
    """The reference's own output recovers the injected integer delays (F4 sign convention)."""
    d, C, B, nb, noheader, nfft, sent_len = load_case(name)
    hdr = 0 if noheader else 16 + 4 * (C + 1)
    assert d["packets"].shape == (nb, hdr + (C + 1) * B)
    assert sent_len == hdr + 2 * (C + 1) * B          # the reference publishes twice the payload (cpacketizer.cc:91-96)
    meas = d["measured"].astype(bool)
    assert meas[:, 1:].sum(axis=1).max() <= nfft - 1   # ccoherent.cc:124 slot cap
    want = np.broadcast_to(np.concatenate([[0], d["delays"]]).astype(np.float32), d["lag"].shape).copy()
    if "delay_matrix" in d.files:                      # delays that step during the run (drift / re-sync)
        want[:, 1:] = d["delay_matrix"]
    dead = (d["iq"] == 0x80).all(axis=2)               # a channel that is not streaming: every |c|^2 ties at 0
    want[dead] = -(B // 2)                             # => first index (volk_32f_index_max_32u), lag = 0 - L
    assert (d["lag"][meas] == want[meas]).all()
    if dead.any():
        b, c = np.argwhere(dead)[0]
        assert np.isnan(d["phase"][b:, c]).all()       # |corr| = 0 => NaN for good (src/csdrdevice.cc:63-67)
        assert (d["packets"][b:, hdr + c * B:hdr + (c + 1) * B] == 0).all()
    if not noheader:
        h = d["packets"][:, :16].copy().view(np.uint32)
        assert (h[:, 0] == np.arange(nb)).all() and (h[:, 1] == C + 1).all() and (h[:, 2] == B // 2).all()
        rc = d["packets"][:, 16:hdr].copy().view(np.uint32)
        assert (rc == (8 + np.arange(nb))[:, None]).all()   # asyncbufn priming buffers came first
    # column 0 is the reference channel's raw bytes with the sign bit flipped (cpacketizer.cc:151)
    col0 = d["packets"][:, hdr:hdr + B].view(np.uint8)
    np.testing.assert_array_equal(col0, d["iq"][:, 0, :] ^ 0x80)


@pytest.mark.parametrize("name", ["four_cfg", "mixed_slots", "raw_mode"])
def test_reference_build_reproduces_golden(oracle, name):
    """Only where /root/reference was compiled (oracle/_ref): the fixtures are reproducible."""
    if not oracle.ref_available():
        pytest.skip("oracle/_ref not built on this machine")
    d, C, B, nb, noheader, nfft, sent_len = load_case(name)
    with oracle.RefCdsp(C, B, noheader=bool(noheader), nfft=nfft) as R:
        out = R.run_batch(d["iq"], lag_request=d["lag_request"], refnoise=d["refnoise"])
        assert R.sent_len == sent_len
    for k in ("packets", "lag", "measured"):
        np.testing.assert_array_equal(out[k], d[k])
    np.testing.assert_array_equal(out["phase"].view(np.uint32), d["phase"].view(np.uint32))


def test_reference_primitives_vs_oracle(oracle):
    """cdsp:: statics of the reference build against the restated primitives."""
    if not oracle.ref_available():
        pytest.skip("oracle/_ref not built on this machine")
    R, L = oracle.ref_lib(), oracle.lib()
    rng = np.random.default_rng(3)
    p = lambda a: a.ctypes.data
    u8 = rng.integers(0, 256, 4096, dtype=np.uint8)
    a, b = np.empty_like(u8), np.empty_like(u8)
    R.refh_cdsp_convtosigned(p(u8), p(a), 4096); L.orc_convtosigned(p(u8), p(b), 4096)
    np.testing.assert_array_equal(a, b)
    fa, fb = np.empty(4096, np.float32), np.empty(4096, np.float32)
    R.refh_cdsp_convtofloat(p(fa), p(a), 4096); L.orc_convtofloat(p(fb), p(b), 4096)
    np.testing.assert_array_equal(fa.view(np.uint32), fb.view(np.uint32))
    x = (rng.standard_normal(2048) * 0.7).astype(np.float32)
    x[:8] = [1.5, -1.5, 0.5 / 127, 1.5 / 127, 2.5 / 127, -0.5 / 127, 1.0, -1.0079]
    qa, qb = np.empty(2048, np.int8), np.empty(2048, np.int8)
    R.refh_cdsp_convto8bit(p(qa), p(x), 1024); L.orc_convto8bit(p(qb), p(x), 2048)
    np.testing.assert_array_equal(qa, qb)
    z = (rng.standard_normal(1024) + 1j * rng.standard_normal(1024)).astype(np.complex64)
    w = (rng.standard_normal(1024) + 1j * rng.standard_normal(1024)).astype(np.complex64)
    oa, ob = np.empty_like(z), np.empty_like(z)
    R.refh_cdsp_scalarmul(p(oa), p(z), 0.6, -0.8, 1024)
    L.orc_scalarmul(p(ob), p(z), oracle.Cf32(0.6, -0.8), 1024)
    np.testing.assert_array_equal(oa.view(np.uint32), ob.view(np.uint32))
    R.refh_cdsp_conjugatemul(p(oa), p(z), p(w), 1024); L.orc_conjugatemul(p(ob), p(z), p(w), 1024)
    np.testing.assert_array_equal(oa.view(np.uint32), ob.view(np.uint32))
    res = np.empty(2, np.float32)
    R.refh_cdsp_conj_dotproduct(p(res), p(z), p(w), 1023)
    r2 = L.orc_conj_dotproduct(p(z), p(w), 1023)
    assert (res[0], res[1]) == (r2.re, r2.im)
    ma, mb = np.empty(1024, np.float32), np.empty(1024, np.float32)
    R.refh_cdsp_magsquared(p(ma), p(z), 1024); L.orc_magsquared(p(mb), p(z), 1024)
    np.testing.assert_array_equal(ma, mb)
    mb[100] = mb[700] = mb.max() + 1                     # tie -> first index
    assert R.refh_cdsp_indexofmax(p(mb), 1024) == L.orc_indexofmax(p(mb), 1024) == 100


# ---------------------------------------------------------------- independent restatement ----
def test_convtofloat_all_codes(oracle):
    """F5: (u8 ^ 0x80) * float(1/127); multiplying by the reciprocal differs from dividing."""
    u8 = np.arange(256, dtype=np.uint8)
    s8 = np.empty(256, np.uint8); f = np.empty(256, np.float32)
    oracle.lib().orc_convtosigned(u8.ctypes.data, s8.ctypes.data, 256)
    oracle.lib().orc_convtofloat(f.ctypes.data, s8.ctypes.data, 256)
    np.testing.assert_array_equal(s8.view(np.int8), (u8.astype(np.int16) - 128).astype(np.int8))
    want = oracle.np_convtofloat(u8)
    np.testing.assert_array_equal(f.view(np.uint32), want.view(np.uint32))
    assert float(np.float32(1.0 / 127.0)).hex() == "0x1.0204080000000p-7"
    div = (u8.astype(np.int16) - 128).astype(np.float32) / np.float32(127.0)
    assert (f != div).sum() == 16


def test_convto8bit_rounding(oracle):
    v = np.array([0.5, 1.5, 2.5, -0.5, -1.5, 126.5, 127.4, 127.6, 200.0, -127.5, -128.5, -300.0, 0.0],
                 dtype=np.float32) / np.float32(127.0)
    out = np.empty(len(v), np.int8)
    oracle.lib().orc_convto8bit(out.ctypes.data, v.ctypes.data, len(v))
    np.testing.assert_array_equal(out, oracle.np_pack8(v))
    rng = np.random.default_rng(11)
    r = (rng.standard_normal(1 << 16) * 0.6).astype(np.float32)
    out = np.empty(len(r), np.int8)
    oracle.lib().orc_convto8bit(out.ctypes.data, r.ctypes.data, len(r))
    np.testing.assert_array_equal(out, oracle.np_pack8(r))
    assert out.max() == 127 and out.min() == -128       # saturation exercised


@pytest.mark.parametrize("n", [2, 4, 8, 32, 64, 512, 2048, 16384, 131072])
@pytest.mark.parametrize("sign", [-1, 1])
def test_fft_vs_numpy(oracle, n, sign):
    rng = np.random.default_rng(n)
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
    got = oracle.fft(x, sign)
    want = np.fft.fft(x.astype(np.complex128)) if sign < 0 else np.fft.ifft(x.astype(np.complex128)) * n
    err = np.linalg.norm(got - want) / np.linalg.norm(want)
    assert err < 2e-6, err


@pytest.mark.parametrize("L,d", [(8192, 0), (8192, 17), (8192, -250), (8192, 4095), (8192, -4096),
                                 (1024, 1), (1024, -1), (64, 5), (32, -3)])
def test_computelag_known_answer(oracle, synth, L, d):
    iq, _ = synth.make_blocks(1, 2 * L, 2, seed=100 + L + d, delays=np.array([d]))
    x = oracle.np_to_complex(oracle.np_convtofloat(iq[1, 1]))
    r = oracle.np_to_complex(oracle.np_convtofloat(iq[1, 0]))
    idx, lag, mag, margin = oracle.computelag_pair(x, r)
    nidx, nlag, nmag, c = oracle.np_computelag(x, r)
    assert idx == nidx and lag == nlag == float(d) and idx == d + L
    assert abs(mag - nmag) / nmag < 1e-5
    assert margin < 0.5                                  # second-highest bin well below the peak


def test_shift_identity_and_phase_recovery(oracle, synth):
    """With every delay zero the software-shift extension reduces exactly to the reference
    path, and conj(p) converges to the injected carrier phase."""
    C, B, nb = 4, 4096, 30
    iq, info = synth.make_blocks(C, B, nb, seed=5, delays=np.zeros(C, dtype=int))
    a = oracle.process_batch(oracle.make_config(C, B, apply_shift=0), iq, lag_mode=1)
    b = oracle.process_batch(oracle.make_config(C, B, apply_shift=1), iq, lag_mode=1)
    for k in ("packets", "lag", "mag"):
        np.testing.assert_array_equal(a[k], b[k])
    np.testing.assert_array_equal(a["phase"].view(np.uint32), b["phase"].view(np.uint32))
    est = np.angle(np.conj(a["phase"][-1, 1:]))
    err = np.angle(np.exp(1j * (est - info["phases"])))
    assert np.abs(err).max() < 1e-2
    # integer-dot specification vs the reference's float dot: inside the north-star tolerances
    c = oracle.process_batch(oracle.make_config(C, B, int_dot=1), iq, lag_mode=1)
    assert np.abs(np.angle(c["phase"] * np.conj(a["phase"]))).max() < 1e-5
    assert (c["packets"] != a["packets"]).mean() < 1e-3


def test_software_shift_aligns_channels(oracle, synth):
    C, B, nb = 3, 2048, 16
    delays = np.array([9, -40, 100])
    iq, _ = synth.make_blocks(C, B, nb, seed=9, delays=delays)
    out = oracle.process_batch(oracle.make_config(C, B, apply_shift=1), iq, lag_mode=1, want_cf32=True)
    assert (out["lag"][:, 1:] == delays).all()
    ref = out["cf32"][-1:, 0, :]                          # last block: phase IIR has converged
    for k, d in enumerate(delays):
        y = out["cf32"][-1:, k + 1, :]
        lo, hi = max(0, -d), min(B // 2, B // 2 - d)
        assert (y[:, :lo] == 0).all() and (y[:, hi:] == 0).all()       # shifted-in samples are zero
        rho = np.vdot(ref[:, lo:hi], y[:, lo:hi]) / (np.linalg.norm(ref[:, lo:hi]) * np.linalg.norm(y[:, lo:hi]))
        assert rho.real > 0.95 and abs(rho.imag) < 0.05                # aligned and phase-corrected


def test_state_carry_equals_one_batch(oracle, synth):
    C, B, nb = 3, 1024, 20
    iq, _ = synth.make_blocks(C, B, nb, seed=21)
    cfg = oracle.make_config(C, B, apply_shift=1)
    whole = oracle.process_batch(cfg, iq, lag_mode=1)
    st = oracle.new_state(C + 1)
    first = oracle.process_batch(cfg, iq[:7], state=st, lag_mode=1)
    second = oracle.process_batch(cfg, iq[7:], state=st, lag_mode=0)     # held lags, IIR state carried
    np.testing.assert_array_equal(np.concatenate([first["phase"], second["phase"]]).view(np.uint32),
                                  whole["phase"].view(np.uint32))
    np.testing.assert_array_equal(np.concatenate([first["packets"], second["packets"]])[:, 16 + 4 * (C + 1):],
                                  whole["packets"][:, 16 + 4 * (C + 1):])


def test_refnoise_gating_holds_phase(oracle, synth):
    C, B, nb = 2, 1024, 12
    iq, _ = synth.make_blocks(C, B, nb, seed=33)
    noise = np.ones(nb, np.uint8); noise[5:9] = 0
    out = oracle.process_batch(oracle.make_config(C, B), iq, lag_mode=0, refnoise=noise)
    for b in range(5, 9):                                 # ccoherent.cc:271: no estimate, still applied
        np.testing.assert_array_equal(out["phase"][b].view(np.uint32), out["phase"][4].view(np.uint32))
    assert (out["phase"][9, 1:] != out["phase"][8, 1:]).all()
