"""Drop-in proof: the reference's object graph with src/cdsp.cc and src/ccoherent.cc swapped for
the product's host/cdsp.cc and host/ccoherent.cc (over libcoherent_cuda.so).

oracle/_ref/libdropin_cuda.so is built by oracle/Makefile from the reference's unmodified
csdrdevice.cc / crtlsdr.cc / cpacketizer.cc / common.cc / ccontrol.cc + the two replacement TUs +
the same hardware-free harness that produced tests/golden/*.npz from the unmodified reference.
Blocks enter through csdrdevice::swapbuffer (what librtlsdr's callback calls) and leave through
cpacketize::send(), exactly as in the live program; only the DSP in between runs on the GPU.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

_GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["four_cfg", "edge_delay", "mixed_slots", "raw_mode", "track_only", "ura21_nfft8", "ura21_nfft22"]


@pytest.fixture(scope="module")
def dropin(oracle):
    if not oracle.dropin_available():
        pytest.fail("oracle/_ref/libdropin_cuda.so is missing: run __graft_entry__.build() where /root/reference exists")
    return oracle.DROPIN_LIB_PATH


@pytest.mark.parametrize("name", CASES)
def test_dropin_tus_reproduce_reference_packets(oracle, dropin, name):
    g = np.load(os.path.join(_GOLD, name + ".npz"))
    C, B, nb, noheader, nfft, _ = (int(v) for v in g["meta"])
    with oracle.RefCdsp(C, B, noheader=bool(noheader), nfft=nfft, lib_path=dropin) as R:
        out = R.run_batch(g["iq"], lag_request=g["lag_request"], refnoise=g["refnoise"])
    np.testing.assert_array_equal(out["measured"], g["measured"])       # same set_lag() calls
    np.testing.assert_array_equal(out["lag"], g["lag"])                 # integer lag bit-exact
    np.testing.assert_allclose(out["mag"], g["mag"], rtol=1e-4)
    dphi = np.angle(out["phase"] * np.conj(g["phase"]))
    assert np.abs(dphi).max() < 1e-4                                    # phase within 1e-4 rad
    hdr = 0 if noheader else 16 + 4 * (C + 1)
    np.testing.assert_array_equal(out["packets"][:, :hdr], g["packets"][:, :hdr])   # hdr0 + readcnt
    a, b = out["packets"][:, hdr:].astype(int), g["packets"][:, hdr:].astype(int)
    np.testing.assert_array_equal(a[:, :B], b[:, :B])                   # reference column: raw signed bytes
    assert np.abs(a - b).max() <= 1                                     # isolated +-1 LSB (exact vs float dot)
    assert (a != b).mean() < 1e-3


def test_inplace_convtosigned_keeps_its_contract_for_cdsp_readers(oracle, dropin):
    """cbuffer::setbufferptr converts the USB buffer in place (include/common.h:121).  The drop-in defers
    that XOR to the kernels, so the bytes stay offset-binary -- but whoever reads the buffer through cdsp::
    (console `phase`: csdrdevice::convtofloat -> cdsp::convtofloat, src/console.cc:295-299) must still get
    the signed samples the reference would have produced."""
    import ctypes as Ct
    D = oracle.ref_lib(dropin)
    D.cohcuda_host_xor_pending.argtypes = [Ct.c_void_p, Ct.c_size_t]
    rng = np.random.default_rng(9)
    D.cohcuda_host_xor_forget.argtypes = [Ct.c_void_p]
    p = lambda a: a.ctypes.data_as(Ct.c_void_p)
    for n in (16384, 4096 + 8):
        ring = rng.integers(0, 256, n, dtype=np.uint8)
        raw = ring.copy()
        want = (raw ^ 0x80).view(np.int8).astype(np.float32) * np.float32(1.0 / 127.0)
        assert D.cohcuda_host_xor_pending(p(ring), n) == 0
        D.refh_cdsp_convtosigned(p(ring), p(ring), n)                      # in place: deferred, recorded
        assert D.cohcuda_host_xor_pending(p(ring), n) == 1
        f = np.zeros(n, np.float32); D.refh_cdsp_convtofloat(p(f), p(ring), n)
        np.testing.assert_array_equal(f, want)                            # bit-exact, as src/cdsp.cc:41-44 on signed bytes
        half = np.zeros(n // 2 - 3, np.float32)                           # a sub-range (odd length) of the same buffer
        D.refh_cdsp_convtofloat(p(half), Ct.c_void_p(ring.ctypes.data + 5), half.size)
        np.testing.assert_array_equal(half, want[5:5 + half.size])
        ring[:] = rng.integers(0, 256, n, dtype=np.uint8)                  # next USB transfer lands in the same buffer
        D.refh_cdsp_convtosigned(p(ring), p(ring), n)
        f2 = np.zeros(n, np.float32); D.refh_cdsp_convtofloat(p(f2), p(ring), n)
        np.testing.assert_array_equal(f2, (ring ^ 0x80).view(np.int8).astype(np.float32) * np.float32(1.0 / 127.0))
        # a buffer that was never converted in place is taken as signed, as before
        other = rng.integers(0, 256, n, dtype=np.uint8)
        f3 = np.zeros(n, np.float32); D.refh_cdsp_convtofloat(p(f3), p(other), n)
        np.testing.assert_array_equal(f3, other.view(np.int8).astype(np.float32) * np.float32(1.0 / 127.0))
        D.cohcuda_host_xor_forget(p(ring))                                # the buffer is released: its record goes with it
        assert D.cohcuda_host_xor_pending(p(ring), n) == 0


def test_dropin_cdsp_statics_match_reference(oracle, dropin):
    """The 17 cdsp:: statics of host/cdsp.cc against the reference's own src/cdsp.cc."""
    import ctypes as Ct
    D = oracle.ref_lib(dropin)
    Rf = oracle.ref_lib() if oracle.ref_available() else None
    rng = np.random.default_rng(5)
    n = 4096
    u8 = rng.integers(0, 256, n, dtype=np.uint8)
    p = lambda a: a.ctypes.data_as(Ct.c_void_p)
    s8 = np.zeros(n, np.uint8); D.refh_cdsp_convtosigned(p(u8), p(s8), n)
    np.testing.assert_array_equal(s8, u8 ^ 0x80)
    f = np.zeros(n, np.float32); D.refh_cdsp_convtofloat(p(f), p(s8), n)
    np.testing.assert_array_equal(f, s8.view(np.int8).astype(np.float32) * np.float32(1.0 / 127.0))
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
    y = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
    o = np.zeros(n, np.complex64); D.refh_cdsp_conjugatemul(p(o), p(x), p(y), n)
    np.testing.assert_allclose(o, x * np.conj(y), rtol=1e-5, atol=1e-6)
    m = np.zeros(n, np.float32); D.refh_cdsp_magsquared(p(m), p(x), n)
    np.testing.assert_allclose(m, x.real ** 2 + x.imag ** 2, rtol=1e-6)
    assert D.refh_cdsp_indexofmax(p(m), n) == int(np.argmax(m))
    r2 = np.zeros(2, np.float32); D.refh_cdsp_conj_dotproduct(p(r2), p(x), p(y), n)
    want = np.sum(x.astype(np.complex128) * np.conj(y.astype(np.complex128)))
    assert abs(complex(r2[0], r2[1]) - want) < 1e-3 * abs(want) + 1e-2
    q = np.zeros(2 * n, np.int8); D.refh_cdsp_convto8bit(p(q), p(x.copy()), n)
    if Rf is not None:
        q2 = np.zeros(2 * n, np.int8); Rf.refh_cdsp_convto8bit(p(q2), p(x.copy()), n)
        np.testing.assert_array_equal(q, q2)
        o1 = np.zeros(n, np.complex64); o2 = np.zeros(n, np.complex64)
        D.refh_cdsp_scalarmul(p(o1), p(x), Ct.c_float(0.6), Ct.c_float(-0.8), n)
        Rf.refh_cdsp_scalarmul(p(o2), p(x), Ct.c_float(0.6), Ct.c_float(-0.8), n)
        np.testing.assert_array_equal(o1.view(np.uint32), o2.view(np.uint32))
        fl = np.abs(rng.standard_normal(n)).astype(np.float32)
        assert abs(D.refh_cdsp_rms(p(fl), n) - Rf.refh_cdsp_rms(p(fl), n)) < 1e-5
        assert abs(D.refh_cdsp_crestfactor(p(fl), n) - Rf.refh_cdsp_crestfactor(p(fl), n)) < 1e-4
