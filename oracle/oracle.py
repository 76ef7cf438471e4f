"""ctypes front-end for oracle/liboracle.so plus an independent numpy restatement.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs -- never by the product package.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle.so")
REF_LIB_PATH = os.path.join(HERE, "_ref", "libref_cdsp.so")


class Cf32(C.Structure):
    _fields_ = [("re", C.c_float), ("im", C.c_float)]


class OrcConfig(C.Structure):
    _fields_ = [("n_signal", C.c_uint32), ("blocksize", C.c_uint32), ("alpha", C.c_float),
                ("noheader", C.c_int), ("apply_shift", C.c_int), ("int_dot", C.c_int),
                ("max_lag_slots", C.c_int)]


class OrcChanState(C.Structure):
    _fields_ = [("phasecorr", Cf32), ("phasecorrprev", Cf32), ("lag", C.c_float),
                ("mag", C.c_float), ("held_lag", C.c_int32)]


def build(force: bool = False) -> None:
    """Compile liboracle.so (and _ref when /root/reference is present)."""
    if force or not os.path.exists(LIB_PATH):
        subprocess.check_call(["make", "-s", "-C", HERE, LIB_PATH])
    subprocess.call(["make", "-s", "-C", HERE, "ref"], stdout=subprocess.DEVNULL)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        vp, u32, i = C.c_void_p, C.c_uint32, C.c_int
        L.orc_packet_len.restype = C.c_size_t
        L.orc_packet_len.argtypes = [C.POINTER(OrcConfig)]
        L.orc_state_init.argtypes = [vp, u32]
        L.orc_process_batch.argtypes = [C.POINTER(OrcConfig), vp, vp, u32, vp, i, vp, i, u32,
                                        vp, vp, vp, vp, vp]
        L.orc_computelag_pair.argtypes = [vp, vp, u32, vp, vp, vp, vp]
        L.orc_fft.argtypes = [vp, vp, i, i]
        L.orc_convtosigned.argtypes = [vp, vp, i]
        L.orc_convtofloat.argtypes = [vp, vp, i]
        L.orc_convto8bit.argtypes = [vp, vp, i]
        L.orc_scalarmul.argtypes = [vp, vp, Cf32, i]
        L.orc_conj_dotproduct.argtypes = [vp, vp, i]
        L.orc_conj_dotproduct.restype = Cf32
        L.orc_magsquared.argtypes = [vp, vp, i]
        L.orc_conjugatemul.argtypes = [vp, vp, vp, i]
        L.orc_indexofmax.argtypes = [vp, i]
        L.orc_indexofmax.restype = C.c_uint32
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def make_config(n_signal, blocksize, alpha=0.5, noheader=0, apply_shift=0, int_dot=0,
                max_lag_slots=0):
    return OrcConfig(n_signal, blocksize, alpha, noheader, apply_shift, int_dot, max_lag_slots)


def new_state(n_chan):
    st = (OrcChanState * n_chan)()
    lib().orc_state_init(st, n_chan)
    return st


def process_batch(cfg, iq, state=None, lag_request=None, lag_mode=1, refnoise=None,
                  input_is_signed=0, readcnt0=0, want_cf32=False):
    """Run orc_process_batch.  iq: uint8 [n_blocks][C+1][B].  Returns dict of numpy arrays."""
    iq = np.ascontiguousarray(iq, dtype=np.uint8)
    nb, nch, B = iq.shape
    assert nch == cfg.n_signal + 1 and B == cfg.blocksize
    st = state if state is not None else new_state(nch)
    plen = lib().orc_packet_len(C.byref(cfg))
    packets = np.zeros((nb, plen), dtype=np.int8)
    lag = np.zeros((nb, nch), dtype=np.float32)
    mag = np.zeros((nb, nch), dtype=np.float32)
    phase = np.zeros((nb, nch), dtype=np.complex64)
    cf = np.zeros((nb, nch, B // 2), dtype=np.complex64) if want_cf32 else None
    lr = None if lag_request is None else np.ascontiguousarray(lag_request, dtype=np.uint8)
    rn = None if refnoise is None else np.ascontiguousarray(refnoise, dtype=np.uint8)
    lib().orc_process_batch(C.byref(cfg), st, _p(iq), nb, _p(lr), int(lag_mode), _p(rn),
                            int(input_is_signed), int(readcnt0), _p(packets), _p(lag), _p(mag),
                            _p(phase), _p(cf))
    return {"packets": packets, "lag": lag, "mag": mag, "phase": phase, "cf32": cf, "state": st}


def computelag_pair(x, r):
    x = np.ascontiguousarray(x, dtype=np.complex64)
    r = np.ascontiguousarray(r, dtype=np.complex64)
    idx = C.c_uint32(); lag = C.c_float(); mag = C.c_float(); margin = C.c_float()
    lib().orc_computelag_pair(_p(x), _p(r), len(x), C.byref(idx), C.byref(lag), C.byref(mag),
                              C.byref(margin))
    return idx.value, lag.value, mag.value, margin.value


def fft(x, sign=-1):
    x = np.ascontiguousarray(x, dtype=np.complex64)
    out = np.empty_like(x)
    lib().orc_fft(_p(out), _p(x), len(x), sign)
    return out


# ----------------------------------------------------------------- numpy restatement -----
# Independent cross-check of the C oracle (different FFT, float64 accumulation).

def np_convtofloat(u8, input_is_signed=False):
    """cdsp.cc:21-44: (u8 ^ 0x80) as int8, times float32(1/127)."""
    s8 = (u8 if input_is_signed else (u8 ^ np.uint8(0x80))).view(np.int8)
    return s8.astype(np.float32) * np.float32(1.0 / 127.0)


def np_to_complex(f):
    return f.reshape(f.shape[:-1] + (-1, 2)).view(np.complex64)[..., 0]


def np_computelag(x, r):
    """ccoherent.cc:154-239 with the crtlsdr.cc:205-218 buffer layout, in float64."""
    L = len(x)
    sx = np.zeros(2 * L, dtype=np.complex128); sx[:L] = x
    sr = np.zeros(2 * L, dtype=np.complex128); sr[L:] = r
    c = np.fft.ifft(np.fft.fft(sx) * np.conj(np.fft.fft(sr))) * (2 * L)
    m = (c.real ** 2 + c.imag ** 2)
    idx = int(np.argmax(m))
    return idx, float(idx - L), float(np.sqrt(m[idx] / L)), c


def np_pack8(v):
    """cdsp.cc:51-54 / volk_32f_s32f_convert_8i: clamp(rint(127*v))."""
    r = (v.astype(np.float32) * np.float32(127.0))
    r = np.clip(r, np.float32(-128.0), np.float32(127.0))
    return np.rint(r).astype(np.int8)


def np_process_batch_c64(iq, lag_mode=1, alpha=0.5, workers=1):
    """the whole per-block path in numpy / scipy.fft complex64 (pocketfft), vectorised
    over the channels of a block -- reference layout (crtlsdr.cc:205-218), fft -> conjugatemul -> ifft ->
    magsquared -> indexofmax (ccoherent.cc:174-192), conjugate dot + IIR (csdrdevice.cc:58-69), scalarmul
    (:80-84), convto8bit (cpacketizer.cc:158-172).  A cross-check and a timed CPU baseline, never the
    checker of record (that is the C port held bit-for-bit to the reference's golden vectors)."""
    import scipy.fft as sf
    nb, nch, B = iq.shape
    L = B // 2
    lag = np.zeros((nb, nch), np.float32); mag = np.zeros((nb, nch), np.float32)
    pk = np.empty((nb, nch, B), np.int8)
    p = np.ones(nch - 1, np.complex64)
    a, na = np.float32(alpha), np.float32(1.0 - alpha)
    buf = np.zeros((nch, 2 * L), np.complex64)
    for b in range(nb):
        f = np_to_complex(np_convtofloat(iq[b]))                      # [nch][L] complex64
        if lag_mode:
            buf[0, :L] = 0; buf[0, L:] = f[0]                         # ref in the second half (crtlsdr.cc:215-218)
            buf[1:, :L] = f[1:]; buf[1:, L:] = 0
            S = sf.fft(buf, axis=1, workers=workers)
            c = sf.ifft(S[1:] * np.conj(S[0]), axis=1, norm="forward", workers=workers)
            m = c.real * c.real + c.imag * c.imag
            idx = np.argmax(m, axis=1)
            lag[b, 1:] = idx.astype(np.float32) - np.float32(L)
            mag[b, 1:] = np.sqrt(m[np.arange(nch - 1), idx] / np.float32(L))
        corr = (f[1:] * np.conj(f[0])).sum(axis=1, dtype=np.complex64)
        q = np.conj(corr) * (np.float32(1.0) / np.abs(corr))
        p = (a * q + na * p).astype(np.complex64)
        y = f[1:] * p[:, None]
        pk[b, 0] = (iq[b, 0] ^ np.uint8(0x80)).view(np.int8)
        pk[b, 1:] = np_pack8(y.view(np.float32).reshape(nch - 1, B))
    return {"packets": pk, "lag": lag, "mag": mag, "phase": p}


def fft_us(n=16384, reps=200):
    """micro-timing printed beside the CPU arms: one n-point complex64 forward transform, 1 thread, through
    (a) the shim FFT that stands in for fftw3f inside oracle/_ref and the C port, (b) scipy's pocketfft."""
    import time
    import scipy.fft as sf
    rng = np.random.default_rng(0)
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
    y = np.empty_like(x)
    lib().orc_fft(_p(y), _p(x), n, -1); sf.fft(x)
    t0 = time.perf_counter()
    for _ in range(reps):
        lib().orc_fft(_p(y), _p(x), n, -1)
    t_shim = (time.perf_counter() - t0) / reps
    t0 = time.perf_counter()
    for _ in range(reps):
        sf.fft(x)
    t_pf = (time.perf_counter() - t0) / reps
    return {"n": n, "shim_fft_us": t_shim * 1e6, "pocketfft_us": t_pf * 1e6}


# ------------------------------------------------- the reference's own TUs (oracle/_ref) -----
# libref_cdsp.so = /root/reference's cdsp.cc, ccoherent.cc, csdrdevice.cc, crtlsdr.cc,
# cpacketizer.cc, common.cc, ccontrol.cc compiled unmodified against oracle/shim/ plus
# oracle/ref_harness.cc.  Built in the build container only; the GPU box uses the prebuilt file.

# libdropin_cuda.so = the same harness and reference TUs, except that src/cdsp.cc and
# src/ccoherent.cc are replaced by the product's host/cdsp.cc and host/ccoherent.cc over
# libcoherent_cuda.so (needs a GPU at run time; tests/test_dropin.py).
DROPIN_LIB_PATH = os.path.join(HERE, "_ref", "libdropin_cuda.so")

_ref = {}


def ref_available() -> bool:
    return os.path.exists(REF_LIB_PATH)


def dropin_available() -> bool:
    return os.path.exists(DROPIN_LIB_PATH)


def ref_lib(path=None):
    path = path or REF_LIB_PATH
    if path not in _ref:
        if not os.path.exists(path):
            raise RuntimeError(f"{path} is not built (needs /root/reference: make -C oracle ref dropin)")
        L = C.CDLL(path)
        vp, u32, i = C.c_void_p, C.c_uint32, C.c_int
        L.refh_create.restype = vp
        L.refh_create.argtypes = [u32, u32, i, i]
        L.refh_destroy.argtypes = [vp]
        L.refh_packet_len.restype = C.c_size_t
        L.refh_packet_len.argtypes = [vp]
        L.refh_sent_len.restype = C.c_size_t
        L.refh_sent_len.argtypes = [vp]
        L.refh_run_batch.argtypes = [vp, vp, u32, vp, i, vp, vp, vp, vp, vp, vp]
        L.refh_get_state.argtypes = [vp, vp, vp]
        L.refh_cdsp_convtosigned.argtypes = [vp, vp, i]
        L.refh_cdsp_convtofloat.argtypes = [vp, vp, i]
        L.refh_cdsp_convto8bit.argtypes = [vp, vp, i]
        L.refh_cdsp_scalarmul.argtypes = [vp, vp, C.c_float, C.c_float, i]
        L.refh_cdsp_conj_dotproduct.argtypes = [vp, vp, vp, i]
        L.refh_cdsp_rms.argtypes = [vp, i]
        L.refh_cdsp_rms.restype = C.c_float
        L.refh_cdsp_crestfactor.argtypes = [vp, i]
        L.refh_cdsp_crestfactor.restype = C.c_float
        L.refh_cdsp_magsquared.argtypes = [vp, vp, i]
        L.refh_cdsp_conjugatemul.argtypes = [vp, vp, vp, i]
        L.refh_cdsp_indexofmax.argtypes = [vp, i]
        L.refh_cdsp_indexofmax.restype = C.c_uint32
        if hasattr(L, "refh_live_create"):
            L.refh_live_create.restype = vp
            L.refh_live_create.argtypes = [u32, u32, u32, vp, vp, vp, C.c_uint64, i]
            L.refh_live_poll.restype = C.c_uint64
            L.refh_live_poll.argtypes = [vp, vp, vp, vp, vp, vp, vp]
            L.refh_live_kick.argtypes = [vp, u32, C.c_double]
            L.refh_live_last_packet.restype = C.c_size_t
            L.refh_live_last_packet.argtypes = [vp, vp, C.c_size_t]
            L.refh_live_destroy.argtypes = [vp]
        _ref[path] = L
    return _ref[path]


class RefCdsp:
    """The reference's own ccoherent/csdrdevice/cpacketize objects, fed block by block
    (oracle/ref_harness.cc).  One live instance per process (cpacketize is all-static).
    nfft=8 is the reference's compile-time value (src/main.cc:165): at most 7 lags per block."""

    def __init__(self, n_signal, blocksize, noheader=False, nfft=8, lib_path=None):
        self.lib = ref_lib(lib_path)
        self.h = self.lib.refh_create(n_signal, blocksize, int(noheader), nfft)
        if not self.h:
            raise RuntimeError("refh_create failed (another instance live, or bad geometry)")
        self.nch, self.B = n_signal + 1, blocksize
        self.packet_len = self.lib.refh_packet_len(self.h)
        self.sent_len = self.lib.refh_sent_len(self.h)

    def run_batch(self, iq, lag_request=None, lag_mode=1, refnoise=None):
        iq = np.ascontiguousarray(iq, dtype=np.uint8)
        nb = iq.shape[0]
        assert iq.shape[1:] == (self.nch, self.B)
        out = {"packets": np.zeros((nb, self.packet_len), dtype=np.int8),
               "lag": np.zeros((nb, self.nch), dtype=np.float32),
               "mag": np.zeros((nb, self.nch), dtype=np.float32),
               "phase": np.zeros((nb, self.nch), dtype=np.complex64),
               "measured": np.zeros((nb, self.nch), dtype=np.uint8)}
        lr = None if lag_request is None else np.ascontiguousarray(lag_request, dtype=np.uint8)
        rn = None if refnoise is None else np.ascontiguousarray(refnoise, dtype=np.uint8)
        rc = self.lib.refh_run_batch(self.h, _p(iq), nb, _p(lr), int(lag_mode), _p(rn), _p(out["packets"]),
                                     _p(out["lag"]), _p(out["mag"]), _p(out["phase"]), _p(out["measured"]))
        if rc:
            raise RuntimeError(f"refh_run_batch failed ({rc})")
        return out

    def close(self):
        if getattr(self, "h", None):
            self.lib.refh_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class RefLive:
    """The reference's whole object graph free-running against simulated dongles (live mode of
    oracle/ref_harness.cc): read threads, ccontrol threads, DSP thread, sender.  The fake
    librtlsdr's delay follows the correction ccontrol sets (oracle/shim/shim_impl.cc), so the
    loop lag -> ppm law -> resampler -> delay is closed.  With lib_path=DROPIN_LIB_PATH the DSP
    thread is the product's host/ccoherent.cc on the GPU; everything else is unchanged."""

    def __init__(self, delays, blocksize=16384, fs=1024000, gains=None, phases=None, seed=0xC0DE, nfft=8,
                 lib_path=None):
        self.lib = ref_lib(lib_path)
        n_signal = len(delays)
        self.nch, self.B = n_signal + 1, blocksize
        d0 = np.zeros(self.nch, dtype=np.float64); d0[1:] = delays
        g = np.ones(self.nch, dtype=np.float32)
        ph = np.zeros(self.nch, dtype=np.float32)
        if gains is not None: g[1:] = gains
        if phases is not None: ph[1:] = phases
        self.h = self.lib.refh_live_create(n_signal, blocksize, fs, _p(d0), _p(g), _p(ph), seed, nfft)
        if not self.h:
            raise RuntimeError("refh_live_create failed (another instance live, or bad geometry)")

    def poll(self):
        out = {"lag": np.zeros(self.nch, dtype=np.float32), "synced": np.zeros(self.nch, dtype=np.uint8),
               "delay": np.zeros(self.nch, dtype=np.float64), "corr": np.zeros(self.nch, dtype=np.float32),
               "n_corrections": np.zeros(self.nch, dtype=np.uint32), "blocks": np.zeros(self.nch, dtype=np.uint64)}
        out["packets"] = self.lib.refh_live_poll(self.h, _p(out["lag"]), _p(out["synced"]), _p(out["delay"]),
                                                 _p(out["corr"]), _p(out["n_corrections"]), _p(out["blocks"]))
        return out

    def kick(self, chan, delta):
        if self.lib.refh_live_kick(self.h, chan, float(delta)):
            raise ValueError("bad channel")

    def last_packet(self):
        n = self.lib.refh_live_last_packet(self.h, None, 0)
        buf = np.zeros(n, dtype=np.uint8)
        self.lib.refh_live_last_packet(self.h, _p(buf), n)
        return buf

    def close(self):
        if getattr(self, "h", None):
            self.lib.refh_live_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
