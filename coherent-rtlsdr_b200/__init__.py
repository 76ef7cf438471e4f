"""coherent-rtlsdr_b200 -- Python face of libcoherent_cuda.so (ctypes over the C ABI).

The product is the shared library in ``lib/`` (sources in ``csrc/``, ABI in
``include/coherent_cuda.h``).  This module only loads it and marshals numpy / raw pointers; it
contains no arithmetic and has no CPU fallback: if the library is missing or no CUDA device is
usable, construction raises.

The directory name carries a hyphen (it mirrors the reference's repository name), so import it
through ``cohpkg.load()`` at the repository root.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# COHCUDA_LIB overrides the library path (kernel-variant experiments); default = the in-tree build
LIB_PATH = os.environ.get("COHCUDA_LIB") or os.path.join(HERE, "lib", "libcoherent_cuda.so")

LAG_NONE, LAG_ALL = 0, 1
SHARD_CHANNELS, SHARD_BLOCKS = 0, 1
GATHER_NCCL, GATHER_PEER_COPY, GATHER_PEER_STORE = 0, 1, 2
GATHER_BLOB_MAX = 256
TRANSPORTS = {"nccl": GATHER_NCCL, "peer_copy": GATHER_PEER_COPY, "peer_store": GATHER_PEER_STORE}

# every symbol include/coherent_cuda.h declares
ABI_SYMBOLS = [
    "cohcuda_create", "cohcuda_destroy", "cohcuda_last_error", "cohcuda_register_host",
    "cohcuda_unregister_host", "cohcuda_packet_layout", "cohcuda_submit_block",
    "cohcuda_wait_block", "cohcuda_batch", "cohcuda_sync", "cohcuda_flush", "cohcuda_covariance", "cohcuda_fraclag", "cohcuda_launch_count",
    "cohcuda_device_packets", "cohcuda_get_state", "cohcuda_set_state",
    "cohcuda_cdsp_convtosigned", "cohcuda_cdsp_convtofloat", "cohcuda_cdsp_convto8bit",
    "cohcuda_cdsp_scalarmul", "cohcuda_cdsp_conj_dotproduct", "cohcuda_cdsp_magsquared",
    "cohcuda_cdsp_conjugatemul", "cohcuda_cdsp_indexofmax", "cohcuda_cdsp_rms",
    "cohcuda_cdsp_fft", "cohcuda_profile", "cohcuda_profile_read",
    "cohcuda_gather_plan", "cohcuda_gather_init", "cohcuda_gather_export", "cohcuda_gather_attach",
    "cohcuda_gather_geometry_of", "cohcuda_gather_packets", "cohcuda_gather_fetch", "cohcuda_gather_bytes",
    "cohcuda_gather_destroy", "cohcuda_lags_static",
]
# limits of cohcuda_lags_static (include/coherent_cuda.h)
STATIC_MAX_BLOCKSIZE, STATIC_MAX_SPECTRUM_BYTES, STATIC_MAX_ITEMS = 16384, 4 << 20, 4096


class Config(C.Structure):
    _fields_ = [("n_signal", C.c_uint32), ("blocksize", C.c_uint32), ("max_lag_slots", C.c_uint32),
                ("alpha", C.c_float), ("noheader", C.c_int), ("apply_shift", C.c_int),
                ("device", C.c_int), ("max_batch_blocks", C.c_uint32), ("stream", C.c_void_p),
                ("pipe_chunk_blocks", C.c_uint32)]


class ChanState(C.Structure):
    _fields_ = [("phasecorr", C.c_float * 2), ("phasecorrprev", C.c_float * 2),
                ("lag", C.c_float), ("mag", C.c_float), ("held_lag", C.c_int32)]


class Result(C.Structure):
    _fields_ = [("lag", C.POINTER(C.c_float)), ("mag", C.POINTER(C.c_float)),
                ("phasecorr", C.POINTER(C.c_float)), ("measured", C.POINTER(C.c_uint8)),
                ("packet", C.POINTER(C.c_int8)), ("packet_len", C.c_size_t),
                ("globalseqn", C.c_uint32)]


class BatchArgs(C.Structure):
    _fields_ = [("iq", C.c_void_p), ("iq_on_device", C.c_int), ("n_blocks", C.c_uint32),
                ("lag_request", C.c_void_p), ("lag_mode", C.c_int), ("refnoise", C.c_void_p),
                ("refnoise_enabled", C.c_int), ("input_is_signed", C.c_int),
                ("readcnt", C.c_void_p), ("readcnt0", C.c_uint32), ("globalseqn0", C.c_uint32),
                ("packets", C.c_void_p), ("packets_on_device", C.c_int), ("lag", C.c_void_p),
                ("mag", C.c_void_p), ("phasecorr", C.c_void_p), ("cf32", C.c_void_p),
                ("gather", C.c_int), ("gather_skip_blocks", C.c_uint32), ("sticky_requests", C.c_int)]


class GatherConfig(C.Structure):
    _fields_ = [("rank", C.c_int), ("n_ranks", C.c_int), ("root", C.c_int), ("shard", C.c_int),
                ("transport", C.c_int), ("rank_count", C.POINTER(C.c_uint32)), ("max_blocks", C.c_uint32),
                ("chunk_blocks", C.c_uint32), ("store_ctas", C.c_uint32), ("reserve_sms", C.c_int32)]


class GatherGeometry(C.Structure):
    _fields_ = [("packet_len", C.c_size_t), ("dev_pitch", C.c_size_t), ("dev_lead", C.c_size_t),
                ("total_channels", C.c_uint32), ("total_blocks", C.c_uint32), ("first_column", C.c_uint32),
                ("first_block", C.c_uint32), ("run_offset", C.c_size_t), ("run_bytes", C.c_size_t)]


def rxx_from_sums(S, m, rows, drop_reference=True, remove_mean=True, scale=1.0 / 128.0):
    """The clients' covariance estimate from the exact sums of ``covariance_sums``:
    X = scale * columns (matlabclient/zmqsdr.c:126 scales by 1/128), reference column dropped
    (heatmap2d2.cpp:189 ``rightCols(cols-1)``), column means removed (heatmap2d2.cpp:190,
    pmusic.m:13), Rxx = X^H X / rows (heatmap2d2.cpp:199, pmusic.m:14).  sum conj(x_i - mu_i)(x_j - mu_j)
    = S_ij - conj(m_i) m_j / rows."""
    S = np.asarray(S, dtype=np.complex128); m = np.asarray(m, dtype=np.complex128)
    if remove_mean:
        S = S - np.conj(m)[..., :, None] * m[..., None, :] / rows
    if drop_reference:
        S = S[..., 1:, 1:]
    return (scale * scale / rows) * S


class CohCudaError(RuntimeError):
    pass


_lib = None


def load_library():
    """dlopen libcoherent_cuda.so and declare prototypes.  Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CohCudaError(f"{LIB_PATH} is missing: run `make` (or __graft_entry__.build()); "
                           "there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, i, u32 = C.c_void_p, C.c_int, C.c_uint32
    L.cohcuda_create.argtypes = [C.POINTER(Config), C.POINTER(vp)]
    L.cohcuda_destroy.argtypes = [vp]; L.cohcuda_destroy.restype = None
    L.cohcuda_last_error.argtypes = [vp]; L.cohcuda_last_error.restype = C.c_char_p
    L.cohcuda_register_host.argtypes = [vp, vp, C.c_size_t]
    L.cohcuda_unregister_host.argtypes = [vp, vp]
    L.cohcuda_packet_layout.argtypes = [vp, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]
    L.cohcuda_submit_block.argtypes = [vp, C.POINTER(vp), vp, vp, i, i]
    L.cohcuda_wait_block.argtypes = [vp, C.POINTER(Result)]
    L.cohcuda_batch.argtypes = [vp, C.POINTER(BatchArgs)]
    L.cohcuda_sync.argtypes = [vp]
    L.cohcuda_flush.argtypes = [vp]
    L.cohcuda_covariance.argtypes = [vp, vp, C.c_uint32, vp, vp]
    L.cohcuda_fraclag.argtypes = [vp, vp, C.c_uint32, C.c_int, vp, vp]
    L.cohcuda_launch_count.argtypes = [vp]; L.cohcuda_launch_count.restype = C.c_uint64
    L.cohcuda_device_packets.argtypes = [vp]; L.cohcuda_device_packets.restype = vp
    L.cohcuda_get_state.argtypes = [vp, vp, u32]
    L.cohcuda_set_state.argtypes = [vp, vp, u32]
    L.cohcuda_cdsp_convtosigned.argtypes = [vp, vp, vp, i]
    L.cohcuda_cdsp_convtofloat.argtypes = [vp, vp, vp, i]
    L.cohcuda_cdsp_convto8bit.argtypes = [vp, vp, vp, i]
    L.cohcuda_cdsp_scalarmul.argtypes = [vp, vp, vp, C.c_float, C.c_float, i]
    L.cohcuda_cdsp_conj_dotproduct.argtypes = [vp, vp, vp, vp, i]
    L.cohcuda_cdsp_magsquared.argtypes = [vp, vp, vp, i]
    L.cohcuda_cdsp_conjugatemul.argtypes = [vp, vp, vp, vp, i]
    L.cohcuda_cdsp_indexofmax.argtypes = [vp, vp, vp, i]
    L.cohcuda_cdsp_rms.argtypes = [vp, vp, vp, i]
    L.cohcuda_cdsp_fft.argtypes = [vp, vp, vp, i, i, i]
    L.cohcuda_profile.argtypes = [vp, i]
    L.cohcuda_profile_read.argtypes = [vp, vp, vp]
    if hasattr(L, "cohcuda_lags_static"):
        L.cohcuda_lags_static.argtypes = [vp, u32, u32, u32, i, vp, vp, vp, C.c_char_p, C.c_size_t]
    if not hasattr(L, "cohcuda_gather_plan"):        # an older build of the library (kernel A/B experiments via COHCUDA_LIB)
        _lib = L
        return L
    L.cohcuda_gather_plan.argtypes = [C.POINTER(GatherConfig), i, u32, i, u32, C.POINTER(GatherGeometry)]
    L.cohcuda_gather_init.argtypes = [vp, C.POINTER(GatherConfig)]
    L.cohcuda_gather_export.argtypes = [vp, vp, C.POINTER(C.c_size_t)]
    L.cohcuda_gather_attach.argtypes = [vp, vp, C.c_size_t]
    L.cohcuda_gather_geometry_of.argtypes = [vp, C.POINTER(GatherGeometry)]
    L.cohcuda_gather_packets.argtypes = [vp]; L.cohcuda_gather_packets.restype = vp
    L.cohcuda_gather_fetch.argtypes = [vp, vp, u32, u32]
    L.cohcuda_gather_bytes.argtypes = [vp]; L.cohcuda_gather_bytes.restype = C.c_uint64
    L.cohcuda_gather_destroy.argtypes = [vp]; L.cohcuda_gather_destroy.restype = None
    _lib = L
    return L


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, int):
        return a
    return a.ctypes.data


def lags_static(iq_dev, n_signal, blocksize, n_blocks, lag_dev, mag_dev, input_is_signed=False, stream=None):
    """cohcuda_lags_static: lag / mag of every (block, channel) without a handle and without any
    allocation inside the library (its working memory is a fixed array in the library image; see
    include/coherent_cuda.h for who needs that, the size limits and the sharing rules).  All three
    buffers are DEVICE pointers (ints) on the current device: iq [n_blocks][n_signal+1][blocksize]
    uint8, lag / mag [n_blocks][n_signal+1] float32.  Enqueues on ``stream`` (int cudaStream_t or
    None) and returns; synchronise that stream before reading the outputs."""
    lib = load_library()
    err = C.create_string_buffer(512)
    rc = lib.cohcuda_lags_static(iq_dev, n_signal, blocksize, n_blocks, int(bool(input_is_signed)),
                                 lag_dev, mag_dev, stream, err, len(err))
    if rc != 0:
        raise CohCudaError(f"cohcuda_lags_static failed ({rc}): {err.value.decode()}")


class CoherentCuda:
    """One handle = one GPU rank's block sync + correction engine (mirrors ccoherent +
    the DSP members of csdrdevice + cpacketize::write in the reference)."""

    def __init__(self, n_signal, blocksize, alpha=0.5, noheader=False, apply_shift=False,
                 max_lag_slots=0, device=0, max_batch_blocks=1, stream=None, pipe_chunk_blocks=0):
        self.lib = load_library()
        self.cfg = Config(n_signal, blocksize, max_lag_slots, alpha, int(noheader), int(apply_shift),
                          device, max_batch_blocks, stream, pipe_chunk_blocks)
        self.h = C.c_void_p()
        rc = self.lib.cohcuda_create(C.byref(self.cfg), C.byref(self.h))
        if rc != 0:
            self.h = None
            raise CohCudaError(f"cohcuda_create failed ({rc}): {self.lib.cohcuda_last_error(None).decode()}")
        self.n_signal, self.nch, self.B, self.L = n_signal, n_signal + 1, blocksize, blocksize // 2
        pl, pitch, lead = C.c_size_t(), C.c_size_t(), C.c_size_t()
        self.lib.cohcuda_packet_layout(self.h, C.byref(pl), C.byref(pitch), C.byref(lead))
        self.packet_len, self.dev_pitch, self.dev_lead = pl.value, pitch.value, lead.value
        self.hdr_len = 0 if noheader else 16 + 4 * self.nch
        self._keep = []

    def _check(self, rc, what):
        if rc != 0:
            raise CohCudaError(f"{what} failed ({rc}): {self.lib.cohcuda_last_error(self.h).decode()}")

    def close(self):
        if getattr(self, "h", None):
            self.lib.cohcuda_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- batch ------------------------------------------------------------------------------
    def batch_raw(self, **kw):
        """Thin call: keyword names are cohcuda_batch_args fields; pointers as ints/ndarrays."""
        a = BatchArgs()
        for k, v in kw.items():
            setattr(a, k, _ptr(v) if k in ("iq", "lag_request", "refnoise", "readcnt", "packets",
                                            "lag", "mag", "phasecorr", "cf32") else v)
        self._check(self.lib.cohcuda_batch(self.h, C.byref(a)), "cohcuda_batch")

    def sync(self):
        self._check(self.lib.cohcuda_sync(self.h), "cohcuda_sync")

    def flush(self):
        """Stream-ordered join of the packet read-backs into the handle's stream (no host wait)."""
        self._check(self.lib.cohcuda_flush(self.h), "cohcuda_flush")

    def batch(self, iq, lag_mode=LAG_ALL, lag_request=None, refnoise=None, refnoise_enabled=True,
              input_is_signed=False, readcnt=None, readcnt0=0, globalseqn0=0, cf32_dev=None,
              gather=False, gather_skip_blocks=0, sticky_requests=False):
        """Host-array convenience: iq uint8 [n_blocks][C+1][B] -> dict of numpy outputs.
        cf32_dev: optional DEVICE pointer (int) receiving the corrected float matrix
        [n_blocks][C+1][L] complex64."""
        iq = np.ascontiguousarray(iq, dtype=np.uint8)
        nb = iq.shape[0]
        assert iq.shape[1:] == (self.nch, self.B), iq.shape
        out = {
            "packets": None if gather else np.zeros((nb, self.packet_len), dtype=np.int8),
            "lag": np.zeros((nb, self.nch), dtype=np.float32),
            "mag": np.zeros((nb, self.nch), dtype=np.float32),
            "phase": np.zeros((nb, self.nch), dtype=np.complex64),
        }
        lr = None if lag_request is None else np.ascontiguousarray(lag_request, dtype=np.uint8)
        rn = None if refnoise is None else np.ascontiguousarray(refnoise, dtype=np.uint8)
        rc_ = None if readcnt is None else np.ascontiguousarray(readcnt, dtype=np.uint32)
        self.batch_raw(iq=iq, iq_on_device=0, n_blocks=nb, lag_request=lr, lag_mode=int(lag_mode),
                       refnoise=rn, refnoise_enabled=int(refnoise_enabled),
                       input_is_signed=int(input_is_signed), readcnt=rc_, readcnt0=readcnt0,
                       globalseqn0=globalseqn0, packets=out["packets"], packets_on_device=0,
                       lag=out["lag"], mag=out["mag"], phasecorr=out["phase"], cf32=cf32_dev,
                       gather=int(bool(gather)), gather_skip_blocks=int(gather_skip_blocks),
                       sticky_requests=int(bool(sticky_requests)))
        self.sync()
        return out

    # ---- multi-GPU gather (one handle per rank; the blob travels through the caller's plumbing) -----
    def gather_init(self, rank, n_ranks, rank_count, shard, transport=GATHER_NCCL, max_blocks=1, root=0,
                    chunk_blocks=0, store_ctas=0, reserve_sms=0):
        cnt = (C.c_uint32 * n_ranks)(*[int(v) for v in rank_count])
        cfg = GatherConfig(rank, n_ranks, root, shard, transport, cnt, max_blocks, chunk_blocks, store_ctas, reserve_sms)
        self._check(self.lib.cohcuda_gather_init(self.h, C.byref(cfg)), "cohcuda_gather_init")
        g = GatherGeometry()
        self._check(self.lib.cohcuda_gather_geometry_of(self.h, C.byref(g)), "cohcuda_gather_geometry_of")
        self.gather_geo = g
        return g

    def gather_export(self):
        buf = C.create_string_buffer(GATHER_BLOB_MAX); n = C.c_size_t()
        self._check(self.lib.cohcuda_gather_export(self.h, buf, C.byref(n)), "cohcuda_gather_export")
        return buf.raw[:n.value]

    def gather_attach(self, blob):
        self._check(self.lib.cohcuda_gather_attach(self.h, blob, len(blob)), "cohcuda_gather_attach")

    def gather_fetch(self, block0, n_blocks):
        out = np.zeros((n_blocks, self.gather_geo.packet_len), dtype=np.int8)
        self._check(self.lib.cohcuda_gather_fetch(self.h, _ptr(out), block0, n_blocks), "cohcuda_gather_fetch")
        return out

    def gather_bytes(self):
        return int(self.lib.cohcuda_gather_bytes(self.h))

    def gather_packets_ptr(self):
        return self.lib.cohcuda_gather_packets(self.h)

    # ---- downstream: sample covariance of the packet columns ------------------
    def covariance_sums(self, n_blocks, packets_dev=None):
        """Exact integer sums over the int8 packet columns of the last batch (or of device packets in
        the dev_pitch layout): S[b][i][j] = sum conj(x_i) x_j (complex, Hermitian, from int64) and
        m[b][i] = sum x_i."""
        n = self.nch
        S = np.zeros((n_blocks, n, n, 2), dtype=np.int64)
        m = np.zeros((n_blocks, n, 2), dtype=np.int32)
        self._check(self.lib.cohcuda_covariance(self.h, packets_dev, n_blocks, _ptr(S), _ptr(m)), "cohcuda_covariance")
        Sc = S[..., 0] + 1j * S[..., 1]
        iu = np.triu_indices(n, 1)
        Sc[:, iu[1], iu[0]] = np.conj(Sc[:, iu[0], iu[1]])          # lower triangle by symmetry
        return Sc, m[..., 0] + 1j * m[..., 1]

    def fraclag(self, n_blocks, iq_dev=None, input_is_signed=False, want_y3=False):
        """Fractional delay of every (block, channel) of the last batch (parabola through the exact
        integer |c|^2 at lag-1, lag, lag+1); total delay = lag + frac."""
        frac = np.zeros((n_blocks, self.nch), dtype=np.float32)
        y3 = np.zeros((n_blocks, self.nch, 3), dtype=np.float64) if want_y3 else None
        self._check(self.lib.cohcuda_fraclag(self.h, iq_dev, n_blocks, int(input_is_signed), _ptr(frac), _ptr(y3)),
                    "cohcuda_fraclag")
        return (frac, y3) if want_y3 else frac

    # ---- per-block streaming ------------------------------------------------------------------
    def submit_block(self, chans, readcnt=None, lag_request=None, refnoise_enabled=True,
                     input_is_signed=False):
        arr = (C.c_void_p * self.nch)(*[c.ctypes.data for c in chans])
        rc_ = None if readcnt is None else np.ascontiguousarray(readcnt, dtype=np.uint32)
        lr = None if lag_request is None else np.ascontiguousarray(lag_request, dtype=np.uint8)
        self._keep = [chans, rc_, lr, arr]
        self._check(self.lib.cohcuda_submit_block(self.h, arr, _ptr(rc_), _ptr(lr),
                                                  int(refnoise_enabled), int(input_is_signed)),
                    "cohcuda_submit_block")

    def wait_block(self):
        r = Result()
        self._check(self.lib.cohcuda_wait_block(self.h, C.byref(r)), "cohcuda_wait_block")
        n = self.nch
        return {
            "lag": np.ctypeslib.as_array(r.lag, (n,)).copy(),
            "mag": np.ctypeslib.as_array(r.mag, (n,)).copy(),
            "phase": np.ctypeslib.as_array(r.phasecorr, (2 * n,)).copy().view(np.complex64),
            "measured": np.ctypeslib.as_array(r.measured, (n,)).copy(),
            "packet": np.ctypeslib.as_array(r.packet, (r.packet_len,)).copy(),
            "globalseqn": r.globalseqn,
        }

    # ---- state ----------------------------------------------------------------------------------
    def get_state(self):
        st = (ChanState * self.nch)()
        self._check(self.lib.cohcuda_get_state(self.h, st, self.nch), "cohcuda_get_state")
        return st

    def set_state(self, st):
        self._check(self.lib.cohcuda_set_state(self.h, st, self.nch), "cohcuda_set_state")

    def profile(self, enable):
        self._check(self.lib.cohcuda_profile(self.h, int(enable)), "cohcuda_profile")

    def profile_read(self):
        ms = (C.c_double * 3)(); n = (C.c_uint32 * 3)()
        self._check(self.lib.cohcuda_profile_read(self.h, ms, n), "cohcuda_profile_read")
        names = ("k_ref_spectrum", "k_correlate", "k_apply")
        return {k: {"ms": ms[i], "launches": n[i]} for i, k in enumerate(names)}

    def launch_count(self):
        return int(self.lib.cohcuda_launch_count(self.h))

    def register_host(self, arr):
        self._check(self.lib.cohcuda_register_host(self.h, arr.ctypes.data, arr.nbytes), "cohcuda_register_host")

    def unregister_host(self, arr):
        self._check(self.lib.cohcuda_unregister_host(self.h, arr.ctypes.data), "cohcuda_unregister_host")
