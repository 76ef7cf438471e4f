"""Seeded synthetic IQ for the block sync + correction path.

A continuous complex-Gaussian reference-noise stream r[t] (sigma = 40 LSB per rail) is shared
by every channel; signal channel k sees  x_k[t] = g_k * exp(j*phi_k) * r[t - d_k] + w_k[t]
(w_k Gaussian, sigma = 4 LSB), is quantised to offset-binary u8 exactly like an RTL-SDR
(``u8 = clip(rint(v + 127.5), 0, 255)``) and cut into blocks of L complex samples.  Delays
draw samples from the neighbouring block, as hardware would.  With this sign convention the
reference's ``lag = idx - L`` (src/ccoherent.cc:232) comes back as ``+d_k``.

Layout returned: ``iq[n_blocks][C+1][B]`` uint8, channel 0 = reference, B = 2L bytes
(interleaved I,Q) -- one row per librtlsdr buffer (src/crtlsdr.cc:54).
"""
from __future__ import annotations

import numpy as np

REF_SIGMA = 40.0
NOISE_SIGMA = 4.0


def make_params(n_signal: int, seed: int, max_delay: int = 64):
    """Per-channel gain, phase and integer delay (g~U[0.7,1], phi~U(-pi,pi))."""
    rng = np.random.Generator(np.random.PCG64(seed))
    gains = rng.uniform(0.7, 1.0, n_signal)
    phases = rng.uniform(-np.pi, np.pi, n_signal)
    delays = rng.integers(-max_delay, max_delay + 1, n_signal)
    return gains, phases, delays


def make_blocks(n_signal: int, blocksize: int, n_blocks: int, seed: int,
                delays=None, phases=None, gains=None, max_delay: int = 64,
                delay_steps=None, noise_sigma: float = NOISE_SIGMA):
    """Return (iq, info).  ``delay_steps``: optional list of (block_index, channel, new_delay)
    applied from that block on (drift / re-sync scenario, config 5)."""
    L = blocksize // 2
    g0, p0, d0 = make_params(n_signal, seed, max_delay)
    gains = g0 if gains is None else np.asarray(gains, dtype=np.float64)
    phases = p0 if phases is None else np.asarray(phases, dtype=np.float64)
    delays = d0 if delays is None else np.asarray(delays, dtype=np.int64)
    rng = np.random.Generator(np.random.PCG64(seed + 0x9E3779B9))
    dmat = np.tile(np.asarray(delays, dtype=np.int64), (n_blocks, 1))
    for (b0, ch, nd) in (delay_steps or []):
        dmat[b0:, ch - 1] = nd
    pad = int(np.abs(dmat).max()) + 1 if n_signal else 1
    T = n_blocks * L + 2 * pad
    r = (rng.standard_normal(T) + 1j * rng.standard_normal(T)) * REF_SIGMA

    def quant(v):
        out = np.empty(v.shape + (2,), dtype=np.uint8)
        out[..., 0] = np.clip(np.rint(v.real + 127.5), 0, 255).astype(np.uint8)
        out[..., 1] = np.clip(np.rint(v.imag + 127.5), 0, 255).astype(np.uint8)
        return out.reshape(v.shape[:-1] + (2 * v.shape[-1],))

    iq = np.empty((n_blocks, n_signal + 1, blocksize), dtype=np.uint8)
    iq[:, 0, :] = quant(r[pad:pad + n_blocks * L].reshape(n_blocks, L))
    for k in range(n_signal):
        w = (rng.standard_normal(n_blocks * L) + 1j * rng.standard_normal(n_blocks * L)) * noise_sigma
        rot = gains[k] * np.exp(1j * phases[k])
        for b in range(n_blocks):
            d = int(dmat[b, k])
            s = pad + b * L - d
            iq[b, k + 1, :] = quant(rot * r[s:s + L] + w[b * L:(b + 1) * L])
    info = {"gains": gains, "phases": phases, "delays": dmat, "L": L}
    return iq, info
