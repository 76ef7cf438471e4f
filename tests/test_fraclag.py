"""Fractional delay: the parabolic interpolation ccoherent::computelag prepares and
leaves disabled (src/ccoherent.cc:206-218), enabled with the curvature sign corrected and fed with
correlation values recomputed exactly in integers.  Oracle: the same three sums in numpy int64."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def exact_corr(xrow_u8, rrow_u8, m):
    """c[m] = sum_n x[n+m] conj(r[n]) over the overlap, exact (int64)."""
    L = xrow_u8.size // 2
    xs = (xrow_u8 ^ 0x80).view(np.int8).astype(np.int64).reshape(L, 2)
    rs = (rrow_u8 ^ 0x80).view(np.int8).astype(np.int64).reshape(L, 2)
    n0, n1 = max(0, -m), min(L, L - m)
    if n1 <= n0:
        return 0, 0
    xr, xi = xs[n0 + m:n1 + m, 0], xs[n0 + m:n1 + m, 1]
    rr, ri = rs[n0:n1, 0], rs[n0:n1, 1]
    return int((xr * rr + xi * ri).sum()), int((xi * rr - xr * ri).sum())


def parabola(y, idx, L):
    if not (1 < idx < 2 * L - 1):
        return 0.0
    a = y[0] - 2.0 * y[1] + y[2]
    return 0.0 if a == 0.0 else 0.5 * (y[0] - y[2]) / a


def bandlimited_blocks(C, B, nb, delays, seed, bw=0.25):
    """Low-pass complex noise (so that the correlation peak is wider than one sample), channels
    delayed by FRACTIONAL amounts with an FFT phase ramp on the continuous stream, then quantised
    like an RTL-SDR and cut into blocks."""
    L = B // 2
    n = nb * L
    rng = np.random.default_rng(seed)
    f = np.fft.fftfreq(n)
    spec = (rng.standard_normal(n) + 1j * rng.standard_normal(n)) * (np.abs(f) < bw / 2)
    r = np.fft.ifft(spec)
    r *= 40.0 / r.real.std()
    iq = np.zeros((nb, C + 1, B), dtype=np.uint8)

    def quant(v):
        q = np.clip(np.rint(np.stack([v.real, v.imag], -1) + 127.5), 0, 255).astype(np.uint8)
        return q.reshape(nb, B)
    iq[:, 0] = quant(r)
    R = np.fft.fft(r)
    for k, d in enumerate(delays):
        xk = np.fft.ifft(R * np.exp(-2j * np.pi * f * d)) * np.exp(1j * (0.4 + k))
        xk = xk + 2.0 * (rng.standard_normal(n) + 1j * rng.standard_normal(n))
        iq[:, k + 1] = quant(xk)
    return iq


def test_fraclag_matches_exact_integer_oracle(pkg, synth):
    C, B, nb = 4, 4096, 5
    L = B // 2
    delays = np.array([0, 17, -250, 1023])
    iq, _ = synth.make_blocks(C, B, nb, seed=5, delays=delays)
    eng = pkg.CoherentCuda(C, B)
    out = eng.batch(iq, lag_mode=pkg.LAG_ALL)
    frac, y3 = eng.fraclag(nb, want_y3=True)
    eng.close()
    assert (out["lag"][:, 1:] == delays[None, :]).all()
    for b in range(nb):
        assert frac[b, 0] == 0
        for c in range(1, C + 1):
            lag = int(out["lag"][b, c])
            y = []
            for m in (lag - 1, lag, lag + 1):
                re, im = exact_corr(iq[b, c], iq[b, 0], m)
                y.append(float(re) * float(re) + float(im) * float(im))
            np.testing.assert_allclose(y3[b, c], y, rtol=1e-14)
            np.testing.assert_allclose(frac[b, c], parabola(y, lag + L, L), rtol=1e-5, atol=1e-7)
            assert abs(frac[b, c]) < 0.05          # white noise, integer delay: a one-sample peak, no fraction


def test_fraclag_tracks_injected_fractional_delays(pkg):
    C, B, nb = 5, 16384, 3
    fr = np.array([-0.4, -0.2, 0.0, 0.2, 0.4])
    delays = np.array([12, -7, 3, 40, -90]) + fr
    iq = bandlimited_blocks(C, B, nb, delays, seed=9)
    eng = pkg.CoherentCuda(C, B)
    out = eng.batch(iq, lag_mode=pkg.LAG_ALL)
    frac = eng.fraclag(nb)
    eng.close()
    est = out["lag"][:, 1:] + frac[:, 1:]
    # the integer peak is the nearest sample; integer + fraction recovers the injected delay.  The
    # three-point parabola on a sinc^2-shaped peak is biased towards the sample, hence the loose bound.
    assert (out["lag"][:, 1:] == np.rint(delays)[None, :]).all()
    assert np.abs(est - delays[None, :]).max() < 0.12
    assert (np.sign(frac[:, 1:][:, fr != 0]) == np.sign(fr[fr != 0])[None, :]).all()
    mean_frac = frac[:, 1:].mean(axis=0)
    assert (np.diff(mean_frac) > 0).all()       # monotonic in the injected fraction
