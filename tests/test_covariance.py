"""Downstream of the path: the sample covariance the reference's clients form from
a packet -- Rxx = X^H X / rows (beamformclient/heatmap2d2.cpp:189-199, matlabclient/functions/
pmusic.m:12-14).  The GPU forms the sums exactly in integers over the int8 packet columns; the check is
numpy int64 on the same packets, bit for bit, then the clients' float formula on top."""
import numpy as np
import pytest


def client_rxx(packet_cols_int8, L):
    """heatmap2d2.cpp:189-199 / pmusic.m:12-14 on one packet: columns scaled by 1/128 (zmqsdr.c:126),
    reference column dropped, column means removed, X^H X / rows."""
    z = packet_cols_int8.astype(np.float64)
    X = ((z[..., 0] + 1j * z[..., 1]) / 128.0).T                 # rows x columns
    X = X[:, 1:]
    X = X - X.mean(axis=0)
    return X.conj().T @ X / L


def test_rxx_from_sums_is_the_client_formula(pkg):
    rng = np.random.default_rng(3)
    nch, L = 5, 256
    cols = rng.integers(-128, 128, (nch, L, 2)).astype(np.int8)
    z = cols[..., 0].astype(np.int64) + 1j * cols[..., 1].astype(np.int64)
    S = np.conj(z) @ z.T
    m = z.sum(axis=1)
    np.testing.assert_allclose(pkg.rxx_from_sums(S, m, L), client_rxx(cols, L), rtol=1e-12, atol=1e-12)


@pytest.mark.gpu
@pytest.mark.parametrize("C,B,nb", [(21, 16384, 12), (3, 4096, 5), (2, 131072, 2),
                                    (40, 1024, 3), (30, 16384, 4), (22, 4096, 3)])
def test_covariance_sums_exact(pkg, synth, C, B, nb):
    """Both kernels of cov_kernel.cu: int8 tensor-core MMA (up to 22 columns, the first three cases) and the
    dp4a kernel that takes over above that (the last three)."""
    L = B // 2
    iq, _ = synth.make_blocks(C, B, nb, seed=77 + C, delays=np.zeros(C, dtype=int))
    # full-scale bytes on one channel: the extremes (-128) must not wrap anywhere in the integer path
    iq[:, 1, :64] = 0
    eng = pkg.CoherentCuda(C, B)
    out = eng.batch(iq, lag_mode=pkg.LAG_NONE)
    S, m = eng.covariance_sums(nb)
    eng.close()
    hdr = 16 + 4 * (C + 1)
    cols = out["packets"][:, hdr:].reshape(nb, C + 1, L, 2).astype(np.int64)
    zr, zi = cols[..., 0], cols[..., 1]
    re = np.einsum("bin,bjn->bij", zr, zr) + np.einsum("bin,bjn->bij", zi, zi)
    im = np.einsum("bin,bjn->bij", zr, zi) - np.einsum("bin,bjn->bij", zi, zr)
    np.testing.assert_array_equal(S.real.astype(np.int64), re)
    np.testing.assert_array_equal(S.imag.astype(np.int64), im)
    np.testing.assert_array_equal(m.real.astype(np.int64), zr.sum(axis=2))
    np.testing.assert_array_equal(m.imag.astype(np.int64), zi.sum(axis=2))
    got = pkg.rxx_from_sums(S[0], m[0], L)
    np.testing.assert_allclose(got, client_rxx(out["packets"][0, hdr:].reshape(C + 1, L, 2), L), rtol=1e-9, atol=1e-12)
    # coherent, phase-corrected columns of one noise source: the covariance is (nearly) rank one
    if C >= 3:
        w = np.linalg.eigvalsh(got)
        assert w[-1] > 20 * w[-2]
