"""Timing helper for the two downstream kernels (covariance, fractional lag) on 1024 URA21 blocks;
run under ncu --metrics gpu__time_duration.sum -k regex:k_covariance|k_fraclag for the kernel times."""
import sys, time; sys.path.insert(0, '.')
import numpy as np, torch, cohpkg
pkg = cohpkg.load(); synth = cohpkg.synth()
C, B, nb = 21, 16384, 1024
iq, _ = synth.make_blocks(C, B, 128, seed=1, delays=np.zeros(C, dtype=int))
iq = np.ascontiguousarray(np.tile(iq, (8, 1, 1)))
eng = pkg.CoherentCuda(C, B)
out = eng.batch(iq, lag_mode=pkg.LAG_NONE)
for i in range(3):
    torch.cuda.synchronize(); t0 = time.time(); S, m = eng.covariance_sums(nb); t1 = time.time()
    print("covariance_sums 1024 blocks incl. alloc + D2H: %.2f ms" % ((t1 - t0) * 1e3))
t0 = time.time(); fr = eng.fraclag(nb); print('fraclag 1024 blocks incl. alloc + D2H: %.2f ms' % ((time.time() - t0) * 1e3))
eng.close()
