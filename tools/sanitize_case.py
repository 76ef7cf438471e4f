#!/usr/bin/env python
"""Small parity cases for compute-sanitizer (memcheck / racecheck / synccheck), one tool per run:

    compute-sanitizer --tool racecheck python tools/sanitize_case.py

four.cfg geometry (3+1 channels x 8192 samples) through k_ref_spectrum / k_correlate (persistent form) /
k_apply (TMA row fetch rewritten in place, decoupled look-back), SYNC with software shift, TRACK, mixed
requests with the noise source gated, and one block size above 8192 samples (k_big_*).  Every case is
checked against the oracle, so a sanitizer-clean run is also a correct one."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cohpkg                                   # noqa: E402
from oracle import oracle as O                  # noqa: E402


def main():
    pkg, synth = cohpkg.load(), cohpkg.synth()
    O.build()
    cases = [(3, 16384, 6), (2, 32768, 2), (5, 2048, 60)]
    for C, B, nb in cases:
        delays = np.array([0, 17, -250, 5, -3][:C])
        iq, _ = synth.make_blocks(C, B, nb, seed=0xC0DE, delays=delays)
        eng = pkg.CoherentCuda(C, B, apply_shift=True)
        cfg = O.make_config(C, B, apply_shift=1, int_dot=1)
        out = eng.batch(iq, lag_mode=pkg.LAG_ALL)
        ref = O.process_batch(cfg, iq, lag_mode=1)
        assert (out["lag"] == ref["lag"]).all() and (out["packets"] == ref["packets"]).all(), ("sync", C, B)
        req = np.zeros((nb, C + 1), np.uint8); req[0, 1:] = 1; req[nb // 2, 1] = 1
        noise = np.ones(nb, np.uint8); noise[1:2] = 0
        eng2 = pkg.CoherentCuda(C, B, apply_shift=True)
        out = eng2.batch(iq, lag_request=req, refnoise=noise)
        ref = O.process_batch(cfg, iq, lag_request=req, refnoise=noise)
        assert (out["packets"] == ref["packets"]).all(), ("mixed", C, B)
        out = eng.batch(iq, lag_mode=pkg.LAG_NONE)             # TRACK on the held lags of the first batch
        assert out["packets"].shape == ref["packets"].shape
        eng.close(); eng2.close()
        print("case", C, B, nb, "ok", flush=True)


if __name__ == "__main__":
    main()
