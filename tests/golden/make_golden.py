#!/usr/bin/env python
"""Generates tests/golden/*.npz from the REFERENCE ITSELF.

Runs in the build container only (needs /root/reference): oracle/_ref/libref_cdsp.so is the
reference's own cdsp.cc / ccoherent.cc / csdrdevice.cc / crtlsdr.cc / cpacketizer.cc compiled
unmodified against the declaration shims in oracle/shim/ and driven block by block by
oracle/ref_harness.cc.  Each fixture stores the raw offset-binary input blocks, the request
flags, and what the reference produced: the published packet bytes, lagpoint.lag / .mag,
csdrdevice::phasecorr and which channels got a set_lag() in each block.

    python tests/golden/make_golden.py          # rewrites the .npz files next to this script
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import cohpkg                              # noqa: E402
from oracle import oracle as O             # noqa: E402

CASES = {
    # name: (C, B, n_blocks, noheader, delays, request pattern, refnoise pattern)
    "four_cfg":   dict(C=3, B=16384, nb=3, noheader=0, delays=[0, 17, -250], req="all", noise="on"),
    "edge_delay": dict(C=6, B=2048, nb=6, noheader=0, delays=[1, -1, 512, -512, 640, -640], req="all", noise="on"),
    "mixed_slots": dict(C=9, B=1024, nb=24, noheader=0, delays=[0, 3, -3, 40, -77, 128, -200, 5, -9], req="random", noise="gated"),
    "raw_mode":   dict(C=2, B=512, nb=10, noheader=1, delays=[7, -31], req="first", noise="on"),
    "track_only": dict(C=4, B=2048, nb=10, noheader=0, delays=[0, 0, 0, 0], req="none", noise="on"),
    # round 2: a dongle that is not streaming (all-0x80 rows: zero energy => |corr| = 0 => NaN phase factor for good,
    # src/csdrdevice.cc:63; every correlation value ties at 0 => first index), it comes back at block 5
    "dead_channel": dict(C=3, B=1024, nb=9, noheader=0, delays=[5, -9, 33], req="all", noise="on", dead=(2, 0, 5)),
    # drift / re-sync (BASELINE config 5, src/console.cc:285-289): delays step at known blocks; the channel
    # re-requests its lag at the step and, for one channel, three blocks late
    "delay_steps": dict(C=4, B=2048, nb=16, noheader=0, delays=[3, -20, 64, 0], req="resync", noise="on",
                        steps=[(5, 1, 11), (9, 3, -70), (12, 1, -2)], late={9: 3}),
    # a block longer than one SM's shared memory holds (L = 16384 > 8192: the split-transform path)
    "long_block": dict(C=2, B=32768, nb=3, noheader=0, delays=[100, -3000], req="all", noise="on"),
    # round 3: the two geometries the benchmarks run.  URA21 (examplecfg/URA21.cfg:5-26: 21 + 1 dongles) with the stock
    # nfft = 8 (seven channels served per block, src/ccoherent.cc:124) and with nfft = 22 (every channel, every block:
    # the bench's SYNC mode); BASELINE config 3's 65536-sample blocks
    "ura21_nfft8": dict(C=21, B=16384, nb=3, noheader=0, req="all", noise="on", nfft=8,
                        delays=[0, 17, -250, 64, -64, 1, -1, 33, -47, 12, -12, 5, -5, 60, -60, 25, -25, 8, -8, 40, -40]),
    "ura21_nfft22": dict(C=21, B=16384, nb=3, noheader=0, req="all", noise="on", nfft=22,
                         delays=[0, 17, -250, 64, -64, 1, -1, 33, -47, 12, -12, 5, -5, 60, -60, 25, -25, 8, -8, 40, -40]),
    "big65536": dict(C=2, B=131072, nb=2, noheader=0, delays=[1000, -20000], req="all", noise="on"),
}


def build_case(name, spec):
    synth = cohpkg.synth()
    C, B, nb = spec["C"], spec["B"], spec["nb"]
    seed = 0x601D + sum(map(ord, name))
    iq, info = synth.make_blocks(C, B, nb, seed=seed, delays=np.array(spec["delays"]), delay_steps=spec.get("steps"))
    if "dead" in spec:
        ch, b0, b1 = spec["dead"]
        iq[b0:b1, ch, :] = 0x80
    rng = np.random.default_rng(seed)
    req = np.zeros((nb, C + 1), dtype=np.uint8)
    if spec["req"] == "all":
        req[:, 1:] = 1
    elif spec["req"] == "random":
        req[:, 1:] = rng.random((nb, C)) < 0.6
        req[0, 1:] = 1
    elif spec["req"] == "first":
        req[0, 1:] = 1
    elif spec["req"] == "resync":
        req[0, 1:] = 1
        for (b0, ch, _) in spec["steps"]:
            req[b0 + spec["late"].get(b0, 0), ch] = 1
    noise = np.ones(nb, dtype=np.uint8)
    if spec["noise"] == "gated":
        noise[nb // 4: nb // 2] = 0
        noise[-3:] = 0
    nfft = spec.get("nfft", 8)
    with O.RefCdsp(C, B, noheader=bool(spec["noheader"]), nfft=nfft) as R:
        out = R.run_batch(iq, lag_request=req, refnoise=noise)
        sent_len = R.sent_len
    return dict(iq=iq, lag_request=req, refnoise=noise, packets=out["packets"], lag=out["lag"],
                mag=out["mag"], phase=out["phase"], measured=out["measured"],
                meta=np.array([C, B, nb, spec["noheader"], nfft, sent_len], dtype=np.int64),
                delays=np.array(spec["delays"], dtype=np.int64), phases=info["phases"], delay_matrix=info["delays"])


def main():
    O.build()
    if not O.ref_available():
        sys.exit("oracle/_ref/libref_cdsp.so is missing: needs /root/reference (make -C oracle ref)")
    only = sys.argv[1:]
    for name, spec in CASES.items():
        if only and name not in only:
            continue
        d = build_case(name, spec)
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **d)
        print(f"{name}: {os.path.getsize(path)} bytes, lag[-1]={d['lag'][-1].tolist()}")


if __name__ == "__main__":
    main()
