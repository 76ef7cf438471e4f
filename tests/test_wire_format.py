"""Wire format after the path: packets built by the path, published the way
cpacketize::send does (src/cpacketizer.cc:109-129) and read back by a subscriber that restates the
Matlab client's receive branch (matlabclient/zmqsdr.c:118-144), over a real ZMQ PUB/SUB pair.
"""
import importlib
import time

import numpy as np
import pytest

zmq = pytest.importorskip("zmq")


def zmqsdr_receive(buf: bytes):
    """matlabclient/zmqsdr.c:118-144: gseq, c, r from the first three u32; samples from byte
    16 + 4c on, r*c complex values scaled by 1/128 into an r x c column-major complex matrix;
    c per-channel sequence numbers from u32[4..]."""
    u32 = np.frombuffer(buf, dtype=np.uint32, count=3)
    gseq, c, r = int(u32[0]), int(u32[1]), int(u32[2])
    offset = 16 + 4 * c
    s8 = np.frombuffer(buf, dtype=np.int8, count=2 * r * c, offset=offset)
    flat = (np.float32(1.0 / 128.0) * s8.astype(np.float32)).view(np.complex64)
    m = flat.reshape(c, r).T                                 # Matlab r x c, column-major
    seq = np.frombuffer(buf, dtype=np.uint32, count=c, offset=16)
    return m, gseq, seq


def _wire(pkg):
    return importlib.import_module(pkg.__name__ + ".wire")


def _roundtrip(wire, packets, phases, nch, B, noheader=False):
    port = 15555 + int(time.time() * 1000) % 20000
    got, gotp = [], []
    with wire.PacketPublisher(nch, B, noheader, address=f"tcp://127.0.0.1:{port}",
                              debug_address=f"tcp://127.0.0.1:{port + 1}") as pub:
        ctx = zmq.Context(1)
        sub = ctx.socket(zmq.SUB); sub.setsockopt(zmq.SUBSCRIBE, b""); sub.setsockopt(zmq.RCVTIMEO, 5000)
        subp = ctx.socket(zmq.SUB); subp.setsockopt(zmq.SUBSCRIBE, b""); subp.setsockopt(zmq.RCVTIMEO, 5000)
        sub.connect(f"tcp://127.0.0.1:{port}"); subp.connect(f"tcp://127.0.0.1:{port + 1}")
        # PUB/SUB joins are asynchronous: send a primer until the subscriber hears one
        for _ in range(200):
            pub.send(packets[0], phases[0])
            try:
                sub.recv(zmq.NOBLOCK); break
            except zmq.Again:
                time.sleep(0.01)
        time.sleep(0.1)
        while True:                                          # drain primers
            try:
                sub.recv(zmq.NOBLOCK)
            except zmq.Again:
                break
        while True:
            try:
                subp.recv(zmq.NOBLOCK)
            except zmq.Again:
                break
        for b in range(len(packets)):
            pub.send(packets[b], phases[b])
            got.append(sub.recv()); gotp.append(subp.recv())
        sub.close(0); subp.close(0); ctx.term()
    return got, gotp


def _check_messages(wire, got, gotp, packets, phases, iq_ref_col, nch, B, seq0, readcnt0):
    L = B // 2
    for b, msg in enumerate(got):
        assert len(msg) == wire.wire_length(nch, B) == 16 + 4 * nch + 2 * nch * B   # the 2x quirk
        m, gseq, seq = zmqsdr_receive(msg)
        assert m.shape == (L, nch) and gseq == seq0 + b
        np.testing.assert_array_equal(seq, np.full(nch, readcnt0 + b, np.uint32))
        cols = packets[b][16 + 4 * nch:].view(np.int8).reshape(nch, L, 2).astype(np.float32) / np.float32(128.0)
        np.testing.assert_array_equal(m, (cols[..., 0] + 1j * cols[..., 1]).T.astype(np.complex64))
        # reference column is the raw signed input (cpacketizer.cc:137-156)
        np.testing.assert_array_equal((m[:, 0].view(np.float32) * 128).astype(np.int8), (iq_ref_col[b] ^ 0x80).view(np.int8))
        assert not any(msg[16 + 4 * nch + nch * B:])         # never-written tail
        np.testing.assert_array_equal(np.frombuffer(gotp[b], np.complex64), phases[b])


def test_wire_length_and_padding(pkg):
    wire = _wire(pkg)
    assert wire.wire_length(22, 16384) == 16 + 88 + 2 * 22 * 16384
    assert wire.wire_length(4, 16384, noheader=True) == 2 * 4 * 16384
    pk = np.arange(16 + 16 + 4 * 64, dtype=np.uint8)
    msg = wire.wire_message(pk, 4, 64)
    assert len(msg) == 16 + 16 + 2 * 4 * 64 and msg[:pk.size] == pk.tobytes() and not any(msg[pk.size:])
    with pytest.raises(ValueError):
        wire.wire_message(pk[:-1], 4, 64)


def test_oracle_packets_reach_a_zmqsdr_style_subscriber(pkg, synth, oracle):
    """CPU: packets from the oracle through the product's publisher to the restated client."""
    wire = _wire(pkg)
    C, B, nb = 3, 4096, 4
    iq, _ = synth.make_blocks(C, B, nb, seed=11, delays=np.array([0, 5, -9]))
    out = oracle.process_batch(oracle.make_config(C, B), iq, readcnt0=8)
    pk = out["packets"].view(np.uint8)
    got, gotp = _roundtrip(wire, pk, out["phase"], C + 1, B)
    _check_messages(wire, got, gotp, pk, out["phase"], iq[:, 0], C + 1, B, 0, 8)


@pytest.mark.gpu
def test_gpu_packets_reach_a_zmqsdr_style_subscriber(pkg, synth):
    """GPU: per-block streaming calls (what ccoherent::threadf makes) -> publisher -> client."""
    wire = _wire(pkg)
    C, B, nb = 3, 16384, 5
    iq, _ = synth.make_blocks(C, B, nb, seed=12, delays=np.array([1, -2, 30]))
    eng = pkg.CoherentCuda(C, B)
    pk, ph = [], []
    for b in range(nb):
        eng.submit_block([iq[b, c] for c in range(C + 1)], readcnt=np.full(C + 1, 8 + b, np.uint32),
                         lag_request=np.ones(C + 1, np.uint8))
        r = eng.wait_block()
        pk.append(r["packet"].view(np.uint8)); ph.append(r["phase"])
    eng.close()
    got, gotp = _roundtrip(wire, pk, ph, C + 1, B)
    _check_messages(wire, got, gotp, pk, ph, iq[:, 0], C + 1, B, 0, 8)
