"""Rank-0 wire side of the path: what leaves the machine after the packets were built on the GPU.

The drop-in build keeps the reference's own cpacketize transport (src/cpacketizer.cc:58-74,
109-129) -- nothing here is needed there.  This module is for the standalone host API
(``CoherentCuda.batch`` / ``wait_block`` and the gathered packets of ``sharding.gather_packets``):
it publishes the packet bytes so that the reference's clients (matlabclient/zmqsdr.c,
beamformclient) keep working unchanged.

Wire format (include/cpacketizer.h:32-37, src/cpacketizer.cc:91-96,109-129):
  samples, PUB tcp://*:5555, one message per block:
    u32 globalseqn, u32 N (= C+1 columns), u32 L (complex samples), u32 unused = 0,
    u32 readcnt[N], then N columns of L interleaved int8 (I, Q) pairs;
    the message is ``packetlength(N, blocksize)`` bytes long with blocksize in BYTES, i.e.
    ``16 + 4N + 2*N*B`` -- twice the column payload; the second half is never written (zeros).
    ``-R`` raw mode: no header, ``2*N*B`` bytes.
  phase factors, PUB tcp://*:5557: N complex64, one message per block (cpacketizer.cc:127).
"""
from __future__ import annotations

import numpy as np

SAMPLE_PORT = 5555          # src/main.cc:258
DEBUG_PORT = 5557           # src/cpacketizer.cc:65


def wire_length(nch: int, blocksize: int, noheader: bool = False) -> int:
    """cpacketize::packetlength (src/cpacketizer.cc:91-96) as the reference calls it from init():
    L is passed the block size in bytes, hence the factor-2 over the written payload."""
    return 2 * nch * blocksize if noheader else 16 + 4 * nch + 2 * nch * blocksize


def wire_message(packet: np.ndarray, nch: int, blocksize: int, noheader: bool = False) -> bytes:
    """One packet as produced by the path (``[hdr0][readcnt][columns]``, or columns only in raw
    mode) padded to the length the reference publishes."""
    packet = np.ascontiguousarray(packet).view(np.uint8).ravel()
    written = (0 if noheader else 16 + 4 * nch) + nch * blocksize
    if packet.size != written:
        raise ValueError(f"packet is {packet.size} bytes, expected {written}")
    msg = np.zeros(wire_length(nch, blocksize, noheader), dtype=np.uint8)
    msg[:written] = packet
    return msg.tobytes()


class PacketPublisher:
    """ZMQ PUB pair mirroring cpacketize::init / send."""

    def __init__(self, nch: int, blocksize: int, noheader: bool = False,
                 address: str = f"tcp://*:{SAMPLE_PORT}", debug_address: str | None = f"tcp://*:{DEBUG_PORT}"):
        import zmq
        self.nch, self.B, self.noheader = nch, blocksize, noheader
        self.ctx = zmq.Context(1)
        self.sock = self.ctx.socket(zmq.PUB)
        self.sock.bind(address)
        self.debug = None
        if debug_address:
            self.debug = self.ctx.socket(zmq.PUB)
            self.debug.bind(debug_address)

    def send(self, packet: np.ndarray, phase: np.ndarray | None = None) -> None:
        self.sock.send(wire_message(packet, self.nch, self.B, self.noheader))
        if self.debug is not None:
            p = np.zeros(self.nch, dtype=np.complex64) if phase is None else np.ascontiguousarray(phase, dtype=np.complex64)
            self.debug.send(p.tobytes())

    def close(self) -> None:
        self.sock.close(0)
        if self.debug is not None:
            self.debug.close(0)
        self.ctx.term()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
