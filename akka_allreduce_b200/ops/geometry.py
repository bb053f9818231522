"""Host-side mirror of ``include/aar/protocol_math.h`` / ``device_layout.h`` for launch set-up.

Pure Python + numpy float32 so the thresholds come out exactly as the C++/CUDA code computes
them: ``int(float32(th) * float32(n))`` (ScatteredDataBuffer.scala:9, ReducedDataBuffer.scala:16).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

ALIGN_ELEMS = 128


def ceil_div(a: int, b: int) -> int:
    return -(-a // b) if b > 0 else 0


def threshold_count(th: float, n: int) -> int:
    return int(np.float32(th) * np.float32(n))


def gate_count(th_allreduce: float, total: int) -> int:
    """Smallest numComplete with float32(numComplete) >= float32(total)*thAllreduce (AllreduceMaster.scala:58)."""
    need = np.float32(total) * np.float32(th_allreduce)
    k = 0
    while np.float32(k) < need:
        k += 1
    return k


@dataclass(frozen=True)
class Geometry:
    n: int
    max_chunk: int
    world: int
    slots: int            # maxLag + 1
    step: int             # block step actually used
    reference_partition: bool

    @staticmethod
    def make(n: int, max_chunk: int, world: int, max_lag: int, align_blocks: bool = False) -> "Geometry":
        step = ceil_div(n, world)
        if align_blocks:
            step = ceil_div(step, ALIGN_ELEMS) * ALIGN_ELEMS
        return Geometry(n, max_chunk, world, max_lag + 1, step, not align_blocks)

    def start(self, i: int) -> int:
        return min(i * self.step, self.n)

    def end(self, i: int) -> int:
        return self.n if i >= self.world - 1 else self.start(i + 1)

    def size(self, i: int) -> int:
        return self.end(i) - self.start(i)

    def chunks_of(self, i: int) -> int:
        return ceil_div(self.size(i), self.max_chunk)

    @property
    def total_chunks(self) -> int:
        return sum(self.chunks_of(i) for i in range(self.world))

    @property
    def block_step_arg(self) -> int:
        """Value for ``AarLaunch.block_step``: 0 selects the reference partition."""
        return 0 if self.reference_partition else self.step

    def req_reduce(self, th_reduce: float) -> int:
        return threshold_count(th_reduce, self.world)

    def req_complete(self, th_complete: float) -> int:
        return threshold_count(th_complete, self.total_chunks)

    def chunk_ranges(self):
        """(owner, chunk, global_start, global_end) for every reduced chunk of a round."""
        for o in range(self.world):
            b = self.size(o)
            for c in range(self.chunks_of(o)):
                yield o, c, self.start(o) + min(c * self.max_chunk, b), self.start(o) + min((c + 1) * self.max_chunk, b)


def default_ctas(message_bytes: int, stream: bool = False) -> int:
    """Work CTAs per rank.  The generic kernel holds one 512-thread CTA per SM, the stream kernel two (its unit of
    work is a warp tile, so it wants about one warp per 2-32 KB of the largest phase); the library clamps the
    number to what the device keeps resident."""
    if stream:
        if message_bytes <= 64 << 10:
            return 4
        if message_bytes <= 1 << 20:
            return 16
        if message_bytes <= 4 << 20:
            return 64
        if message_bytes <= 16 << 20:
            return 128
        return 512
    if message_bytes <= 64 << 10:
        return 4
    if message_bytes <= 1 << 20:
        return 16
    if message_bytes <= 16 << 20:
        return 32
    return 128  # measured best on 8 x B200 for 64 MB - 1 GB (profiles/sweep_8gpu.md)
