"""Ring-buffer unit tests (no engine, no transport).

Scenarios of the reference's ScatteredDataBufferSpec.scala:1-109 and
ReducedDataBufferSpec.scala:1-166 against the C++ buffers.
"""
import numpy as np
import pytest

from akka_allreduce_b200 import host as h

rng = np.random.default_rng(0)


def rand(n):
    return rng.random(n, dtype=np.float32)


class TestScatterBuffer:
    block, peers, slots, th, mc, row = 5, 4, 4, 0.75, 3, 1

    def make(self):
        return h.ScatterBuffer(self.block, self.peers, self.slots, self.th, self.mc)

    def test_shape(self):  # ScatteredDataBufferSpec.scala:24-30
        b = self.make()
        assert (b.slots, b.peers, b.width) == (self.slots, self.peers, self.block)
        assert b.chunks == 2 and b.required == 3

    def test_last_chunk_overflow_raises(self):  # :32-42
        b = self.make()
        last = b.chunks - 1
        with pytest.raises(IndexError):
            b.store(rand(self.mc), self.row, 0, last)
        excess = b.chunks * self.mc - self.block
        b.store(rand(self.mc - excess), self.row, 0, last)

    def test_threshold_fires_at_exactly_three(self):  # :44-54
        b = self.make()
        seen = []
        for i in range(3):
            b.store(rand(self.mc), self.row, i, 0)
            seen.append(b.at_threshold(self.row, 0))
        assert seen == [False, False, True]
        b.store(rand(self.mc), self.row, 3, 0)
        assert not b.at_threshold(self.row, 0)  # equality: fires once

    def test_reduce_counts(self):  # :56-64
        b = self.make()
        s, c = b.reduce(0, 0)
        assert c == 0 and s.sum() == 0
        for i in range(3):
            b.store(rand(self.mc), self.row, i, 0)
        assert b.reduce(self.row, 0)[1] == 3

    def test_summation_and_row_isolation(self):  # :66-102
        b = h.ScatterBuffer(2, 2, 2, 1.0, 3)
        for i in range(2):
            b.store(np.full(2, i, np.float32), 0, i, 0)
            assert b.reduce(0, 0)[1] == i + 1
        assert b.reduce(0, 0)[0].tolist() == [1.0, 1.0]
        s, c = b.reduce(1, 0)
        assert c == 0 and s.tolist() == [0.0, 0.0]

    def test_advance_recycles_oldest_slot(self):  # AllReduceBuffer.scala:38-42
        b = h.ScatterBuffer(2, 2, 2, 1.0, 2)
        b.store(np.array([1, 2], np.float32), 0, 0, 0)
        b.store(np.array([5, 6], np.float32), 1, 1, 0)
        b.advance()
        assert b.reduce(0, 0)[0].tolist() == [5.0, 6.0] and b.arrivals(0, 0) == 1
        assert b.reduce(1, 0)[0].tolist() == [0.0, 0.0] and b.arrivals(1, 0) == 0


class TestReduceBufferEven:
    maxb, minb, peers, slots, th, mc, total, row = 5, 5, 3, 4, 0.7, 2, 15, 1

    def make(self):
        return h.ReduceBuffer(self.maxb, self.minb, self.total, self.peers, self.slots, self.th, self.mc)

    def test_shape_and_zero_counts(self):  # ReducedDataBufferSpec.scala:24-38
        b = self.make()
        assert (b.slots, b.peers, b.width) == (self.slots, self.peers, self.maxb)
        d, c = b.assemble(self.row)
        assert d.sum() == 0 and c.sum() == 0 and len(d) == len(c) == self.total

    def test_store_first_chunk(self):  # :40-52
        b = self.make()
        x = rand(self.mc)
        b.store(x, self.row, 0, 0, self.peers)
        d, c = b.assemble(self.row)
        assert d[: self.mc].tolist() == x.tolist()
        assert c[: self.mc].tolist() == [self.peers] * self.mc

    def test_short_last_chunk(self):  # :54-70
        b = self.make()
        last = b.chunks - 1
        with pytest.raises(IndexError):
            b.store(rand(self.mc), self.row, self.peers - 1, last, self.peers)
        n = self.maxb - last * self.mc
        x = rand(n)
        b.store(x, self.row, self.peers - 1, last, self.peers)
        d, _ = b.assemble(self.row)
        assert d[self.total - n:].tolist() == x.tolist()

    def test_completion_and_missing_chunks(self):  # :72-119 (golden vector B8)
        b = self.make()
        assert b.required == 6
        b.store(rand(2), 1, 0, 0, 3)
        for src, chunk in [(0, 1), (1, 0), (1, 1)]:
            b.store(rand(2), 1, src, chunk, 3)
            assert not b.at_threshold(1)
        b.store(rand(1), 1, 2, 2, 3)
        assert not b.at_threshold(1)
        b.store(rand(2), 1, 2, 1, 3)
        assert b.at_threshold(1)
        d, c = b.assemble(1)
        missing = [4, 9, 10, 11]
        assert all(d[i] == 0 and c[i] == 0 for i in missing)
        assert all(c[i] == 3 for i in range(self.total) if i not in missing)


def test_reduce_buffer_uneven_blocks():  # ReducedDataBufferSpec.scala:124-160
    b = h.ReduceBuffer(6, 4, 16, 3, 4, 1.0, 2)
    assert b.required == 8
    for chunk in range(3):
        for peer in range(2):
            b.store(rand(2), 1, peer, chunk, 3)
            assert not b.at_threshold(1)
    b.store(rand(2), 1, 2, 0, 3)
    assert not b.at_threshold(1)
    b.store(rand(2), 1, 2, 1, 3)
    assert b.at_threshold(1)


def test_partition_matches_reference_formula():
    """AllreduceWorker.scala:240-250 — ceil step, short last block."""
    for n, P in [(8, 4), (3, 2), (9, 3), (778, 4), (10, 2), (6, 2), (4, 4), (1 << 24, 8)]:
        p = h.Partition(n, P)
        step = -(-n // P)
        starts = list(range(0, n, step))
        assert p.reference_defined() and len(starts) == P
        for i in range(P):
            end = n if i >= P - 1 else starts[i + 1]
            assert (p.start(i), p.end(i)) == (starts[i], end)
        assert p.max_block() == p.size(0) and p.min_block() == p.size(P - 1)
    assert not h.Partition(9, 4).reference_defined()
    assert [h.Partition(9, 4).size(i) for i in range(4)] == [3, 3, 3, 0]
