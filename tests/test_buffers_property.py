"""Property tests of the lag-ring buffers against a few-line numpy model (hypothesis).

The reference's buffer specs check a handful of fixed cases; here random sequences of store / advance /
reduce / assemble are replayed on the C++ buffers and on a direct transcription of the semantics
(sum over ALL peer rows, arrival counter per chunk id, duplicates overwrite and still
count, recycled slot zeroed, chunk counts expanded per element)."""
import numpy as np
from hypothesis import given, settings, strategies as st

from akka_allreduce_b200 import host as h


class ModelRing:
    def __init__(self, width, peers, slots, mc):
        self.w, self.p, self.s, self.mc = width, peers, slots, mc
        self.chunks = -(-width // mc)
        self.data = np.zeros((slots, peers, width), np.float32)
        self.arr = np.zeros((slots, self.chunks), np.int64)
        self.cnt = np.zeros((slots, peers, self.chunks), np.int64)
        self.head = 0

    def slot(self, row):
        return (row + self.head) % self.s

    def store(self, vals, row, src, chunk, count=None):
        at = chunk * self.mc
        self.data[self.slot(row), src, at:at + len(vals)] = vals
        self.arr[self.slot(row), chunk] += 1
        if count is not None:
            self.cnt[self.slot(row), src, chunk] = count

    def advance(self):
        old = self.head
        self.head = (self.head + 1) % self.s
        self.data[old] = 0
        self.arr[old] = 0
        self.cnt[old] = 0


ops = st.lists(st.tuples(st.sampled_from(["store", "store", "store", "advance", "check"]), st.integers(0, 10 ** 6)),
               min_size=1, max_size=60)


@settings(max_examples=60, deadline=None)
@given(width=st.integers(1, 12), peers=st.integers(1, 5), slots=st.integers(1, 4), mc=st.integers(1, 5), seq=ops)
def test_scatter_buffer_matches_model(width, peers, slots, mc, seq):
    buf = h.ScatterBuffer(width, peers, slots, 0.75, mc)
    m = ModelRing(width, peers, slots, mc)
    assert buf.chunks == m.chunks and buf.required == h.threshold_count(0.75, peers)
    for op, r in seq:
        rng = np.random.default_rng(r)
        row, src, chunk = int(rng.integers(slots)), int(rng.integers(peers)), int(rng.integers(m.chunks))
        if op == "store":
            n = min(mc, width - chunk * mc)
            vals = rng.integers(-8, 8, n).astype(np.float32)
            buf.store(vals, row, src, chunk)
            m.store(vals, row, src, chunk)
        elif op == "advance":
            buf.advance()
            m.advance()
        else:
            got, cnt = buf.reduce(row, chunk)
            lo, hi = chunk * mc, min(width, (chunk + 1) * mc)
            assert np.array_equal(got, m.data[m.slot(row), :, lo:hi].sum(axis=0))
            assert cnt == m.arr[m.slot(row), chunk] == buf.arrivals(row, chunk)
            assert buf.at_threshold(row, chunk) == (cnt == buf.required)


@settings(max_examples=60, deadline=None)
@given(maxb=st.integers(1, 10), peers=st.integers(1, 5), slots=st.integers(1, 4), mc=st.integers(1, 4),
       short=st.integers(0, 9), seq=ops)
def test_reduce_buffer_matches_model(maxb, peers, slots, mc, short, seq):
    minb = max(1, maxb - short % maxb) if maxb > 1 else 1
    total = maxb * (peers - 1) + minb
    buf = h.ReduceBuffer(maxb, minb, total, peers, slots, 0.7, mc)
    m = ModelRing(maxb, peers, slots, mc)
    assert buf.required == h.threshold_count(0.7, m.chunks * (peers - 1) + -(-minb // mc))
    for op, r in seq:
        rng = np.random.default_rng(r)
        row, src, chunk = int(rng.integers(slots)), int(rng.integers(peers)), int(rng.integers(m.chunks))
        if op == "store":
            n = min(mc, maxb - chunk * mc)
            vals = rng.integers(-8, 8, n).astype(np.float32)
            count = int(rng.integers(0, peers + 1))
            buf.store(vals, row, src, chunk, count)
            m.store(vals, row, src, chunk, count)
        elif op == "advance":
            buf.advance()
            m.advance()
        else:
            d, c = buf.assemble(row)
            want_d = m.data[m.slot(row)].reshape(-1)[:total]
            per_elem = np.repeat(m.cnt[m.slot(row)], mc, axis=1)[:, :maxb].reshape(-1)[:total]
            assert np.array_equal(d, want_d) and np.array_equal(c, per_elem)
            assert buf.at_threshold(row) == (m.arr[m.slot(row)].sum() == buf.required)
