"""Single-worker protocol harness: one real engine, every peer and the master are a probe.

Mirrors the technique of the reference's AllreduceSpec (src/test/scala/AllreduceSpec.scala:
774-818): outgoing messages are captured in order and asserted one by one; incoming messages
are injected by hand.
"""
from collections import deque

import numpy as np

from akka_allreduce_b200 import host as h


def basic_source(size):
    return lambda it: np.arange(size, dtype=np.float32) + np.float32(it)


class Harness:
    def __init__(self, source=None, sink=None, size=8, options=None):
        self.bus = h.LocalBus()
        self.probe = self.bus.add_probe()
        self.outputs = []
        src = source or basic_source(size)

        def default_sink(data, count, iteration):
            self.outputs.append((data.tolist(), count.tolist(), iteration))

        self._sink = sink
        def both(data, count, iteration):
            default_sink(data, count, iteration)
            if self._sink:
                self._sink(data, count, iteration)

        self.ep, self.worker = self.bus.add_worker(src, both, options or h.WorkerOptions())
        self.fifo = deque()

    # -- helpers ------------------------------------------------------------------------
    def peers_as_probe(self, n, real=None):
        m = {i: self.probe for i in range(n)}
        if real is not None:
            m[real] = self.ep
        return m

    def init(self, workers, worker_num, idx, th_reduce, th_complete, max_lag, data_size, max_chunk):
        self.tell(h.InitWorkers(workers, worker_num, self.probe, idx, th_reduce, th_complete, max_lag,
                                data_size, max_chunk))

    def tell(self, msg):
        self.bus.send(self.ep, msg)
        self.bus.run()
        self.fifo.extend(self.bus.take(self.probe))

    def next(self):
        assert self.fifo, "expected a message, probe received none"
        return self.fifo.popleft()

    def expect_scatter(self, value, src, dest, chunk, rnd):
        m = self.next()
        assert isinstance(m, h.ScatterBlock), m
        assert (m.src_id, m.dest_id, m.chunk_id, m.round) == (src, dest, chunk, rnd), m
        assert m.value.tolist() == [float(v) for v in value], m

    def expect_reduce(self, value, src, dest, chunk, rnd, count):
        m = self.next()
        assert isinstance(m, h.ReduceBlock), m
        assert (m.src_id, m.dest_id, m.chunk_id, m.round, m.count) == (src, dest, chunk, rnd, count), m
        assert m.value.tolist() == [float(v) for v in value], m

    def expect_complete(self, src, rnd):
        m = self.next()
        assert isinstance(m, h.CompleteAllreduce), m
        assert (m.src_id, m.round) == (src, rnd), m

    def expect_no_msg(self):
        assert not self.fifo, f"unexpected messages: {list(self.fifo)}"

    def fish_complete(self, src, rnd):
        while self.fifo:
            m = self.fifo.popleft()
            if isinstance(m, h.CompleteAllreduce) and (m.src_id, m.round) == (src, rnd):
                return
        raise AssertionError(f"CompleteAllreduce({src},{rnd}) never arrived")
