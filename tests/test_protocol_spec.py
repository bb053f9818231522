"""Message-level scenarios of the worker protocol, against the C++ host engine.

Every scenario of the reference's single-actor spec (src/test/scala/AllreduceSpec.scala) is
replayed here with the same parameters and the same expected message sequence; line ranges in
the docstrings point at the scenario being reproduced.
"""
import numpy as np

from akka_allreduce_b200 import host as h
from harness import Harness, basic_source

S, R = h.ScatterBlock, h.ReduceBlock


def test_flushed_output_sums_correctly():
    """AllreduceSpec.scala:46-97 — P=2, id=1, peer 1 is the real worker (inline self-delivery)."""
    gen = lambda i, it: i + float(it)
    t = Harness(size=3)
    workers = t.peers_as_probe(2, real=1)
    t.init(workers, 2, 1, 1.0, 1.0, 5, 3, 2)
    t.tell(h.StartAllreduce(0))
    t.tell(S([2.0], 0, 1, 0, 0))
    t.tell(R([0.0, 2.0], 0, 1, 0, 0, 2))
    t.tell(h.StartAllreduce(1))
    t.tell(S([3.0], 0, 1, 0, 1))
    t.tell(R([2.0, 4.0], 0, 1, 0, 1, 2))
    t.fish_complete(1, 0)
    t.fish_complete(1, 1)
    assert t.outputs[0] == ([2 * gen(i, 0) for i in range(3)], [2, 2, 2], 0)
    assert t.outputs[1] == ([2 * gen(i, 1) for i in range(3)], [2, 2, 2], 1)


def test_early_receiving_reduce():
    """AllreduceSpec.scala:99-139 — reduces for round 3 self-start it; later scatters are ignored."""
    t = Harness()
    t.init(t.peers_as_probe(4), 4, 0, 1.0, 0.8, 5, 8, 2)
    t.tell(h.StartAllreduce(0))
    fut = 3
    for v, s in [([12, 15], 0), ([11, 10], 1), ([10, 20], 2), ([9, 10], 3)]:
        t.tell(R(v, s, 0, 0, fut, 4))
    t.fish_complete(0, fut)
    t.fifo.clear()
    for i in range(4):
        t.tell(S([2.0 * i, 2.0 * i], i, 0, 0, fut))
    t.expect_no_msg()


def test_nodes_live_at_different_times():
    """AllreduceSpec.scala:141-172 — scatter only to known peers; re-init refreshes the peer map."""
    t = Harness()
    workers = t.peers_as_probe(4)
    t.init({0: workers[0]}, 4, 0, 1.0, 1.0, 5, 8, 2)
    t.tell(h.StartAllreduce(0))
    t.expect_scatter([0, 1], 0, 0, 0, 0)
    t.expect_no_msg()
    t.init(workers, 4, 0, 1.0, 1.0, 5, 8, 2)
    t.tell(h.StartAllreduce(1))
    for i in range(4):
        t.expect_scatter([2 * i + 1, 2 * i + 2], 0, i, 0, 1)


def test_single_round():
    """AllreduceSpec.scala:175-213 — completion fires on the 3rd of 4 reduced chunks."""
    t = Harness()
    t.init(t.peers_as_probe(4), 4, 0, 1.0, 0.75, 5, 8, 2)
    t.tell(h.StartAllreduce(0))
    for i in range(4):
        t.expect_scatter([2 * i, 2 * i + 1], 0, i, 0, 0)
    for i in range(4):
        t.tell(S([2.0 * i, 2.0 * i], i, 0, 0, 0))
    for d in range(4):
        t.expect_reduce([12, 12], 0, d, 0, 0, 4)
    for v, s in [([12, 15], 0), ([11, 10], 1), ([10, 20], 2), ([9, 10], 3)]:
        t.tell(R(v, s, 0, 0, 0, 4))
    t.expect_complete(0, 0)
    t.expect_no_msg()  # exactly one completion; the 4th reduce is stale
    assert t.outputs == [([12, 15, 11, 10, 10, 20, 0, 0], [4, 4, 4, 4, 4, 4, 0, 0], 0)]


def test_uneven_size_self_first():
    """AllreduceSpec.scala:215-238 — destination rotation starts at self; blocks [0,2),[2,3)."""
    t = Harness(size=3)
    t.init(t.peers_as_probe(2), 2, 1, 1, 1, 1, 3, 1)
    t.tell(h.StartAllreduce(0))
    t.expect_scatter([2], 1, 1, 0, 0)
    t.expect_scatter([0], 1, 0, 0, 0)
    t.expect_scatter([1], 1, 0, 1, 0)


def test_nasty_chunk_size():
    """AllreduceSpec.scala:240-284 — thReduce .9 -> 1 arrival, thComplete .8 -> 3 of 4 chunks."""
    t = Harness(size=6)
    t.init(t.peers_as_probe(2), 2, 0, 0.9, 0.8, 5, 6, 2)
    t.tell(h.StartAllreduce(0))
    t.expect_scatter([0, 1], 0, 0, 0, 0)
    t.expect_scatter([2], 0, 0, 1, 0)
    t.expect_scatter([3, 4], 0, 1, 0, 0)
    t.expect_scatter([5], 0, 1, 1, 0)
    t.tell(S([0, 1], 0, 0, 0, 0))
    t.tell(S([2], 0, 0, 1, 0))
    t.tell(S([0, 1], 1, 0, 0, 0))
    t.tell(S([2], 1, 0, 1, 0))
    t.expect_reduce([0, 1], 0, 0, 0, 0, 1)
    t.expect_reduce([0, 1], 0, 1, 0, 0, 1)
    t.expect_reduce([2], 0, 0, 1, 0, 1)
    t.expect_reduce([2], 0, 1, 1, 0, 1)
    t.tell(R([0, 2], 0, 0, 0, 0, 1))
    t.tell(R([4], 0, 0, 1, 0, 1))
    t.tell(R([6, 8], 1, 0, 0, 0, 1))
    t.expect_complete(0, 0)
    t.tell(R([10], 1, 0, 1, 0, 1))
    t.expect_no_msg()
    assert t.outputs == [([0, 2, 4, 6, 8, 0], [1, 1, 1, 1, 1, 0], 0)]


def test_nasty_chunk_size_contd():
    """AllreduceSpec.scala:286-349 — int(0.7f*3)=2 arrivals reduce; int(0.7f*9)=6 chunks complete."""
    t = Harness(size=9)
    t.init(t.peers_as_probe(3), 3, 0, 0.7, 0.7, 5, 9, 1)
    t.tell(h.StartAllreduce(0))
    for d in range(3):
        for c in range(3):
            t.expect_scatter([3 * d + c], 0, d, c, 0)
    for s in range(3):
        for c in range(3):
            t.tell(S([float(c)], s, 0, c, 0))
    for c, v in [(0, 0), (1, 2), (2, 4)]:
        for d in range(3):
            t.expect_reduce([v], 0, d, c, 0, 2)
    t.expect_no_msg()  # the third arrival is stored, never rebroadcast
    vals = [0, 3, 6, 9, 12, 15, 18]
    for k, v in enumerate(vals):
        t.tell(R([float(v)], k // 3, 0, k % 3, 0, 2))
        if k < 5:
            t.expect_no_msg()
    t.expect_complete(0, 0)
    t.tell(R([21.0], 2, 0, 1, 0, 2))
    t.tell(R([24.0], 2, 0, 2, 0, 2))
    t.expect_no_msg()
    assert t.outputs == [([0, 3, 6, 9, 12, 15, 0, 0, 0], [2, 2, 2, 2, 2, 2, 0, 0, 0], 0)]


def test_multi_round():
    """AllreduceSpec.scala:351-385 — 10 rounds, thReduce .8 (3 of 4), thComplete .5 (2 of 4)."""
    t = Harness()
    t.init(t.peers_as_probe(4), 4, 0, 0.8, 0.5, 5, 8, 2)
    for i in range(10):
        t.tell(h.StartAllreduce(i))
        for d in range(4):
            t.expect_scatter([2 * d + i, 2 * d + 1 + i], 0, d, 0, i)
        for s in range(4):
            t.tell(S([0.0 + i, 1.0 + i], s, 0, 0, i))
        for d in range(4):
            t.expect_reduce([3 * i, 3 + 3 * i], 0, d, 0, i, 3)
        t.tell(R([1, 2], 0, 0, 0, i, 3))
        t.tell(R([1, 2], 1, 0, 0, i, 3))
        t.expect_complete(0, i)
        t.tell(R([1, 2], 2, 0, 0, i, 3))
        t.tell(R([1, 2], 3, 0, 0, i, 3))
        t.expect_no_msg()
    assert t.worker.round == 10


def test_multi_round_v2():
    """AllreduceSpec.scala:387-422 — P=2, thReduce .6 (1 of 2), thComplete .8 (3 of 4)."""
    t = Harness()
    t.init(t.peers_as_probe(2), 2, 0, 0.6, 0.8, 5, 8, 2)
    for i in range(10):
        t.tell(h.StartAllreduce(i))
        t.expect_scatter([0 + i, 1 + i], 0, 0, 0, i)
        t.expect_scatter([2 + i, 3 + i], 0, 0, 1, i)
        t.expect_scatter([4 + i, 5 + i], 0, 1, 0, i)
        t.expect_scatter([6 + i, 7 + i], 0, 1, 1, i)
        t.tell(S([0 + i, 1 + i], 0, 0, 0, i))
        t.tell(S([2 + i, 3 + i], 0, 0, 1, i))
        t.tell(S([10 + i, 11 + i], 1, 0, 0, i))
        t.tell(S([12 + i, 13 + i], 1, 0, 1, i))
        t.expect_reduce([i, 1 + i], 0, 0, 0, i, 1)
        t.expect_reduce([i, 1 + i], 0, 1, 0, i, 1)
        t.expect_reduce([2 + i, 3 + i], 0, 0, 1, i, 1)
        t.expect_reduce([2 + i, 3 + i], 0, 1, 1, i, 1)
        t.tell(R([1, 2], 0, 0, 0, i, 1))
        t.tell(R([1, 2], 0, 0, 1, i, 1))
        t.tell(R([1, 2], 1, 0, 0, i, 1))
        t.expect_complete(0, i)
        t.tell(R([1, 2], 1, 0, 1, i, 1))
        t.expect_no_msg()


def test_missed_scatter():
    """AllreduceSpec.scala:424-459 — reduce only when exactly 3 of 4 arrived."""
    t = Harness(size=4)
    t.init(t.peers_as_probe(4), 4, 0, 0.75, 0.75, 5, 4, 2)
    t.tell(h.StartAllreduce(0))
    for d in range(4):
        t.expect_scatter([d], 0, d, 0, 0)
    t.tell(S([0], 0, 0, 0, 0))
    t.expect_no_msg()
    t.tell(S([2], 1, 0, 0, 0))
    t.expect_no_msg()
    t.tell(S([4], 2, 0, 0, 0))
    t.tell(S([6], 3, 0, 0, 0))
    for d in range(4):
        t.expect_reduce([6], 0, d, 0, 0, 3)
    t.tell(R([12], 0, 0, 0, 0, 3))
    t.tell(R([11], 1, 0, 0, 0, 3))
    t.tell(R([10], 2, 0, 0, 0, 3))
    t.expect_complete(0, 0)
    t.tell(R([9], 3, 0, 0, 0, 3))
    t.expect_no_msg()


def test_future_scatter():
    """AllreduceSpec.scala:461-513 — round 1 reduces before round 0; a duplicate contribution
    overwrites and bumps the count without re-triggering."""
    t = Harness(size=4)
    t.init(t.peers_as_probe(4), 4, 0, 0.75, 0.75, 5, 4, 2)
    t.tell(h.StartAllreduce(0))
    for d in range(4):
        t.expect_scatter([d], 0, d, 0, 0)
    t.tell(S([2], 1, 0, 0, 0))
    t.tell(S([4], 2, 0, 0, 0))
    t.tell(R([11], 1, 0, 0, 0, 3))
    t.tell(R([10], 2, 0, 0, 0, 3))
    t.tell(h.StartAllreduce(1))
    t.tell(S([2], 1, 0, 0, 1))
    t.tell(S([4], 2, 0, 0, 1))
    t.tell(S([6], 3, 0, 0, 1))
    for d in range(4):
        t.expect_scatter([d + 1], 0, d, 0, 1)
    for d in range(4):
        t.expect_reduce([12], 0, d, 0, 1, 3)
    t.tell(S([0], 3, 0, 0, 0))
    t.tell(S([6], 3, 0, 0, 0))  # duplicate: stored, no second reduce
    for d in range(4):
        t.expect_reduce([6], 0, d, 0, 0, 3)
    t.expect_no_msg()
    t.tell(R([9], 3, 0, 0, 0, 3))
    t.expect_complete(0, 0)
    t.tell(R([11], 1, 0, 0, 1, 3))
    t.tell(R([10], 2, 0, 0, 1, 3))
    t.tell(R([9], 3, 0, 0, 1, 3))
    t.expect_complete(0, 1)


def test_missed_reduce():
    """AllreduceSpec.scala:515-548 — thReduce 1, thComplete .75, one big chunk."""
    t = Harness(size=4)
    t.init(t.peers_as_probe(4), 4, 0, 1.0, 0.75, 5, 4, 100)
    t.tell(h.StartAllreduce(0))
    for d in range(4):
        t.expect_scatter([d], 0, d, 0, 0)
    for s in range(4):
        t.tell(S([2.0 * s], s, 0, 0, 0))
    for d in range(4):
        t.expect_reduce([12], 0, d, 0, 0, 4)
    t.tell(R([12], 0, 0, 0, 0, 4))
    t.expect_no_msg()
    t.tell(R([11], 1, 0, 0, 0, 4))
    t.expect_no_msg()
    t.tell(R([10], 2, 0, 0, 0, 4))
    t.expect_complete(0, 0)


def test_delayed_future_reduce():
    """AllreduceSpec.scala:550-599 — two rounds interleaved: Complete(0) then Complete(1)."""
    t = Harness(size=4)
    t.init(t.peers_as_probe(4), 4, 0, 0.75, 0.75, 5, 4, 100)
    t.tell(h.StartAllreduce(0))
    for d in range(4):
        t.expect_scatter([d], 0, d, 0, 0)
    t.tell(S([2], 1, 0, 0, 0))
    t.tell(S([4], 2, 0, 0, 0))
    t.tell(S([6], 3, 0, 0, 0))
    for d in range(4):
        t.expect_reduce([12], 0, d, 0, 0, 3)
    t.tell(h.StartAllreduce(1))
    t.tell(S([3], 1, 0, 0, 1))
    t.tell(S([5], 2, 0, 0, 1))
    t.tell(S([7], 3, 0, 0, 1))
    for d in range(4):
        t.expect_scatter([d + 1], 0, d, 0, 1)
    for d in range(4):
        t.expect_reduce([15], 0, d, 0, 1, 3)
    t.tell(R([11], 1, 0, 0, 0, 3))
    t.tell(R([11], 1, 0, 0, 1, 3))
    t.tell(R([10], 2, 0, 0, 0, 3))
    t.tell(R([10], 2, 0, 0, 1, 3))
    t.tell(R([9], 3, 0, 0, 0, 3))
    t.tell(R([9], 3, 0, 0, 1, 3))
    t.expect_complete(0, 0)
    t.expect_complete(0, 1)


def _expect_basic_scatter(t, i):
    for d in range(4):
        t.expect_scatter([2 * d + i, 2 * d + 1 + i], 0, d, 0, i)


def test_simple_catchup():
    """AllreduceSpec.scala:605-630,750-767 — round r-(maxLag+1) is force-completed with the
    partial sum 7*(i+1), count 3."""
    t = Harness()
    max_lag = 5
    t.init(t.peers_as_probe(4), 4, 0, 1, 1, max_lag, 8, 2)
    for i in range(6):
        t.tell(h.StartAllreduce(i))
        _expect_basic_scatter(t, i)
        for s, k in [(1, 1.0), (2, 2.0), (3, 4.0)]:
            t.tell(S([k * (i + 1), k * (i + 1)], s, 0, 0, i))
        for s in (1, 2, 3):
            t.tell(R([12.0, 12.0], s, 0, 0, i, 4))
    for catchup in (6, 7, 8):
        t.tell(h.StartAllreduce(catchup))
        done = catchup - (max_lag + 1)
        for d in range(4):
            t.expect_reduce([7.0 * (done + 1)] * 2, 0, d, 0, done, 3)
        t.expect_complete(0, done)
        _expect_basic_scatter(t, catchup)
    assert t.outputs[0] == ([0, 0, 12, 12, 12, 12, 12, 12], [0, 0, 4, 4, 4, 4, 4, 4], 0)


def test_cold_catchup():
    """AllreduceSpec.scala:632-656 — Start(10) on a fresh worker, maxLag 5."""
    t = Harness()
    t.init(t.peers_as_probe(4), 4, 0, 1, 1, 5, 8, 2)
    t.tell(h.StartAllreduce(10))
    for i in range(5):
        for d in range(4):
            t.expect_reduce([0, 0], 0, d, 0, i, 0)
        t.expect_complete(0, i)
    for i in range(11):
        _expect_basic_scatter(t, i)
    t.expect_no_msg()
    assert [o[2] for o in t.outputs] == [0, 1, 2, 3, 4]
    assert all(o[0] == [0.0] * 8 and o[1] == [0] * 8 for o in t.outputs)


def test_multi_round_v3_out_of_order_completion():
    """AllreduceSpec.scala:662-735 — Complete(0,1) is emitted before Complete(0,0); round ends at 2."""
    t = Harness(size=9)
    t.init(t.peers_as_probe(3), 3, 0, 0.75, 0.75, 5, 9, 2)
    t.tell(h.StartAllreduce(0))
    for d in range(3):
        t.expect_scatter([3 * d, 3 * d + 1], 0, d, 0, 0)
        t.expect_scatter([3 * d + 2], 0, d, 1, 0)
    for s in range(3):
        t.tell(S([0, 1], s, 0, 0, 0))
    for s in range(3):
        t.tell(S([2], s, 0, 1, 0))
    for d in range(3):
        t.expect_reduce([0, 2], 0, d, 0, 0, 2)
    for d in range(3):
        t.expect_reduce([4], 0, d, 1, 0, 2)
    t.tell(h.StartAllreduce(1))
    t.tell(S([10, 11], 1, 0, 0, 1))
    t.tell(S([12], 1, 0, 1, 1))
    t.tell(S([10, 11], 2, 0, 0, 1))
    t.tell(S([12], 2, 0, 1, 1))
    for d in range(3):
        t.expect_scatter([3 * d + 1, 3 * d + 2], 0, d, 0, 1)
        t.expect_scatter([3 * d + 3], 0, d, 1, 1)
    for d in range(3):
        t.expect_reduce([20, 22], 0, d, 0, 1, 2)
    for d in range(3):
        t.expect_reduce([24], 0, d, 1, 1, 2)
    t.tell(R([11, 11], 1, 0, 0, 0, 2))
    t.tell(R([11], 1, 0, 1, 1, 2))
    t.tell(R([11, 11], 1, 0, 0, 1, 2))
    t.tell(R([11], 1, 0, 1, 0, 2))
    t.tell(R([11, 11], 2, 0, 0, 0, 2))
    t.tell(R([11], 2, 0, 1, 1, 2))
    t.expect_no_msg()
    t.tell(R([11, 11], 2, 0, 0, 1, 2))
    t.expect_complete(0, 1)
    assert t.worker.round == 0
    t.tell(R([11], 2, 0, 1, 0, 2))
    t.expect_complete(0, 0)
    assert t.worker.round == 2


# ---- behaviours beyond the reference spec ----------------------------------------------------

def test_threshold_arithmetic_is_fp32():
    """Golden vector B9: fp32 multiply, then truncate."""
    assert [h.threshold_count(a, b) for a, b in [(0.7, 10), (0.7, 9), (0.8, 6), (0.8, 4), (0.75, 8)]] == [7, 6, 4, 3, 6]
    assert int(float(np.float32(0.7)) * 10) == 6  # what widening to double first would give


def test_wrong_size_input_is_reported_not_fatal():
    """AllreduceWorker.scala:197-204,287-299 — a source of the wrong size raises inside the handler,
    which is caught and recorded; the engine stays alive."""
    t = Harness(source=basic_source(5), size=5)
    t.init(t.peers_as_probe(2), 2, 0, 1, 1, 1, 8, 2)
    t.tell(h.StartAllreduce(0))
    t.expect_no_msg()
    assert any("different from initialization time" in e for e in t.worker.errors)


def test_oversized_reduce_block_rejected():
    """AllreduceWorker.scala:150-154."""
    t = Harness()
    t.init(t.peers_as_probe(4), 4, 0, 1, 1, 5, 8, 2)
    t.tell(h.StartAllreduce(0))
    t.fifo.clear()
    t.tell(R([1, 2, 3], 1, 0, 0, 0, 4))
    t.tell(R([1, 2], 1, 3, 0, 0, 4))
    assert any("larger than expected" in e for e in t.worker.errors)
    assert any("incorrectly routed" in e for e in t.worker.errors)


def test_partition_with_empty_trailing_block_and_strict_mode():
    """Golden vector B10: n=9, P=4 crashes the reference's partitioner; strict mode reports the same
    error, the default gives the last owner an empty block."""
    opts = h.WorkerOptions()
    opts.strict_partition = True
    t = Harness(size=9, options=opts)
    t.init(t.peers_as_probe(4), 4, 0, 1, 1, 1, 9, 2)
    assert t.worker.id == -1 and t.worker.errors

    t = Harness(size=9)
    t.init(t.peers_as_probe(4), 4, 0, 1, 1, 1, 9, 2)
    t.tell(h.StartAllreduce(0))
    got = [(m.dest_id, m.chunk_id, m.value.tolist()) for m in t.fifo]
    assert got == [(0, 0, [0, 1]), (0, 1, [2]), (1, 0, [3, 4]), (1, 1, [5]), (2, 0, [6, 7]), (2, 1, [8])]
    assert t.worker.complete_required == 6


def test_messages_before_init_are_replayed():
    """AllreduceWorker.scala:95-97 — a worker that is not initialised yet keeps the message for later."""
    t = Harness()
    t.tell(h.StartAllreduce(0))
    t.expect_no_msg()
    t.init(t.peers_as_probe(4), 4, 0, 1, 1, 5, 8, 2)
    _expect_basic_scatter(t, 0)


def test_absurd_round_numbers_are_dropped_not_replayed():
    """Not in the reference (its catch-up loop would run for as many rounds as the message names): a Start, Scatter
    or Reduce naming a round more than `max_round_jump` ahead is recorded as an error and dropped; the worker
    keeps working.  A large but plausible jump still catches up."""
    hs = Harness(size=8)
    hs.init(hs.peers_as_probe(2, real=0), 2, 0, 1.0, 1.0, 1, 8, 2)
    hs.tell(h.StartAllreduce(0))
    while hs.fifo:
        hs.next()
    before = len(hs.worker.errors)
    hs.tell(h.StartAllreduce(2 ** 30))
    hs.tell(h.ScatterBlock(np.ones(2, dtype=np.float32), 1, 0, 0, 2 ** 30))
    hs.tell(h.ReduceBlock(np.ones(2, dtype=np.float32), 1, 0, 0, 2 ** 30, 2))
    assert len(hs.worker.errors) == before + 3 and all("rounds beyond" in e for e in hs.worker.errors[before:])
    assert hs.worker.max_round == 0 and hs.bus.pending() == 0
    hs.expect_no_msg()
    hs.tell(h.StartAllreduce(500))           # far, but legitimate: force-completes 0..498 and scatters up to 500
    assert hs.worker.max_round == 500 and hs.worker.round == 499
