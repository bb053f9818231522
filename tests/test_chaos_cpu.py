"""Long runs of whole clusters on the in-process bus with messages dropped and reordered at random.

The reference has no fault-injection tooling and no property tests; these runs check the
invariants its thresholds are supposed to give under loss: every round is eventually completed everywhere
(by threshold, or by catch-up once a worker is more than maxLag behind), and whatever a sink receives is
self-consistent — each worker's message is the constant 2^id, so a reduced value is the bitmask of its
contributors and its popcount must equal the reported count; count 0 must come with value 0.
"""
import numpy as np
import pytest

from akka_allreduce_b200 import host as h


def _run(P, n, mc, max_lag, rounds, th, seed, drop, delay):
    bus = h.LocalBus()
    c = h.MasterConfig()
    c.total_workers, c.data_size, c.max_chunk_size, c.max_lag, c.max_round = P, n, mc, max_lag, rounds
    c.th_allreduce, c.th_reduce, c.th_complete = th
    _, master = bus.add_master(c)
    outs = [dict() for _ in range(P)]
    workers = []
    for i in range(P):
        ep, w = bus.add_worker(lambda it, i=i: np.full(n, float(1 << i), dtype=np.float32),
                               lambda d, cnt, it, i=i: outs[i].__setitem__(it, (d.copy(), cnt.copy())))
        workers.append(w)
        master.member_up(ep)
    bus.set_chaos(seed, drop, delay)
    bus.run(1 << 26)
    return bus, master, workers, outs


def _consistent(d, cnt):
    bits = d.astype(np.int64)
    pop = sum(((bits >> b) & 1) for b in range(16))
    return (pop == cnt).all() and (bits.astype(np.float32) == d).all()


@pytest.mark.parametrize("seed", [1, 2, 3])
@pytest.mark.parametrize("max_lag", [1, 4])
def test_thresholds_absorb_loss_and_reordering(seed, max_lag):
    P, n, mc, rounds = 4, 24, 3, 400
    bus, master, workers, outs = _run(P, n, mc, max_lag, rounds, (0.75, 0.75, 0.75), seed, drop=0.03, delay=0.10)
    assert bus.chaos_dropped() > 100
    assert master.finished and master.round == rounds       # the online-nodes gate keeps the job moving
    for i, o in enumerate(outs):
        assert len(o) >= rounds * 0.9                        # stragglers are force-completed, not stuck
        for it, (d, cnt) in o.items():
            assert _consistent(d, cnt), f"worker {i} round {it}"
            assert cnt.max() <= P
    assert all(not w.errors for w in workers)
    partial = sum(int((cnt[cnt > 0] < P).any()) for o in outs for _, cnt in o.values())
    assert partial > 0                                       # losses really did produce partial sums


def test_no_loss_long_run_is_exact_for_every_lag():
    """Thousands of rounds, maxLag 1..4, thresholds 1, reordering only: output == P-fold sum, count == P."""
    for max_lag in (1, 2, 3, 4):
        P, n = 3, 10
        bus, master, workers, outs = _run(P, n, 2, max_lag, 1500, (1.0, 1.0, 1.0), seed=5 + max_lag, drop=0.0, delay=0.2)
        assert master.finished
        for o in outs:
            assert sorted(o) == list(range(1501))
            for d, cnt in o.values():
                assert (d == 7.0).all() and (cnt == P).all()


def test_engine_survives_malformed_messages():
    """Wire-level fuzz of one worker engine: ids, chunk numbers, rounds and payload sizes that no peer of this job
    could send.  Every message is either handled or recorded in `errors` (the reference's tryCatch,
    AllreduceWorker.scala:287-299); nothing may reach Python as an exception, let alone crash.  The same fuzz was
    run against the ASan build (profiles/host_sanitizers.md)."""
    import random

    from harness import Harness

    rng = random.Random(7)
    recorded = 0
    for _ in range(120):
        P, n, mc, lag = rng.randint(1, 5), rng.randint(1, 40), rng.randint(1, 9), rng.randint(0, 3)
        idx = rng.randrange(P)
        hs = Harness(size=n)
        hs.init(hs.peers_as_probe(P, real=idx), P, idx, rng.choice([0.5, 0.7, 1.0]), rng.choice([0.5, 0.8, 1.0]),
                lag, n, mc)
        for _ in range(rng.randint(5, 40)):
            kind, rnd = rng.random(), rng.randint(-2, 8)
            val = np.asarray([rng.random() for _ in range(rng.randint(0, mc + 3))], dtype=np.float32)
            src, chunk = rng.randint(-1, P + 1), rng.randint(-1, 12)
            dest = rng.choice([idx, idx, idx, rng.randint(-1, P + 1)])
            if kind < 0.2:
                hs.tell(h.StartAllreduce(rnd))
            elif kind < 0.6:
                hs.tell(h.ScatterBlock(val, src, dest, chunk, rnd))
            elif kind < 0.95:
                hs.tell(h.ReduceBlock(val, src, dest, chunk, rnd, rng.randint(-1, P + 2)))
            else:
                hs.tell(h.CompleteAllreduce(src, rnd))
        recorded += len(hs.worker.errors)
    assert recorded > 100          # the fuzz does hit the guarded paths
