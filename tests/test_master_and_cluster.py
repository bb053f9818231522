"""Master gate and whole-cluster runs on the in-process bus (CPU, deterministic)."""
import numpy as np
import pytest

from akka_allreduce_b200 import host as h


def cfg(**kw):
    c = h.MasterConfig()
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def test_master_waits_for_all_then_inits_and_starts():
    """AllreduceMaster.scala:36-44,66-89 — ids by arrival order, one InitWorkers each, then Start(0)."""
    bus = h.LocalBus()
    ep, master = bus.add_master(cfg(total_workers=3, data_size=9, max_chunk_size=2, max_lag=2, max_round=5))
    probes = [bus.add_probe() for _ in range(3)]
    assert master.member_up(probes[0]) == 0
    assert master.member_up(probes[1]) == 1
    bus.run()
    assert all(not bus.take(p) for p in probes) and master.round == -1
    assert master.member_up(probes[2]) == 2
    bus.run()
    for i, p in enumerate(probes):
        init, start = bus.take(p)
        assert isinstance(init, h.InitWorkers) and init.dest_id == i and init.worker_num == 3
        assert init.workers == {0: probes[0], 1: probes[1], 2: probes[2]} and init.master == ep
        assert (init.data_size, init.max_chunk_size, init.max_lag) == (9, 2, 2)
        assert start == h.StartAllreduce(0)


def test_master_gate_fraction_and_max_round():
    """AllreduceMaster.scala:54-63 — advance at numComplete >= total*thAllreduce; rounds 0..maxRound."""
    bus = h.LocalBus()
    ep, master = bus.add_master(cfg(total_workers=4, th_allreduce=0.75, max_round=2))
    probes = [bus.add_probe() for _ in range(4)]
    for p in probes:
        master.member_up(p)
    bus.run()
    for p in probes:
        bus.take(p)
    bus.send(ep, h.CompleteAllreduce(0, 0)); bus.send(ep, h.CompleteAllreduce(1, 0)); bus.run()
    assert master.round == 0 and master.num_complete == 2
    bus.send(ep, h.CompleteAllreduce(3, 5)); bus.run()          # wrong round: ignored
    assert master.num_complete == 2
    bus.send(ep, h.CompleteAllreduce(2, 0)); bus.run()          # 3 >= 4*0.75
    assert master.round == 1 and master.num_complete == 0
    assert all(bus.take(p) == [h.StartAllreduce(1)] for p in probes)
    bus.send(ep, h.CompleteAllreduce(3, 0)); bus.run()          # late report for round 0: ignored
    assert master.num_complete == 0
    for r in (1, 2):
        for s in range(3):
            bus.send(ep, h.CompleteAllreduce(s, r))
        bus.run()
    assert master.round == 2 and master.finished                # never starts round 3
    assert [m.round for m in bus.take(probes[0])] == [2]


def test_master_drops_terminated_worker_but_keeps_configured_total():
    """AllreduceMaster.scala:46-52,58."""
    bus = h.LocalBus()
    ep, master = bus.add_master(cfg(total_workers=3, th_allreduce=1.0, max_round=10))
    probes = [bus.add_probe() for _ in range(3)]
    for p in probes:
        master.member_up(p)
    bus.run()
    master.terminated(probes[1])
    assert sorted(master.workers) == [0, 2]
    bus.send(ep, h.CompleteAllreduce(0, 0)); bus.send(ep, h.CompleteAllreduce(2, 0)); bus.run()
    assert master.round == 0  # 2 < 3*1.0: the gate still uses the configured total


def run_cluster(P, n, mc, max_lag, max_round, th=(1.0, 1.0, 1.0), source=None):
    bus = h.LocalBus()
    outs = [dict() for _ in range(P)]
    c = cfg(total_workers=P, th_allreduce=th[0], th_reduce=th[1], th_complete=th[2], max_lag=max_lag,
            data_size=n, max_round=max_round, max_chunk_size=mc)
    ep, master = bus.add_master(c)
    src = source or (lambda it: np.arange(n, dtype=np.float32))
    workers = []
    for i in range(P):
        def sink(d, cnt, it, i=i):
            outs[i][it] = (d.copy(), cnt.copy())
        w_ep, w = bus.add_worker(src, sink)
        workers.append(w)
        master.member_up(w_ep)
    bus.run()
    return master, workers, outs


def test_default_smoke_config_two_workers():
    """BASELINE config 1 / AllreduceMaster.scala:98-107: 2 workers, dataSize 10, maxChunkSize 2,
    maxLag 1, maxRound 100, th 1 / 1 / 0.8."""
    master, workers, outs = run_cluster(2, 10, 2, 1, 100, th=(1.0, 1.0, 0.8))
    assert master.finished and master.round == 100
    for o in outs:
        assert sorted(o) == list(range(101))
        for d, c in o.values():
            got = c > 0
            assert (d[got] == 2 * np.arange(10, dtype=np.float32)[got]).all()
            assert (c[got] == 2).all() and got.sum() >= 7  # int(0.8f*6)=4 of the 6 chunks (2,2,1 per block)
    assert all(not w.errors for w in workers)


def test_manual_four_worker_config_assert_multiple():
    """scripts/testAllreduceMaster.sc:7-22 + AllreduceWorker.scala:337-340: dataSize 778, chunk 3,
    maxLag 3, thresholds 1 -> output == 4*input, count == 4."""
    master, workers, outs = run_cluster(4, 778, 3, 3, 30)
    ramp = np.arange(778, dtype=np.float32)
    assert master.finished
    for o in outs:
        assert sorted(o) == list(range(31))
        for d, c in o.values():
            assert (d == 4 * ramp).all() and (c == 4).all()


@pytest.mark.parametrize("P,n,mc", [(3, 9, 2), (4, 9, 2), (8, 1000, 7), (5, 17, 100)])
def test_cluster_thresholds_one_any_shape(P, n, mc):
    master, workers, outs = run_cluster(P, n, mc, 2, 5, source=lambda it: np.arange(n, dtype=np.float32) + it)
    assert master.finished
    for o in outs:
        for it, (d, c) in o.items():
            assert (d == P * (np.arange(n, dtype=np.float32) + it)).all() and (c == P).all()


def test_replacement_worker_joins_running_job():
    """Elastic recovery beyond the reference (a node that joins after round 0 is never
    initialised there).  A worker dies mid-run, thresholds of 0.75 keep rounds going with count 3; a
    replacement registers, inherits the vacated id and block, is initialised AT the current round, and the
    counts go back to 4."""
    P, n, mc = 4, 16, 2
    bus = h.LocalBus()
    c = cfg(total_workers=P, th_allreduce=0.75, th_reduce=0.75, th_complete=0.75, max_lag=2, data_size=n,
            max_round=200, max_chunk_size=mc)
    ep, master = bus.add_master(c)
    outs = [dict() for _ in range(P + 1)]
    src = lambda it: np.ones(n, dtype=np.float32)
    eps, workers = [], []
    for i in range(P):
        w_ep, w = bus.add_worker(src, lambda d, cnt, it, i=i: outs[i].__setitem__(it, (d.copy(), cnt.copy())))
        eps.append(w_ep)
        workers.append(w)
        master.member_up(w_ep)
    while master.round < 20:
        bus.run(200)
    # worker 3 dies: the master drops it; what others still send it becomes dead letters
    master.terminated(eps[3])
    dead = bus.add_probe()
    bus.take(dead)
    died_at = master.round
    while master.round < died_at + 30:
        bus.run(200)
    mid = max(outs[0])
    assert outs[0][mid][1].max() <= 4 and master.round >= died_at + 30
    # a replacement shows up
    new_ep, neww = bus.add_worker(src, lambda d, cnt, it: outs[P].__setitem__(it, (d.copy(), cnt.copy())))
    joined_at = master.round
    assert master.member_up(new_ep) == 3 and sorted(master.workers) == [0, 1, 2, 3]
    bus.run()
    assert master.finished and neww.id == 3
    assert min(outs[P]) >= joined_at            # it started at the current round, not at 0
    # While worker 3 was gone nobody reduced block 3 (elements 12..15): count 0 there on the survivors; once
    # the replacement owns it, it is reduced and broadcast again.  (Counts stay at 3: thReduce 0.75 fires on
    # the 3rd arrival and reports 3, AllreduceSpec "missed scatter".)
    block3 = slice(12, 16)
    gone = [r for r in outs[0] if died_at + 5 < r < joined_at - 3]
    assert gone and all(outs[0][r][1][block3].max() == 0 for r in gone)
    back = [r for r in outs[0] if r > joined_at + 5]
    assert back and any(outs[0][r][1][block3].min() > 0 for r in back)
    for r in back:
        d, cnt = outs[0][r]
        assert (d[cnt > 0] == cnt[cnt > 0]).all()   # sum of ones == contributors
    assert not neww.errors


def test_master_announces_removals_when_asked():
    """Extension for the CUDA backend (MasterConfig.announce_removals): a worker removed from a running job makes the
    master re-send InitWorkers — a peer-map refresh for the survivors (AllreduceWorker.scala:87-89) — so that GPU
    workers can stop pushing to and waiting for the dead rank.  Off by default: the reference sends InitWorkers once."""
    for announce in (False, True):
        bus = h.LocalBus()
        c = cfg(total_workers=3, th_allreduce=0.6, max_round=10)
        c.announce_removals = announce
        ep, master = bus.add_master(c)
        probes = [bus.add_probe() for _ in range(3)]
        for p in probes:
            master.member_up(p)
        bus.run()
        for p in probes:
            bus.take(p)                                  # the initial InitWorkers + StartAllreduce(0)
        master.terminated(probes[1])
        bus.run()
        got = [bus.take(p) for p in probes]
        if not announce:
            assert all(not g for g in got)
        else:
            assert not got[1]                            # nothing goes to the removed worker
            for i in (0, 2):
                (init,) = got[i]
                assert isinstance(init, h.InitWorkers) and sorted(init.workers) == [0, 2] and init.worker_num == 3
                assert init.dest_id == i and init.start_round == 0
        # before the job has started there is nobody to tell
        bus2 = h.LocalBus()
        c2 = cfg(total_workers=3)
        c2.announce_removals = announce
        _, m2 = bus2.add_master(c2)
        p0 = bus2.add_probe()
        m2.member_up(p0)
        m2.terminated(p0)
        bus2.run()
        assert not bus2.take(p0)
