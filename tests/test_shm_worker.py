"""The `shm` backend end to end: AllreduceMaster over TCP + one ShmWorker per worker, data plane = the device protocol
in POSIX shared memory.  In-process here (threads attach the same segment) and as separate worker processes through
the CLI, like the reference's one JVM per worker."""
import os
import subprocess
import sys
import time

import numpy as np
import pytest

from akka_allreduce_b200 import AllReduceInput, AllreduceConfig
from akka_allreduce_b200.master import AllreduceMaster
from akka_allreduce_b200.worker import create_worker

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _cfg(P, n, mc, rounds, th=(1.0, 1.0, 1.0), max_lag=1, gate="device", timeout_ms=2000.0):
    cfg = AllreduceConfig()
    cfg.cluster.port = 0
    cfg.worker.totalSize, cfg.worker.maxLag = P, max_lag
    cfg.data.dataSize, cfg.data.maxChunkSize, cfg.data.maxRound = n, mc, rounds
    cfg.thresholds.thAllreduce, cfg.thresholds.thReduce, cfg.thresholds.thComplete = th
    cfg.device.backend, cfg.device.gate, cfg.device.timeoutMs = "shm", gate, timeout_ms
    return cfg


@pytest.mark.parametrize("gate", ["device", "host"])
def test_shm_backend_python_callbacks(gate):
    """4 workers, the reference's manual configuration (dataSize 778, maxChunkSize 3, maxLag 3), thresholds 1: every
    round on every worker is the exact sum with count 4, under both gates (in shared memory / the coordinator's)."""
    P, n = 4, 778
    cfg = _cfg(P, n, 3, 60, max_lag=3, gate=gate)
    master = AllreduceMaster(cfg)
    port = master.start()
    outs = [dict() for _ in range(P)]
    workers = []
    for i in range(P):
        def source(req):
            return AllReduceInput(np.arange(n, dtype=np.float32) + req.iteration)

        def sink(out, i=i):
            outs[i][out.iteration] = (np.array(out.data), np.array(out.count))

        w = create_worker(source, sink, cfg, master_port=port, shm_name=f"/aar_test_{os.getpid()}_{gate}")
        w.start()
        workers.append(w)
    try:
        assert master.wait(60.0), f"master stuck at round {master.round}"
        assert all(w.wait(30.0) for w in workers)
    finally:
        for w in workers:
            w.stop()
        master.stop()
    assert sorted(w.id for w in workers) == [0, 1, 2, 3]
    for o in outs:
        assert sorted(o) == list(range(61))
        for it, (d, c) in o.items():
            assert (d == P * (np.arange(n, dtype=np.float32) + it)).all() and (c == P).all()
    assert sum(w.timeouts for w in workers) == 0


def test_shm_backend_thresholds_keep_going_without_a_dead_worker():
    """One of 4 workers stops mid-run; thresholds 0.75 keep the rest going (their sums are over the ranks the counts
    name), the master drops the dead one."""
    P, n = 4, 64
    cfg = _cfg(P, n, 8, 150, th=(0.75, 0.75, 0.75), max_lag=2, timeout_ms=200.0)
    master = AllreduceMaster(cfg)
    port = master.start()
    bad, done = [], [0] * P
    workers = []
    for i in range(P):
        def source(req, i=i):
            return AllReduceInput(np.full(n, float(1 << i), dtype=np.float32))

        def sink(out, i=i):
            done[i] = max(done[i], out.iteration)
            bits = np.array(out.data).astype(np.uint32)
            pop = np.array([bin(int(b)).count("1") for b in bits])
            if not (pop == np.array(out.count)).all():
                bad.append((i, out.iteration))

        w = create_worker(source, sink, cfg, master_port=port, shm_name=f"/aar_test_{os.getpid()}_dead")
        w.start()
        workers.append(w)
    try:
        t0 = time.time()
        while master.round < 10 and time.time() - t0 < 20:
            time.sleep(0.005)
        victim = next(w for w in workers if w.id == 3)      # not the segment's creator
        victim.stop()
        assert master.wait(60.0), f"master stuck at round {master.round}"
        assert master.alive == 3
    finally:
        for w in workers:
            w.stop()
        master.stop()
    assert not bad, bad[:5]
    assert sum(d >= 145 for d in done) >= 3



def test_shm_backend_dead_worker_with_full_thresholds_costs_no_timeouts():
    """Every data threshold at 1 and a 5 s wait bound: without the dead-peer mask each round after a worker's death
    would sit out the clock.  The coordinator announces the removal (refreshed InitWorkers), the workers turn the ids
    missing from the peer map into the mask, and the remaining rounds run at full speed with count 3 and zeros /
    count 0 on the dead owner's block — the GPU backend's behaviour (tests/multi_gpu_kill_worker.py) on the CPU."""
    P, n, rounds = 4, 64, 400
    cfg = _cfg(P, n, 8, rounds, th=(0.5, 1.0, 1.0), max_lag=2, timeout_ms=5000.0)
    master = AllreduceMaster(cfg)
    port = master.start()
    seen = [dict() for _ in range(P)]
    workers = []
    for i in range(P):
        def source(req, i=i):
            return AllReduceInput(np.full(n, float(1 << i), dtype=np.float32))

        def sink(out, i=i):
            seen[i][out.iteration] = (time.perf_counter(), np.array(out.data), np.array(out.count))

        w = create_worker(source, sink, cfg, master_port=port, shm_name=f"/aar_test_{os.getpid()}_mask")
        w.start()
        workers.append(w)
    try:
        t0 = time.time()
        while master.round < 20 and time.time() - t0 < 20:
            time.sleep(0.002)
        victim = next(w for w in workers if w.id == 3)      # not the segment's creator
        victim.stop()
        t_kill = time.perf_counter()
        assert master.wait(30.0), f"master stuck at round {master.round}"   # 380 rounds at 5 s each would never make it
        assert master.alive == 3
    finally:
        for w in workers:
            w.stop()
        master.stop()
    survivors = [w for w in workers if w.id != 3]
    for w in survivors:
        o = seen[workers.index(w)]
        assert max(o) == rounds
        tail = [o[r] for r in range(rounds - 50, rounds + 1)]
        for _, d, c in tail:
            assert (c[:48] == 3).all() and (d[:48] == 7.0).all()      # ranks 0-2 on the blocks they own
            assert (c[48:] == 0).all() and (d[48:] == 0.0).all()      # nobody reduces the dead owner's block
        # at most a few rounds around the death may have sat out the 5 s clock
        slow = [r for r in range(1, rounds + 1) if r in o and r - 1 in o and o[r][0] - o[r - 1][0] > 2.0]
        assert len(slow) <= 3, slow
        assert o[rounds][0] - t_kill < 25.0


def test_shm_backend_cli_processes():
    """Master + 3 worker PROCESSES through the launchers (`--backend shm`), built-in ramp source and assert sink."""
    env = dict(os.environ, PYTHONPATH=ROOT)
    port = 29000 + os.getpid() % 2000
    m = subprocess.Popen([sys.executable, "-m", "akka_allreduce_b200.master", str(port), "3", "300", "16",
                          "--maxRound", "200", "--thComplete", "1.0", "--maxLag", "2"], env=env,
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    ws = [subprocess.Popen([sys.executable, "-m", "akka_allreduce_b200.worker", "0", "300", "--master-port", str(port),
                            "--checkpoint", "50", "--assert-multiple", "3", "--backend", "shm"], env=env,
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for _ in range(3)]
    try:
        outs = [w.communicate(timeout=90)[0] for w in ws]
        mout = m.communicate(timeout=30)[0]
    finally:
        for p in ws + [m]:
            if p.poll() is None:
                p.kill()
    assert m.returncode == 0, mout
    for w, o in zip(ws, outs):
        assert w.returncode == 0, o
        assert "MBytes/sec" in o and "----Data output at #200" in o and "assertion failed" not in o


def test_shm_backend_replacement_worker_rejoins():
    """Elastic path on shared memory: one of 4 workers leaves, a replacement registers, the coordinator hands it the
    vacated id and the current round, it attaches to the running job's segment, starts at the job's front and
    contributes again: worker i sends 2^i, so the sums the survivors receive show the vacated id's bit coming back."""
    P, n = 4, 64
    cfg = _cfg(P, n, 8, 2000, th=(0.75, 0.75, 0.75), max_lag=2, timeout_ms=200.0)
    master = AllreduceMaster(cfg)
    port = master.start()
    name = f"/aar_test_{os.getpid()}_rejoin"
    seen = {}

    def make(tag):
        box = {}

        def source(req):
            time.sleep(0.001)                        # >= 2 s of job: time to leave and come back, even on a busy box
            return AllReduceInput(np.full(n, float(1 << box["w"].id), dtype=np.float32))

        def sink(out, tag=tag):
            bits = 0
            for v in np.unique(np.array(out.data)):
                bits |= int(v)
            seen.setdefault(tag, []).append((out.iteration, bits))

        w = create_worker(source, sink, cfg, master_port=port, shm_name=name)
        box["w"] = w
        w.start()
        return w

    workers = [make(i) for i in range(P)]
    try:
        t0 = time.time()
        while master.round < 30 and time.time() - t0 < 20:
            time.sleep(0.002)
        victim = next(w for w in workers if w.id not in (-1, 0))     # not the segment's creator
        vid = victim.id
        victim.stop()
        t0 = time.time()
        while master.alive != 3 and time.time() - t0 < 10:
            time.sleep(0.005)
        assert master.alive == 3
        repl = make("replacement")
        workers.append(repl)
        assert master.wait(60), f"master stuck at round {master.round}"
        assert repl.id == vid and master.alive == 4
    finally:
        for w in workers:
            w.stop()
        master.stop()
    rounds = [r for r, _ in seen["replacement"]]
    assert rounds and min(rounds) >= 30
    survivor = next(i for i in range(P) if seen.get(i) and workers[i] is not victim)
    gone = [r for r, bits in seen[survivor] if not bits & (1 << vid)]
    back = [r for r, bits in seen[survivor] if r > min(rounds) and bits & (1 << vid)]
    assert gone and back, (len(gone), len(back))          # its contribution stopped, then came back
