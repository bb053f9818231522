#!/usr/bin/env python
"""Timed CPU baseline: the reference's default plumbing run on this repo's C++ host engine.

The reference's own JVM build cannot run here (no java/sbt), so the stand-in baseline is the CPU backend:
master + N worker *processes* — `cpu`: C++ engine over loopback TCP with the native ramp source and throughput sink;
`shm`: the device protocol in POSIX shared memory, Python round loop —
(BASELINE.json config 1: 2 workers, dataSize 10, maxChunkSize 2; and the 4-worker manual config).
Reports rounds/s and the reference's own metric, MBytes/sec = len*4*rounds/1e6/elapsed per worker.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run(workers: int, data_size: int, max_chunk: int, max_round: int, max_lag: int, th=(1.0, 1.0, 1.0),
        backend: str = "cpu") -> dict:
    from akka_allreduce_b200 import AllreduceConfig
    from akka_allreduce_b200.master import AllreduceMaster

    cfg = AllreduceConfig()
    cfg.cluster.port = 0
    cfg.worker.totalSize, cfg.worker.maxLag = workers, max_lag
    cfg.data.dataSize, cfg.data.maxChunkSize, cfg.data.maxRound = data_size, max_chunk, max_round
    cfg.thresholds.thAllreduce, cfg.thresholds.thReduce, cfg.thresholds.thComplete = th
    master = AllreduceMaster(cfg)
    port = master.start()
    env = dict(os.environ, PYTHONPATH=ROOT)
    t0 = time.perf_counter()
    procs = [subprocess.Popen([sys.executable, "-m", "akka_allreduce_b200.worker", "0", str(data_size), "--master-port",
                               str(port), "--checkpoint", str(max(1, max_round // 4)), "--assert-multiple",
                               str(workers if th == (1.0, 1.0, 1.0) else 0), "--backend", backend], env=env,
                              stdout=subprocess.DEVNULL)
             for _ in range(workers)]
    ok = master.wait(120)
    elapsed = time.perf_counter() - t0
    for p in procs:
        try:
            p.wait(timeout=20)
        except subprocess.TimeoutExpired:
            p.kill()
    master.stop()
    rounds = max_round + 1
    return {"backend": backend, "workers": workers, "dataSize": data_size, "maxChunkSize": max_chunk, "maxLag": max_lag, "rounds": rounds,
            "finished": bool(ok), "seconds_incl_startup": round(elapsed, 3), "rounds_per_s": round(rounds / elapsed, 1),
            "MBytes_per_s_per_worker": round(data_size * 4.0 * rounds / 1e6 / elapsed, 4),
            "worker_exit_codes": [p.returncode for p in procs]}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    rows = []
    for backend in ("cpu", "shm"):   # C++ engine over loopback TCP / the device protocol in POSIX shared memory
        rows += [run(2, 10, 2, 2000, 1, (1.0, 1.0, 0.8), backend),   # reference defaults (AllreduceMaster.scala:98-107)
                 run(4, 778, 3, 1000, 3, backend=backend),           # scripts/testAllreduceMaster.sc
                 run(4, 1 << 20, 1 << 16, 50, 1, backend=backend)]   # a 4 MB message: the CPU data plane's rate
    for r in rows:
        print(json.dumps(r))
    if a.out:
        with open(a.out, "w") as f:
            for r in rows:
                f.write(json.dumps(r) + "\n")
