#!/usr/bin/env python
"""The device protocol without a GPU: master + 4 workers on the `shm` backend, one of them slow.

    python examples/quickstart_shm.py

Same callbacks as quickstart_cpu.py; only `config.device.backend` differs.  Thresholds are 0.75, so rounds do not
wait for the slowest worker: worker i contributes 2^i everywhere, every output element names its contributors, and
the per-element count says how many there were.
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np

from akka_allreduce_b200 import AllReduceInput, AllreduceConfig
from akka_allreduce_b200.master import AllreduceMaster
from akka_allreduce_b200.worker import create_worker

WORKERS, N, ROUNDS = 4, 4096, 200
POPCOUNT = np.array([bin(v).count("1") for v in range(1 << WORKERS)])

cfg = AllreduceConfig()
cfg.cluster.port = 0
cfg.device.backend = "shm"                # cpu | shm | cuda: the rest of this file would not change
cfg.worker.totalSize, cfg.worker.maxLag = WORKERS, 2
cfg.data.dataSize, cfg.data.maxChunkSize, cfg.data.maxRound = N, 256, ROUNDS
cfg.thresholds.thAllreduce = cfg.thresholds.thReduce = cfg.thresholds.thComplete = 0.75

master = AllreduceMaster(cfg)
port = master.start()
seen = {i: [] for i in range(WORKERS)}


def make_worker(index):
    def source(request):
        if index == WORKERS - 1 and request.iteration % 10 == 0:
            time.sleep(0.002)             # the straggler
        return AllReduceInput(np.full(N, float(1 << index), dtype=np.float32))

    def sink(output):
        named = POPCOUNT[output.data.astype(np.uint32)]          # how many workers each sum names
        assert (named == output.count).all(), "a sum that does not match its count"
        seen[index].append(float(output.count.mean()))

    w = create_worker(source, sink, cfg, master_port=port, shm_name=f"/aar_example_{os.getpid()}")
    w.start()
    return w


workers = [make_worker(i) for i in range(WORKERS)]
assert master.wait(60), f"stuck at round {master.round}"
for w in workers:
    w.wait(10)
    w.stop()
master.stop()
mean = np.mean([np.mean(v) for v in seen.values()])
print(f"ok: {WORKERS} workers x {ROUNDS + 1} rounds of {N} floats on shared memory; "
      f"mean contributors per element {mean:.2f} of {WORKERS}")
