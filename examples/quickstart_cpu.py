#!/usr/bin/env python
"""Master + 3 workers in one process on the CPU backend, with the reference's callback contract.

    python examples/quickstart_cpu.py

Each worker's source returns `iteration + worker_index` everywhere; with thresholds of 1 every sink must see
the sum over workers and a contributor count of 3 for every element.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np

from akka_allreduce_b200 import AllReduceInput, AllreduceConfig
from akka_allreduce_b200.master import AllreduceMaster
from akka_allreduce_b200.worker import AllreduceWorker

WORKERS, N, ROUNDS = 3, 1000, 50

cfg = AllreduceConfig()
cfg.cluster.port = 0                      # any free port
cfg.worker.totalSize, cfg.worker.maxLag = WORKERS, 2
cfg.data.dataSize, cfg.data.maxChunkSize, cfg.data.maxRound = N, 128, ROUNDS
cfg.thresholds.thAllreduce = cfg.thresholds.thReduce = cfg.thresholds.thComplete = 1.0

master = AllreduceMaster(cfg)
port = master.start()
results = {}


def make_worker(index):
    def source(request):                  # AllReduceInputRequest(iteration) -> AllReduceInput(data)
        return AllReduceInput(np.full(N, request.iteration + index, dtype=np.float32))

    def sink(output):                     # AllReduceOutput(data, count, iteration)
        results[(index, output.iteration)] = (float(output.data[0]), int(output.count.min()))

    w = AllreduceWorker(source, sink, master_port=port)
    w.start()
    return w


workers = [make_worker(i) for i in range(WORKERS)]
assert master.wait(60)
for w in workers:
    w.wait(10)
    w.stop()
master.stop()

for it in (0, ROUNDS):
    expect = float(sum(it + i for i in range(WORKERS)))
    assert all(results[(i, it)] == (expect, WORKERS) for i in range(WORKERS)), results[(0, it)]
print(f"ok: {WORKERS} workers x {ROUNDS + 1} rounds of {N} floats; round {ROUNDS} sum = {results[(0, ROUNDS)][0]}, count = {WORKERS}")
