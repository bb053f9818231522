#!/usr/bin/env python
"""Regenerates profiles/protocol_model.md: the configurations of the device flag protocol that were searched
(akka_allreduce_b200/utils/protocol_model.py) and what the search found.  python scripts/protocol_model_table.py > /tmp/t.md"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import time
from akka_allreduce_b200.utils.protocol_model import ModelConfig, check
rows = [
 ("two-shot, thresholds 1, maxLag 1, no gate (flagship)", ModelConfig(P=2, L=2, rounds=4), {}),
 ("same, 3 ranks", ModelConfig(P=3, L=2, rounds=3), {}),
 ("maxLag 0 with the gate", ModelConfig(P=3, L=1, rounds=2, gate_count=3), {}),
 ("maxLag 0 WITHOUT the gate", ModelConfig(P=2, L=1, rounds=3), {}),
 ("thresholds 0.5, timeouts anywhere", ModelConfig(P=2, L=2, rounds=5, req_reduce=1, req_complete=1, gate_count=1, timeouts=True), {}),
 ("thresholds 2/3, <= 2 timeouts, 3 ranks", ModelConfig(P=3, L=2, rounds=2, req_reduce=2, req_complete=2, gate_count=2, timeouts=True, giveups=2), {}),
 ("seeded bug: no back-pressure", ModelConfig(P=2, L=2, rounds=4, req_reduce=1, req_complete=1, gate_count=1, timeouts=True, backpressure=False), {}),
 ("thresholds 2/3, no timeouts (liveness)", ModelConfig(P=3, L=2, rounds=3, req_reduce=2, req_complete=2, gate_count=2), {}),
 ("same with skip notices (shipped kernels)", ModelConfig(P=3, L=2, rounds=3, req_reduce=2, req_complete=2, gate_count=2, notices=True), {}),
 ("one slot, no gate, skip notices", ModelConfig(P=2, L=1, rounds=4, notices=True), {}),
 ("a rank dies after round 1; dead-peer mask, thresholds 1, no timeouts (shipped kernels)", ModelConfig(P=3, L=2, rounds=4, dead_rank=2, dies_after=1), {}),
 ("same, one slot with the full gate (needs the skip notices: the relaxed gate no longer implies free slots)", ModelConfig(P=3, L=1, rounds=3, gate_count=3, notices=True, dead_rank=0, dies_after=1), {}),
 ("seeded bug: the removal is never announced", ModelConfig(P=3, L=2, rounds=4, dead_rank=2, dies_after=1, announce=False), {}),
 ("one-shot, 2 slots", ModelConfig(P=3, L=2, rounds=3, oneshot=True), {}),
 ("one-shot, 1 slot, no gate (refused by the ABI)", ModelConfig(P=2, L=1, rounds=3, oneshot=True), {}),
 ("one-shot under timeouts WITHOUT a destination mask (the round-1 kernel: seeded bug)", ModelConfig(P=2, L=2, rounds=4, oneshot=True, timeouts=True), {}),
 ("one-shot under timeouts, shipped kernel (destination mask + notices)", ModelConfig(P=2, L=2, rounds=5, oneshot=True, timeouts=True, oneshot_backpressure=True), {}),
 ("input_view written after round r-1, thReduce 1, thComplete 2/3 (the library hands the view out only with thresholds 1)", ModelConfig(P=3, L=2, rounds=3, early_write=True, req_complete=2), {}),
 ("input_view, thReduce 0.5", ModelConfig(P=2, L=3, rounds=5, early_write=True, req_reduce=1, req_complete=1), {}),
 ("output_view, thresholds 1", ModelConfig(P=3, L=2, rounds=4, output_view=True), {}),
 ("output_view, thReduce 2/3 (now refused)", ModelConfig(P=3, L=2, rounds=4, output_view=True, req_reduce=2), {}),
 ("8 ranks, thresholds 1, random executions", ModelConfig(P=8, L=2, rounds=10), dict(walks=100, seed=1)),
 ("8 ranks, thresholds 0.75, timeouts, random executions", ModelConfig(P=8, L=3, rounds=10, req_reduce=6, req_complete=6, gate_count=6, timeouts=True), dict(walks=200, seed=2)),
 ("8 ranks, skip notices, no timeouts, random executions", ModelConfig(P=8, L=2, rounds=10, req_reduce=6, req_complete=6, gate_count=6, notices=True), dict(walks=200, seed=3)),
]
print("| configuration | search | states / steps | result |\n|---|---|---|---|")
for name, cfg, kw in rows:
    t0 = time.time(); r = check(cfg, **kw)
    res = "safe, live" if r.ok else r.violation.split(":")[0] if r.violation.startswith("stuck") else "VIOLATION: " + r.violation
    if r.ok and cfg.timeouts: res = "safe"
    print(f"| {name} | {'random x%d' % kw['walks'] if kw else 'exhaustive'} | {r.states} | {res} |", flush=True)
