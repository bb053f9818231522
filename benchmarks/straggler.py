#!/usr/bin/env python
"""Rounds/sec and correctness under thresholds < 1 with injected stragglers (BASELINE.json config 3).

    python benchmarks/straggler.py --emulate 8                 # one GPU: in-kernel delays/mutes, correctness
    torchrun --nproc-per-node 8 benchmarks/straggler.py        # real ranks: AllreduceMaster + DeviceWorker,
                                                               # straggler ranks sleep in their data source

Config: thReduce = thComplete = 0.75, online-nodes threshold thAllreduce = 0.75 (6 of 8), maxLag = 4.
Every rank's message is the constant 2^rank, so a reduced value IS the bitmask of its contributors:
for every element the number of set bits must equal the reported count (partial sums numerically
correct against `count`), and a count of 0 must come with a value of 0.
"""
from __future__ import annotations

import argparse
import json
import os
import random
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch


def check(data: torch.Tensor, count: torch.Tensor) -> bool:
    bits = data.to(torch.int64)
    pop = torch.zeros_like(bits)
    for b in range(16):
        pop += (bits >> b) & 1
    return bool((pop == count.to(torch.int64)).all()) and bool((data == bits.to(data.dtype)).all())


def emulated(args) -> dict:
    from akka_allreduce_b200.ops.fused_allreduce import EmulatedCluster, Injection

    P, n, mc = args.emulate, args.size, args.chunk
    cl = EmulatedCluster(P, n, mc, max_lag=4, th_reduce=0.75, th_complete=0.75, th_allreduce=0.75, timeout_ms=args.timeout_ms)
    srcs = [torch.full((n,), float(1 << r), device="cuda") for r in range(P)]
    rng = random.Random(0)
    ok, hist = True, {}
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for r in range(args.rounds):
        late = rng.sample(range(P), rng.choice([0, 1, 2]))
        inj = Injection(delay_ns={k: args.delay_us * 1000 for k in late[:1]}, mute=sum(1 << k for k in late[1:]))
        outs, counts = cl.run_round(srcs, inj)
        if r % 10 == 0 or late:
            torch.cuda.synchronize()
            for i in range(P):
                ok = ok and check(outs[i], counts[i])
            for c in counts[0].unique().tolist():
                hist[c] = hist.get(c, 0) + 1
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return {"mode": "emulated", "workers": P, "bytes": n * 4, "rounds": args.rounds, "rounds_per_s": round(args.rounds / dt, 1),
            "partial_sums_match_counts": ok, "count_values_seen_on_rank0": {str(k): v for k, v in sorted(hist.items())},
            "timeouts": cl.timeouts_seen()}


def distributed(args) -> dict:
    import torch.distributed as dist

    from akka_allreduce_b200 import AllReduceInput, AllreduceConfig
    from akka_allreduce_b200.master import AllreduceMaster
    from akka_allreduce_b200.parallel.device_worker import DeviceWorker

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    cfg = AllreduceConfig()
    cfg.worker.totalSize, cfg.worker.maxLag = world, 4
    cfg.data.dataSize, cfg.data.maxChunkSize, cfg.data.maxRound = args.size, args.chunk, args.rounds - 1
    cfg.thresholds.thAllreduce = cfg.thresholds.thReduce = cfg.thresholds.thComplete = 0.75
    cfg.device.backend, cfg.device.gate, cfg.device.timeoutMs = "cuda", "device", args.timeout_ms
    cfg.cluster.port = int(os.environ["MASTER_PORT"]) + 59
    master = AllreduceMaster(cfg) if rank == 0 else None
    if master:
        master.start()
    stragglers = set(range(world - args.stragglers, world))      # the last k ranks are slow
    rng = random.Random(rank)
    state = {"done": 0, "t0": None, "t1": None, "last": None, "slow_rounds": 0}
    # Every round is checked, on the device and without a host synchronisation (the sink runs on the round's stream):
    # a table turns the value — the bit mask of the contributors — into its population count.
    lut = torch.tensor([bin(i).count("1") for i in range(1 << world)], dtype=torch.int32, device=dev)
    acc = torch.zeros(4, dtype=torch.int64, device=dev)          # mismatching elements, count-0 elements, elements, -
    min_count = torch.full((1,), 99, dtype=torch.int32, device=dev)
    hist = torch.zeros(world + 1, dtype=torch.int64, device=dev)
    mine = None

    def source(req):
        nonlocal mine
        if mine is None:  # the worker id (= bit of this rank's contribution) is assigned by the master at InitWorkers
            mine = torch.full((args.size,), float(1 << w.id), device=dev)
        if rank in stragglers and rng.random() < args.straggle_prob:
            time.sleep(args.delay_us * 1e-6)                     # a slow data source = a straggling worker
        return AllReduceInput(mine)

    def sink(out):
        now = time.perf_counter()
        if state["t0"] is None:
            state["t0"] = now
        if state["last"] is not None and now - state["last"] > 0.9 * args.timeout_ms * 1e-3:
            state["slow_rounds"] += 1          # a round that sat out (most of) the in-kernel timeout
        state["last"] = state["t1"] = now
        state["done"] += 1
        if out.iteration % args.check_every == 0:
            verify(out.data, out.count, out.iteration % 16 == 0)

    def verify(data, count, sample_hist):
        bits = data.to(torch.int32)
        inb = bits & ((1 << world) - 1)
        bad = (inb.to(data.dtype) != data) | (lut[inb] != count)     # value = bit mask of exactly `count` contributors
        acc[0] += bad.sum()
        acc[1] += (count == 0).sum()
        acc[2] += count.numel()
        torch.minimum(min_count, count.min().reshape(1), out=min_count)
        if sample_hist:       # contributor-count histogram, sampled
            hist.add_(torch.bincount(count, minlength=world + 1)[: world + 1])

    # first use of these ops loads their kernels (tens of ms): do it now, not inside round 0 on every rank at once
    verify(torch.zeros(1024, device=dev), torch.zeros(1024, dtype=torch.int32, device=dev), True)
    acc.zero_(); hist.zero_(); min_count.fill_(99)
    torch.cuda.synchronize()

    w = DeviceWorker(source, sink, cfg, master_port=cfg.cluster.port, device=local)
    dist.barrier()
    w.start()
    finished = w.wait(600)
    w.stop()
    if master:
        master.wait(30)
        master.stop()
    secs = (state["t1"] - state["t0"]) if state["t0"] else float("nan")
    torch.cuda.synchronize()
    a4 = acc.tolist()
    t = torch.tensor([1.0 if a4[0] == 0 else 0.0, state["done"], secs, float(min_count.item()), a4[1], a4[2],
                      state["slow_rounds"], float(w.engine.timeouts_seen() if w.engine else -1)]
                     + [float(x) for x in hist.tolist()], dtype=torch.float64, device=dev)
    allt = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(allt, t)
    dist.barrier()
    dist.destroy_process_group()
    rows = [x.tolist() for x in allt]
    fast = [r for i, r in enumerate(rows) if i not in stragglers]
    return {"mode": "distributed", "workers": world, "straggler_ranks": sorted(stragglers), "bytes": args.size * 4,
            "rounds": args.rounds, "finished": bool(finished),
            "rounds_per_s_fast_ranks": round(min(r[1] / r[2] for r in fast), 1),
            "rounds_per_s_slowest_rank": round(min(r[1] / r[2] for r in rows), 1),
            "partial_sums_match_counts": all(r[0] == 1.0 for r in rows),
            "min_count_seen": int(min(r[3] for r in rows)), "elements_with_count_0": int(sum(r[4] for r in rows)),
            "elements_checked": int(sum(r[5] for r in rows)),
            "count_0_share_fast_ranks": round(sum(r[4] for r in fast) / max(1.0, sum(r[5] for r in fast)), 4),
            "rounds_that_sat_out_a_timeout": {"fast_ranks": int(sum(r[6] for r in fast)), "all_ranks": int(sum(r[6] for r in rows))},
            "last_round_tag_closed_by_timeout_per_rank": [int(r[7]) for r in rows],
            "count_histogram_fast_ranks_sampled": {str(k): int(sum(r[8 + k] for r in fast)) for k in range(world + 1)},
            "checked_every": args.check_every, "timeout_ms": args.timeout_ms, "kernel": os.environ.get("AAR_KERNEL", "auto"),
            "thresholds": "thReduce=thComplete=0.75, thAllreduce=0.75, maxLag=4", "delay_us": args.delay_us,
            "straggle_prob": args.straggle_prob}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--emulate", type=int, default=0)
    ap.add_argument("--size", type=int, default=1 << 22)       # 16 MB fp32
    ap.add_argument("--chunk", type=int, default=1 << 16)      # 256 KB chunks
    ap.add_argument("--rounds", type=int, default=300)
    ap.add_argument("--stragglers", type=int, default=2)
    ap.add_argument("--straggle-prob", type=float, default=0.3)
    ap.add_argument("--delay-us", type=int, default=2000)
    ap.add_argument("--timeout-ms", type=float, default=50.0)
    ap.add_argument("--check-every", type=int, default=1)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    res = emulated(a) if a.emulate else distributed(a)
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps(res), flush=True)
        if a.out:
            with open(a.out, "w") as f:
                f.write(json.dumps(res) + "\n")
    sys.exit(0 if res["partial_sums_match_counts"] else 1)
