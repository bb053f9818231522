#!/usr/bin/env python
"""Long run with EVERY round verified: the check a round-0 comparison cannot give (torchrun, >= 2 GPUs).

    torchrun --nproc-per-node 8 benchmarks/soak.py --rounds 2000 --size-mb 64 [--zero-copy] [--algo nvls]

Round k's input on rank q is ``base * (q + 1) + (k % 251)``, rewritten every round (into the engine's staging slot
with --zero-copy, else into an ordinary tensor), so every round has a different expectation
``base * sum(q + 1) + P * (k % 251)`` and a stale ring slot, a flag that overtook its data or a torn tile shows up as a
mismatch in that very round.  Prints one JSON line: rounds, mismatching rounds (first few listed), worst error,
counts check, timeouts.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--rounds", type=int, default=2000)
    ap.add_argument("--size-mb", type=float, default=64.0)
    ap.add_argument("--chunk-mb", type=float, default=1.0)
    ap.add_argument("--algo", default="auto")
    ap.add_argument("--kernel", default=None)
    ap.add_argument("--zero-copy", action="store_true")
    ap.add_argument("--max-lag", type=int, default=1)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from akka_allreduce_b200.parallel.engine import DeviceAllreduce, default_uds_dir

    n = int(a.size_mb * (1 << 20)) // 4
    mc = max(1, int(a.chunk_mb * (1 << 20)) // 4)
    eng = DeviceAllreduce(rank, world, n, mc, max_lag=a.max_lag, th_allreduce=1.0, device=local, uds_dir=default_uds_dir(),
                          timeout_ms=10_000, algo=a.algo, kernel=a.kernel)
    g = torch.Generator(device="cpu").manual_seed(99)
    base = torch.randn(n, generator=g).to(dev)
    mine = base * float(rank + 1)
    ranksum = float(sum(range(1, world + 1)))
    scale = float((base.abs().max() * ranksum + 251 * world).item())
    src = torch.empty(n, device=dev)
    out = torch.empty(n, device=dev)
    cnt = torch.empty(n, dtype=torch.int32, device=dev)
    zc = a.zero_copy and eng.output_view(0) is not None
    bad, worst, bad_counts = [], 0.0, 0
    errs = torch.zeros(a.rounds, device=dev)
    cnt_ok = torch.zeros(a.rounds, device=dev)
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    for k in range(a.rounds):
        shift = float(k % 251)
        r = eng.round
        iv = eng.input_view(r) if zc else None
        ov = eng.output_view(r) if zc else out
        s = iv if iv is not None else src
        torch.add(mine, shift, out=s)                       # the producer rewrites its message every round
        eng.allreduce(s, ov, cnt)
        errs[k] = (ov - (base * ranksum + shift * world)).abs().max()      # same stream: ordered after the round
        cnt_ok[k] = (cnt == world).all()
    torch.cuda.synchronize()
    secs = time.perf_counter() - t0
    errs = (errs / scale).cpu()
    cnt_ok = cnt_ok.cpu()
    bad = [int(i) for i in torch.nonzero((errs > 1e-5) | (cnt_ok == 0)).flatten()[:8]]
    nbad = int(((errs > 1e-5) | (cnt_ok == 0)).sum())
    t = torch.tensor([nbad, float(errs.max()), float(eng.timeouts_seen())], dtype=torch.float64, device=dev)
    allt = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(allt, t)
    res = {"soak": True, "workers": world, "rounds": a.rounds, "bytes": n * 4, "algo": eng.algo,
           "kernel": "stream" if eng.launcher.stream_expected else "generic", "zero_copy": bool(zc),
           "rounds_with_a_mismatch_per_rank": [int(x[0]) for x in allt], "first_bad_rounds_rank0": bad,
           "max_rel_err": max(float(x[1]) for x in allt), "timeouts_per_rank": [int(x[2]) for x in allt],
           "seconds": round(secs, 2), "ms_per_round_incl_verification": round(secs / a.rounds * 1e3, 4)}
    ok = all(int(x[0]) == 0 for x in allt)
    if rank == 0:
        print(json.dumps(res), flush=True)
        if a.out:
            with open(a.out, "a") as f:
                f.write(json.dumps(res) + "\n")
    eng.close()
    dist.barrier()
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
