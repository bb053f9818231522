#!/usr/bin/env python
"""Small, bounded workload for compute-sanitizer (memcheck / racecheck / synccheck) and for ncu: a few rounds of every
kernel with all ranks emulated on one GPU (one cooperative launch per round), small shapes, long in-kernel timeouts
(the tools slow the kernels down by one to two orders of magnitude; a wait must not give up because of the tool).

    compute-sanitizer --tool memcheck python benchmarks/sanitizer_target.py
    ncu --set full -k regex:aar_stream ... python benchmarks/sanitizer_target.py --big
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

from akka_allreduce_b200.ops.fused_allreduce import EmulatedCluster, Injection


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--big", action="store_true", help="one 8-rank 64 MB-per-rank shape for profiling instead of the small ones")
    ap.add_argument("--rounds", type=int, default=3)
    a = ap.parse_args()
    ok = True
    if a.big:
        cases = [dict(world=8, n=1 << 24, mc=1 << 18, kernel="stream"), dict(world=1, n=1 << 24, mc=1 << 18, kernel="stream")]
    else:
        cases = [dict(world=4, n=1 << 14, mc=1 << 11, kernel="stream"),
                 dict(world=4, n=1 << 14, mc=1 << 11, kernel="generic"),
                 dict(world=3, n=3 * 4096 + 1024, mc=1000, kernel="stream", align_blocks=True),
                 dict(world=4, n=778, mc=3, kernel=None, max_lag=3),                       # scalar paths of the generic kernel
                 dict(world=4, n=1 << 12, mc=1 << 10, kernel=None, algo="oneshot"),
                 dict(world=4, n=1 << 14, mc=1 << 11, kernel="stream", th_reduce=0.75, th_complete=0.75, inj=True),
                 dict(world=4, n=1 << 14, mc=1 << 11, kernel="stream", dtype=torch.bfloat16, align_blocks=True)]
    for case in cases:
        inj = case.pop("inj", False)
        world, n, mc = case.pop("world"), case.pop("n"), case.pop("mc")
        dtype = case.get("dtype", torch.float32)
        cl = EmulatedCluster(world, n, mc, timeout_ms=120_000, **case)
        for r in range(a.rounds):
            srcs = [torch.full((n,), float(i + 1 + r), dtype=dtype, device="cuda") for i in range(world)]
            outs, counts = cl.run_round(srcs, Injection(mute=1 << (world - 1)) if inj else None)
            torch.cuda.synchronize()
            contributors = range(world - 1) if inj else range(world)
            want = float(sum(i + 1 + r for i in contributors))
            if inj:  # thComplete < 1: a chunk may be left out (zeros, count 0); what is there must be the partial sum
                good = all(bool((((o.float() == want) & (c == len(contributors))) | ((o.float() == 0) & (c == 0))).all())
                           for o, c in zip(outs, counts))
            else:
                good = all(bool((o.float() == want).all()) for o in outs) and all(bool((c == world).all()) for c in counts)
            ok = ok and good
            if not good:
                print(f"MISMATCH world={world} n={n} {case} round {r}", flush=True)
        print(f"done world={world} n={n} kernel={cl.kernel_used()} {case}", flush=True)
        del cl
    print("ALL OK" if ok else "FAILED", flush=True)
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
