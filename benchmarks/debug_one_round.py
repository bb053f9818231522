"""One emulated round (8 ranks on one GPU) — smallest reproducer for sanitizer / ncu runs."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from akka_allreduce_b200.ops.fused_allreduce import EmulatedCluster

world = int(os.environ.get("W", 8))
n = int(os.environ.get("N", 1 << 20))
mc = int(os.environ.get("MC", 1 << 14))
cl = EmulatedCluster(world, n, mc, max_lag=1, timeout_ms=3000, algo=os.environ.get("ALGO", "twoshot"))
srcs = [torch.randn(n, device="cuda") for _ in range(world)]
for r in range(int(os.environ.get("ROUNDS", 3))):
    outs, counts = cl.run_round(srcs)
torch.cuda.synchronize()
want = sum(srcs[1:], srcs[0].clone())
print("max err", float((outs[0] - want).abs().max()), "count ok", bool((counts[0] == world).all()), "timeouts", cl.timeouts_seen())
