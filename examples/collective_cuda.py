#!/usr/bin/env python
"""The fused kernel as a stream-ordered collective, one rank per GPU.

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 examples/collective_cuda.py

Shows both ways to feed it: ordinary tensors (the kernel stages and flushes them) and the engine's own
symmetric slots (nothing to copy), the reduce-scatter / all-gather halves of the round on their own, plus partial
completion: with thReduce = thComplete = 0.75 a rank that
stays silent is simply left out and `count` says so.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

from akka_allreduce_b200.ops.fused_allreduce import Injection
from akka_allreduce_b200.parallel.engine import DeviceAllreduce

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))  # only used for barriers / printing
n = 1 << 22

eng = DeviceAllreduce(rank, world, n, max_chunk=1 << 18, max_lag=1)      # thresholds default to 1
x = torch.full((n,), float(rank + 1), device="cuda")
out = torch.empty_like(x)
count = torch.empty(n, dtype=torch.int32, device="cuda")
eng.allreduce(x, out, count)                                             # ordinary tensors
torch.cuda.synchronize()
assert float(out[0]) == world * (world + 1) / 2 and int(count.min()) == world

slot_in, slot_out = eng.input_view(eng.round), eng.output_view(eng.round)
if slot_out is not None:                                                 # zero-copy slots
    src = slot_in if slot_in is not None else x
    src.fill_(float(rank + 1))
    eng.allreduce(src, slot_out, count)
    torch.cuda.synchronize()
    assert float(slot_out[-1]) == world * (world + 1) / 2

if eng.algo in ("twoshot", "nvls") and eng.launcher.stream_expected:     # the two halves of the round on their own
    lo, hi = eng.block_range()
    block = torch.empty(hi - lo, device="cuda")
    eng.reduce_scatter(x, block)                                         # my block of the sum ...
    eng.all_gather(block, out, count)                                    # ... and everybody's blocks back together
    torch.cuda.synchronize()
    assert float(block[0]) == world * (world + 1) / 2 and float(out[-1]) == world * (world + 1) / 2
    assert int(count.min()) == 1                                         # one contributor per element: its owner
eng.close()

if world >= 4:                                                           # partial completion
    th = (world - 1) / world
    part = DeviceAllreduce(rank, world, n, max_chunk=1 << 18, th_reduce=th, th_complete=1.0, timeout_ms=200)
    part.allreduce(x, out, count, inj=Injection(mute=1 << (world - 1)))  # the last rank stays silent
    torch.cuda.synchronize()
    assert int(count.max()) == world - 1 and float(out[0]) == world * (world + 1) / 2 - world
    part.close()
dist.barrier()
if rank == 0:
    print(f"ok on {world} GPUs: algo={eng.algo}, multicast={eng.has_multicast}")
dist.destroy_process_group()
