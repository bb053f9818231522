"""One rank of the CPU DDP test (spawned by tests/test_ddp_cpu.py): a tiny GPT-2 under DDP with the shared-memory
allreduce hook, against the same model under stock DDP (gloo allreduce).  Prints 'ok <max difference>' or fails."""
import os
import sys

import torch
import torch.distributed as dist
from torch.nn.parallel import DistributedDataParallel as DDP

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from akka_allreduce_b200.models.gpt2 import GPT2, GPT2Config  # noqa: E402
from akka_allreduce_b200.parallel.ddp import ShmAllreduceCommHook  # noqa: E402


def train(model, hook, rank, steps=3):
    ddp = DDP(model, bucket_cap_mb=0.05)       # small buckets: several per backward, a few sizes
    if hook is not None:
        ddp.register_comm_hook(None, hook)
    opt = torch.optim.SGD(ddp.parameters(), lr=0.1)
    g = torch.Generator().manual_seed(100 + rank)      # every rank sees different data
    for _ in range(steps):
        x = torch.randint(0, 512, (2, 17), generator=g)
        opt.zero_grad()
        ddp(x[:, :-1], x[:, 1:]).backward()
        opt.step()
    return [p.detach().clone() for p in model.parameters()]


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo")
    torch.manual_seed(0)
    torch.set_num_threads(1)
    cfg = GPT2Config.tiny()
    a, b = GPT2(cfg), GPT2(cfg)
    b.load_state_dict(a.state_dict())
    hook = ShmAllreduceCommHook(rank, world, name=f"/aar_ddp_{os.environ['MASTER_PORT']}")
    pa = train(a, hook, rank)
    pb = train(b, None, rank)
    diff = max(float((x - y).abs().max()) for x, y in zip(pa, pb))
    scale = max(float(y.abs().max()) for y in pb)
    dist.barrier()
    hook.close()
    ok = diff <= 1e-5 * max(1.0, scale) and hook.buckets_reduced >= 3
    print(f"{'ok' if ok else 'FAIL'} rank {rank} max difference {diff:.3e} buckets {hook.buckets_reduced}", flush=True)
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
