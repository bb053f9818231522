#!/usr/bin/env python
"""GPT-2 small, bf16, data parallel: step time with the fused-allreduce comm hook vs stock NCCL DDP.

    torchrun --nproc-per-node N benchmarks/ddp_gpt2.py [--batch 8] [--seq 1024] [--steps 10] [--tiny]

Synthetic tokens, random-init weights (BASELINE.json config 5).  Step = forward + backward (gradient
sync overlapped with backward through the DDP hook) + AdamW step.  CUDA-event timed, max over ranks.
Also checks that both variants produce the same gradients on the first step.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist
from torch.nn.parallel import DistributedDataParallel as DDP

from akka_allreduce_b200.models.gpt2 import GPT2, GPT2Config
from akka_allreduce_b200.parallel.ddp import AllreduceCommHook


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--seq", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--tiny", action="store_true")
    ap.add_argument("--bucket-mb", type=int, default=25)
    ap.add_argument("--out", default="")
    ap.add_argument("--ctas", type=int, default=0, help="work CTAs of the comm kernel (0 = hook default)")
    ap.add_argument("--skip-nccl", action="store_true")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    cfg = GPT2Config.tiny() if args.tiny else GPT2Config.small()
    seq = min(args.seq, cfg.n_positions)

    def build():
        torch.manual_seed(0)
        return GPT2(cfg).to(dev).to(torch.bfloat16)

    def batch(step):
        g = torch.Generator(device="cpu").manual_seed(1000 * step + rank)
        x = torch.randint(0, cfg.vocab_size, (args.batch, seq + 1), generator=g).to(dev)
        return x[:, :-1].contiguous(), x[:, 1:].contiguous()

    def mx(v):
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    results, grads = {}, {}
    for variant in (("fused_hook",) if args.skip_nccl else ("nccl", "fused_hook")):
        model = DDP(build(), device_ids=[local], bucket_cap_mb=args.bucket_mb, gradient_as_bucket_view=True)
        hook = None
        if variant == "fused_hook":
            hook = AllreduceCommHook(rank, world, local, ctas=args.ctas or None)
            model.register_comm_hook(None, hook)
        opt = torch.optim.AdamW(model.parameters(), lr=1e-4, fused=True)
        times = []
        for step in range(args.warmup + args.steps):
            x, y = batch(step)
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            loss = model(x, y)
            loss.backward()
            if step == 0:
                grads[variant] = torch.cat([p.grad.float().flatten()[:4096] for p in model.parameters()]).clone()
            opt.step()
            opt.zero_grad(set_to_none=False)
            e1.record()
            torch.cuda.synchronize()
            if step >= args.warmup:
                times.append(e0.elapsed_time(e1))
        ms = mx(sum(times) / len(times))
        results[variant] = {"ms_per_step": round(ms, 3), "tokens_per_s": round(world * args.batch * seq / (ms * 1e-3), 1),
                            "final_loss": round(float(loss), 4)}
        if hook is not None:
            results[variant]["buckets_per_step"] = hook.buckets_reduced // (args.warmup + args.steps)
            results[variant]["engines"] = {f"{k[0]}x{str(k[1]).replace('torch.', '')}": e.algo for k, e in hook._engines.items()}
            torch.cuda.synchronize()
            hook.close()
        del model, opt
        torch.cuda.empty_cache()
    if args.skip_nccl:
        if rank == 0:
            print(json.dumps({"ctas": args.ctas, **results}), flush=True)
        dist.barrier()
        dist.destroy_process_group()
        return 0
    diff = float((grads["nccl"] - grads["fused_hook"]).abs().max())
    scale = float(grads["nccl"].abs().max())
    out = {"model": "gpt2-tiny" if args.tiny else "gpt2-small", "params": sum(p.numel() for p in build().parameters()),
           "dtype": "bf16", "gpus": world, "batch_per_gpu": args.batch, "seq": seq, "bucket_mb": args.bucket_mb,
           "grad_max_abs_diff_step0": diff, "grad_max_abs": scale, **results,
           "speedup_vs_nccl_ddp": round(results["nccl"]["ms_per_step"] / results["fused_hook"]["ms_per_step"], 3)}
    if rank == 0:
        print(json.dumps(out), flush=True)
        if args.out:
            with open(args.out, "w") as f:
                f.write(json.dumps(out) + "\n")
    dist.barrier()
    dist.destroy_process_group()
    return 0 if diff <= 0.02 * max(scale, 1e-6) + 1e-3 else 1


if __name__ == "__main__":
    sys.exit(main())
