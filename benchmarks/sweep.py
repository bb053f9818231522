#!/usr/bin/env python
"""Message-size / chunk-size / CTA sweep of the fused allreduce vs torch.distributed NCCL all_reduce.

    python benchmarks/sweep.py --emulate 8                       # one GPU, 8 emulated workers (HBM only)
    torchrun --nproc-per-node N benchmarks/sweep.py              # N GPUs over NVLink, with the NCCL column

Every number is device-timed (CUDA events), max over ranks; inputs rotate through > L2 worth of buffers.
Prints one JSON object per (size, chunk, ctas) point; rank 0 also writes them to --out.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist


def parse_sizes(text):
    out = []
    for tok in text.split(","):
        tok = tok.strip().upper()
        mult = 1
        if tok.endswith("K"):
            mult, tok = 1 << 10, tok[:-1]
        elif tok.endswith("M"):
            mult, tok = 1 << 20, tok[:-1]
        elif tok.endswith("G"):
            mult, tok = 1 << 30, tok[:-1]
        out.append(int(float(tok) * mult))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="1K,16K,256K,1M,4M,16M,64M,256M")
    ap.add_argument("--chunks", default="1M")
    ap.add_argument("--ctas", default="0")
    ap.add_argument("--emulate", type=int, default=0)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--dtype", default="float32")
    ap.add_argument("--out", default="")
    ap.add_argument("--no-nccl", action="store_true")
    ap.add_argument("--algo", default="auto")
    ap.add_argument("--kernel", default=None, help="auto | generic | stream (default: AAR_KERNEL or auto)")
    ap.add_argument("--zero-copy", action="store_true", help="source/sink tensors alias the engine's symmetric slots")
    ap.add_argument("--no-counts", action="store_true")
    ap.add_argument("--trace", action="store_true", help="print per-phase device timestamps of one extra round")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    dtype = getattr(torch, args.dtype)
    esz = torch.empty((), dtype=dtype).element_size()
    rows = []

    def mx(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed(fn, iters, warm):
        for i in range(warm):
            fn(i)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(iters):
            fn(warm + i)
        e1.record()
        torch.cuda.synchronize()
        return mx(e0.elapsed_time(e1) / iters)

    for size in parse_sizes(args.sizes):
        n = max(1, size // esz)
        nccl_ms = None
        if world > 1 and not args.no_nccl:
            buf = torch.randn(n, device=dev).to(dtype)
            nccl_ms = timed(lambda i: dist.all_reduce(buf), args.iters, args.warmup)
            del buf
        for chunk in parse_sizes(args.chunks):
            mc = max(1, min(chunk // esz, n))
            for ctas in [int(c) for c in args.ctas.split(",")]:
                P = args.emulate or world
                nsrc = max(1, min(4, (256 << 20) // max(size, 1)))
                if args.emulate:
                    from akka_allreduce_b200.ops.fused_allreduce import EmulatedCluster

                    cl = EmulatedCluster(P, n, mc, max_lag=1, dtype=dtype, ctas=ctas or None, timeout_ms=10_000,
                                         align_blocks=True, kernel=args.kernel)
                    srcs = [torch.randn(n, device=dev).to(dtype) for _ in range(P)]
                    ms = timed(lambda i: cl.run_round(srcs), args.iters, args.warmup)
                    used = cl.launcher.ctas
                    kernel_used = cl.kernel_used()
                    algo_used = "twoshot"
                    zc = False
                    del cl, srcs
                else:
                    from akka_allreduce_b200.parallel.engine import DeviceAllreduce, default_uds_dir

                    eng = DeviceAllreduce(rank, world, n, mc, max_lag=1, dtype=dtype, th_allreduce=1.0, device=local,
                                          uds_dir=default_uds_dir(), ctas=ctas or None, timeout_ms=10_000, algo=args.algo,
                                          kernel=args.kernel)
                    srcs = [torch.randn(n, device=dev).to(dtype) for _ in range(nsrc)]
                    outs = [torch.empty(n, dtype=dtype, device=dev) for _ in range(2)]
                    cnts = [torch.empty(n, dtype=torch.int32, device=dev) for _ in range(2)]
                    zc = False
                    if args.zero_copy and eng.output_view(0) is not None:
                        zc = True
                        L = eng.geo.slots
                        outs = [eng.output_view(r) for r in range(L)]
                        ins = [eng.input_view(r) for r in range(L)]
                        for r in range(L):
                            if ins[r] is not None:
                                ins[r].copy_(srcs[r % nsrc])
                            else:
                                ins[r] = srcs[r % nsrc]
                        srcs, nsrc = ins, L
                    if args.no_counts:
                        cnts = [None, None]
                    if zc:
                        ms = timed(lambda i: eng.allreduce(srcs[eng.round % nsrc], outs[eng.round % nsrc], cnts[i % 2]), args.iters, args.warmup)
                    else:
                        ms = timed(lambda i: eng.allreduce(srcs[i % nsrc], outs[i % 2], cnts[i % 2]), args.iters, args.warmup)
                    used = eng.launcher.ctas
                    algo_used = eng.algo
                    kernel_used = "oneshot" if eng.algo == "oneshot" else ("stream" if eng.launcher.stream_expected else "generic")
                    if args.trace:
                        eng.enable_trace()
                        if world > 1:
                            dist.barrier()
                        eng.allreduce(srcs[eng.round % nsrc] if zc else srcs[0], outs[eng.round % nsrc] if zc else outs[0], cnts[0])
                        print(f"[rank {rank}] trace bytes={n * esz} algo={algo_used}: {json.dumps(eng.trace_summary())}", flush=True)
                    eng.close()
                    del srcs, outs, cnts
                alg = n * esz / (ms * 1e-3) / 1e9
                row = {"bytes": n * esz, "chunk_bytes": mc * esz, "workers": P, "gpus": world, "ctas": used,
                       "ms": round(ms, 5), "algbw_GBps": round(alg, 2), "busbw_GBps": round(alg * 2 * (P - 1) / P, 2),
                       "emulated": bool(args.emulate), "dtype": args.dtype, "algo": algo_used,
                       "zero_copy": bool(args.zero_copy) and not args.emulate and zc, "counts": not args.no_counts,
                       "kernel": kernel_used}
                if nccl_ms is not None:
                    row["nccl_ms"] = round(nccl_ms, 5)
                    row["nccl_busbw_GBps"] = round(n * esz / (nccl_ms * 1e-3) / 1e9 * 2 * (P - 1) / P, 2)
                    row["speedup_vs_nccl"] = round(nccl_ms / ms, 3)
                rows.append(row)
                if rank == 0:
                    print(json.dumps(row), flush=True)
    if rank == 0 and args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as f:
            for r in rows:
                f.write(json.dumps(r) + "\n")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
