#!/usr/bin/env python
"""Headline benchmark: N-worker fp32 allreduce of a 64 MB message in 1 MB chunks, all thresholds 1,
maxLag 1 (BASELINE.json config 2), one worker per GPU, through the fused sm_100a kernel.

    python bench.py --gpus N --steps K --warmup W            # N=1 runs in-process
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Rank 0 prints ONE JSON line.  ``value`` is the whole-job aggregate algorithm bandwidth
(N x message bytes / step time — the reference's own "MBytes/sec" per worker, AllreduceWorker.scala:
329-335, summed over the workers), timed on the device with CUDA events, max over ranks.  The bus
bandwidth (nccl-tests convention), the NCCL all_reduce number for the same message, the end-to-end
number through the worker API (pinned-host source -> H2D -> kernel -> sink -> D2H digest) and the
emulated-8-worker single-GPU number ride along as extra keys.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def reference_arm() -> int:
    # The reference is a Scala/Akka (JVM) project: no setup.py/pyproject, and neither java nor sbt
    # exists in this image (DESIGN.md "Reference arm" records the attempted install).
    print(json.dumps({"impl": "reference",
                      "unavailable": "reference is a Scala/sbt (JVM + Akka 2.5.4) project; no JVM/sbt/ivy cache "
                                     "in the image and pip cannot install it (no setup.py/pyproject)"}))
    return 0


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--size-mb", type=float, default=64.0)
    ap.add_argument("--chunk-mb", type=float, default=1.0)
    ap.add_argument("--ctas", type=int, default=0)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm()
    args.warmup = max(args.warmup, 3)

    import torch
    import torch.distributed as dist

    from akka_allreduce_b200.ops import cuda as K
    from akka_allreduce_b200.parallel.engine import DeviceAllreduce, default_uds_dir
    from akka_allreduce_b200.utils.clocks import ClockSampler

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            sys.stderr.write("bench.py: --gpus N>1 must be launched with torch.distributed.run (one rank per GPU)\n")
            return 2
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # before anything is allocated: run (and first-touch pinned buffers) on the cores next to this GPU
    from akka_allreduce_b200.utils.affinity import bind_to_gpu
    placement = bind_to_gpu(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    K.lib()  # fail loudly here if the native library is missing

    n = int(args.size_mb * (1 << 20)) // 4
    mc = max(1, int(args.chunk_mb * (1 << 20)) // 4)
    msg_bytes = n * 4
    W, S = args.warmup, args.steps

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ------------------------------------------------------------------ device-timed collective
    eng = DeviceAllreduce(rank, world, n, mc, max_lag=1, dtype=torch.float32, th_reduce=1.0, th_complete=1.0,
                          th_allreduce=1.0, device=local, uds_dir=default_uds_dir(), align_blocks=False,
                          ctas=args.ctas or None, timeout_ms=10_000)
    g = torch.Generator(device="cpu").manual_seed(1234)
    NSRC = 4  # 4 x 64 MB of rotating inputs > 126 MB L2
    bases = [torch.randn(n, generator=g) for _ in range(NSRC)]
    srcs = [(b * float(rank + 1)).to(dev) for b in bases]
    cnts = [torch.empty(n, dtype=torch.int32, device=dev) for _ in range(2)]
    # Headline path: source and sink tensors are the engine's own symmetric slots (what a registered /
    # symmetric-memory buffer is for NCCL): the producer writes its message where the kernel needs it.
    zero_copy = eng.output_view(0) is not None and os.environ.get("AAR_BENCH_COPY") is None
    L = eng.geo.slots
    if zero_copy:
        outs = [eng.output_view(r) for r in range(L)]
        ins = [eng.input_view(r) for r in range(L)]
        for r in range(L):
            if ins[r] is not None:
                ins[r].copy_(srcs[r % NSRC])
            else:
                ins[r] = srcs[r % NSRC]
    else:
        outs = [torch.empty(n, dtype=torch.float32, device=dev) for _ in range(2)]

    def step(i: int) -> None:
        if zero_copy:
            eng.allreduce(ins[eng.round % L], outs[eng.round % L], cnts[i % 2])
        else:
            eng.allreduce(srcs[i % NSRC], outs[i % 2], cnts[i % 2])

    step(0)
    torch.cuda.synchronize()
    want = bases[0].to(dev) * float(sum(range(1, world + 1)))
    err = float((outs[0] - want).abs().max() / want.abs().max())
    ok = err < 1e-5 and bool((cnts[0] == world).all())
    del want
    for i in range(1, W):
        step(i)
    barrier()
    launches_before = eng.launcher.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        barrier()
        e0.record()
        for i in range(S):
            step(W + i)
        e1.record()
        torch.cuda.synchronize()
        barrier()
    ms = max_over_ranks(e0.elapsed_time(e1) / S)
    gpu_launches = eng.launcher.launches - launches_before
    algbw = msg_bytes / (ms * 1e-3) / 1e9
    value = algbw * world
    busbw = algbw * 2 * (world - 1) / world

    try:
        roof = roofline(world, eng.algo, msg_bytes, ms, zero_copy, os.path.join(ROOT, "MEASURED_PEAKS.json"))
    except Exception as ex:  # bookkeeping never takes the headline down
        roof = {"error": str(ex)[:200]}
    extras = {"busbw_GBps": round(busbw, 2), "algbw_per_rank_GBps": round(algbw, 2), "roofline": roof, "correct": ok,
              "max_rel_err": err, "multicast": eng.has_multicast, "algo": eng.algo, "ctas": eng.launcher.ctas,
              "zero_copy_buffers": zero_copy,
              "timeouts": eng.timeouts_seen()}

    # ------------------------------------------------------------------ same kernel, arbitrary user tensors
    if zero_copy:
        outs2 = [torch.empty(n, dtype=torch.float32, device=dev) for _ in range(2)]
        for i in range(W):
            eng.allreduce(srcs[i % NSRC], outs2[i % 2], cnts[i % 2])
        barrier()
        e0.record()
        for i in range(S):
            eng.allreduce(srcs[i % NSRC], outs2[i % 2], cnts[i % 2])
        e1.record()
        torch.cuda.synchronize()
        ms2 = max_over_ranks(e0.elapsed_time(e1) / S)
        extras["user_tensor_api"] = {"ms": round(ms2, 4), "aggregate_algbw_GBps": round(world * msg_bytes / (ms2 * 1e-3) / 1e9, 2),
                                     "note": "src/out are ordinary torch tensors: staged into and flushed from the slots by the kernel"}
        del outs2

    # ------------------------------------------------------------------ NCCL all_reduce, same message
    if world > 1:
        buf = srcs[0].clone()
        for _ in range(W):
            dist.all_reduce(buf)
        barrier()
        e0.record()
        for _ in range(S):
            dist.all_reduce(buf)
        e1.record()
        torch.cuda.synchronize()
        nccl_ms = max_over_ranks(e0.elapsed_time(e1) / S)
        nccl_alg = msg_bytes / (nccl_ms * 1e-3) / 1e9
        extras["nccl"] = {"ms": round(nccl_ms, 4), "busbw_GBps": round(nccl_alg * 2 * (world - 1) / world, 2),
                          "aggregate_algbw_GBps": round(nccl_alg * world, 2)}
        extras["vs_nccl"] = round(nccl_ms / ms, 3)
        del buf
    eng.close()
    del srcs, outs, cnts

    # ------------------------------------------------------------------ end to end through the worker API
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(rank, world, local, n, mc, W, S, max_over_ranks, barrier)

    # ------------------------------------------------------------------ extra: 8 emulated workers on one GPU
    if world == 1 and not args.no_extras:
        try:
            extras["emulated_8_workers_1_gpu"] = run_emulated(n, mc, W, min(S, 20))
        except Exception as ex:  # extras never take the headline down
            extras["emulated_8_workers_1_gpu"] = {"error": str(ex)[:200]}

    if rank == 0:
        line = {
            "metric": "allreduce_aggregate_algbw_GBps", "value": round(value, 2), "unit": "GB/s",
            "n_gpus": world, "steps": S, "warmup": W, "ms_per_step": round(ms, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32",
            "data": "synthetic", "impl": "ours",
            "config": {"model": "allreduce 64MB fp32 message, 1MB chunks, thAllreduce=thReduce=thComplete=1.0, maxLag=1"
                       if args.size_mb == 64 else f"allreduce {args.size_mb}MB fp32",
                       "global_batch": world, "seq_len": n, "parallelism": f"dp{world}",
                       "workers": world, "dataSize": n, "maxChunkSize": mc, "maxLag": 1,
                       "cpu_placement": placement,  # rank 0's; None = left to the OS (single node / hidden sysfs)
                       "l2": (f"working set larger than L2: {L} staging + {L} sink slots + 2 count buffers of {args.size_mb:g} MB"
                              if zero_copy else f"inputs larger than L2: {NSRC} rotating {args.size_mb:g} MB sources")},
            "clocks": {"sm_mhz": clocks.summary["sm_mhz"], "sm_max_mhz": clocks.summary["sm_max_mhz"],
                       "reasons": clocks.summary["reasons"], "samples": clocks.summary["samples"]},
            "e2e": e2e, "gpu_launches": gpu_launches,
        }
        line.update(extras)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0 if ok else 1


def roofline(world: int, algo: str, msg_bytes: int, ms: float, zero_copy: bool, peaks_path: str) -> dict:
    """Fraction of the roofline of the variant that actually ran.  One GPU: the round is HBM traffic only (the
    single contribution is read and written once, the counts written once, plus the staging/flush copies when the
    tensors are not the slots themselves), against the MEASURED copy bandwidth in MEASURED_PEAKS.json.  Several GPUs:
    NVLink bus bandwidth against 900 GB/s per direction nominal — B for unicast two-shot, B x 2(P-1)/(P+1) for the
    in-switch (NVLS) variant, and per-GPU ingress (P-1) x N for one-shot."""
    out = {}
    t = ms * 1e-3
    if world == 1:
        hbm = None
        try:
            with open(peaks_path) as f:
                hbm = float(json.load(f)["hbm_gbs"])
        except Exception:
            pass
        moved = msg_bytes * (3 if zero_copy else 7)   # slots: r + w + counts; user tensors: + stage (r+w) + flush (r+w)
        out = {"bound": "HBM", "bytes_per_step": moved, "achieved_GBps": round(moved / t / 1e9, 1),
               "peak_GBps": hbm, "peak_source": "MEASURED_PEAKS.json hbm_gbs" if hbm else "unavailable",
               "fraction": round(moved / t / 1e9 / hbm, 3) if hbm else None}
    else:
        link = 900.0
        busbw = msg_bytes / t / 1e9 * 2 * (world - 1) / world
        if algo == "nvls":
            peak = link * 2 * (world - 1) / (world + 1)
        elif algo == "oneshot":
            peak = link * 2 / world      # ingress (P-1) x N at link rate, expressed as bus bandwidth
        else:
            peak = link
        out = {"bound": "NVLink", "variant": algo, "busbw_GBps": round(busbw, 1), "roofline_busbw_GBps": round(peak, 1),
               "peak_source": "900 GB/s per direction nominal", "fraction": round(busbw / peak, 3)}
    return out


def run_e2e(rank, world, local, n, mc, W, S, max_over_ranks, barrier):
    """Rounds through the public API: AllreduceMaster + DeviceWorker with a pinned-host source."""
    import torch

    from akka_allreduce_b200 import AllReduceInput, AllreduceConfig
    from akka_allreduce_b200.master import AllreduceMaster
    from akka_allreduce_b200.parallel.device_worker import DeviceWorker

    cfg = AllreduceConfig()
    cfg.worker.totalSize, cfg.worker.maxLag = world, 1
    cfg.data.dataSize, cfg.data.maxChunkSize, cfg.data.maxRound = n, mc, W + S - 1
    cfg.thresholds.thAllreduce = cfg.thresholds.thReduce = cfg.thresholds.thComplete = 1.0
    cfg.device.backend, cfg.device.gate, cfg.device.timeoutMs = "cuda", "device", 10_000
    port = int(os.environ.get("MASTER_PORT", "29500")) + 17
    cfg.cluster.port = port
    master = None
    if rank == 0:
        master = AllreduceMaster(cfg)
        master.start()
    pinned = [torch.randn(n).pin_memory() for _ in range(3)]
    marks = {}
    digest = torch.zeros(2, dtype=torch.float32).pin_memory()

    def source(req):
        return AllReduceInput(pinned[req.iteration % len(pinned)])

    def sink(out):
        # the step's result leaves the device: checksum of the reduced vector + smallest contributor count
        d = torch.stack([out.data.sum(), out.count.min().float()])
        digest.copy_(d, non_blocking=False)
        if out.iteration == W - 1:
            marks["t0"] = time.perf_counter()
        if out.iteration == W + S - 1:
            marks["t1"] = time.perf_counter()
            marks["min_count"] = float(digest[1])

    worker = DeviceWorker(source, sink, cfg, master_port=port, device=local)
    barrier()
    worker.start()
    try:
        worker.wait(600)
    finally:
        worker.stop()
        if master is not None:
            master.stop()
    barrier()
    if "t0" not in marks or "t1" not in marks:
        return None
    sec = max_over_ranks((marks["t1"] - marks["t0"]) / S)
    return {"value": round(world * n * 4 / sec / 1e9, 3), "unit": "GB/s", "ms_per_step": round(sec * 1e3, 4),
            "h2d_bytes_per_step": n * 4, "d2h_bytes_per_step": 8, "min_count": marks.get("min_count"),
            "path": "AllreduceMaster + DeviceWorker(source=pinned host tensor, sink=device views) over the coordinator"}


def run_emulated(n, mc, W, S):
    """8 workers emulated on one GPU (one cooperative launch per round): the whole 8-rank protocol on HBM."""
    import torch

    from akka_allreduce_b200.ops.fused_allreduce import EmulatedCluster

    world = 8
    cl = EmulatedCluster(world, n, mc, max_lag=1, timeout_ms=10_000)
    srcs = [torch.randn(n, device="cuda") for _ in range(world)]
    for _ in range(W):
        cl.run_round(srcs)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(S):
        cl.run_round(srcs)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / S
    alg = n * 4 / (ms * 1e-3) / 1e9
    return {"ms_per_round": round(ms, 4), "aggregate_algbw_GBps": round(alg * world, 2),
            "busbw_GBps": round(alg * 2 * (world - 1) / world, 2), "ctas_per_rank": cl.launcher.ctas}


if __name__ == "__main__":
    sys.exit(main())
