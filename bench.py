#!/usr/bin/env python
"""Headline benchmark: N-worker fp32 allreduce of a 64 MB message in 1 MB chunks, all thresholds 1,
maxLag 1 (BASELINE.json config 2), one worker per GPU, through the fused sm_100a kernel.

    python bench.py --gpus N --steps K --warmup W            # N=1 runs in-process
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Rank 0 prints ONE JSON line.  ``value`` is the whole-job aggregate algorithm bandwidth
(N x message bytes / step time — the reference's own "MBytes/sec" per worker, AllreduceWorker.scala:
329-335, summed over the workers), timed on the device with CUDA events, max over ranks.

What is timed as the headline is the ORDINARY-TENSOR call — ``eng.allreduce(src, out, counts)`` with src/out/
counts plain torch tensors, a different source every step — the same contract as ``dist.all_reduce`` on an
ordinary tensor, which is timed next to it (``nccl``, ``vs_nccl``).  The registered-buffer mode (source and sink
ARE the engine's symmetric slots, what a symmetric-memory buffer is for NCCL / torch symm_mem) is reported under
``registered_buffers`` with ITS like-for-like comparators: ``torch.ops.symm_mem.multimem_all_reduce_`` /
``two_shot_all_reduce_`` on a symm_mem buffer and NCCL on an NCCL-registered buffer.

Every mode is verified AFTER its timed loop: the last timed output against the fp32 expectation, then
``VERIFY`` more rounds whose inputs are rewritten every round (so a stale ring slot cannot pass), each checked
against the expectation, the contributor counts and NCCL's result.
"""
from __future__ import annotations

import argparse
import json
import os
import shutil
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
VERIFY = 6  # verification rounds per mode, inputs rewritten every round


def reference_arm() -> int:
    """The reference is a Scala/sbt project (Akka 2.5.4 on the JVM).  Probe for a toolchain and a dependency
    cache on THIS box before saying it cannot run (DESIGN.md "Reference arm" records the pip attempt)."""
    found = {t: shutil.which(t) for t in ("java", "sbt", "scala", "scalac", "cs", "mvn")}
    home = os.path.expanduser("~")
    caches = {d: os.path.isdir(os.path.join(home, d)) for d in (".ivy2", ".sbt", ".m2", ".cache/coursier")}
    ref_installed = os.path.isdir(os.path.join(ROOT, "baseline", "_ref")) and bool(os.listdir(os.path.join(ROOT, "baseline", "_ref")))
    have_jvm = bool(found["java"]) and bool(found["sbt"]) and (caches[".ivy2"] or caches[".cache/coursier"])
    why = ("reference is a Scala/sbt (JVM + Akka 2.5.4) project: probed java/sbt/scala/coursier/mvn -> "
           + ", ".join(f"{k}={'yes' if v else 'no'}" for k, v in found.items())
           + "; dependency caches " + ", ".join(f"{k}={'yes' if v else 'no'}" for k, v in caches.items())
           + f"; baseline/_ref installed={'yes' if ref_installed else 'no'} (pip cannot install it: no setup.py/pyproject)")
    if have_jvm:
        why += "; a JVM and sbt exist but Akka 2.5.4 / Scala 2.12.2 jars still need the network (none here)"
    print(json.dumps({"impl": "reference", "unavailable": why}))
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

    def all_ok(flag: bool) -> bool:
        if world == 1:
            return flag
        t = torch.tensor([0 if flag else 1], dtype=torch.int32, device=dev)
        dist.all_reduce(t)
        return int(t.item()) == 0

    def timed(fn, steps=S, warm=W):
        """W untimed + S timed calls of fn(i), CUDA events on the launching stream, max over ranks (ms per call)."""
        for i in range(warm):
            fn(i)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(warm + i)
        e1.record()
        torch.cuda.synchronize()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1) / steps)

    # ------------------------------------------------------------------ inputs
    eng = DeviceAllreduce(rank, world, n, mc, max_lag=1, dtype=torch.float32, th_reduce=1.0, th_complete=1.0,
                          th_allreduce=1.0, device=local, uds_dir=default_uds_dir(), align_blocks=False,
                          ctas=args.ctas or None, timeout_ms=10_000)
    g = torch.Generator(device="cpu").manual_seed(1234)
    NSRC = 4  # 4 x 64 MB of rotating inputs > 126 MB L2
    bases = [torch.randn(n, generator=g).to(dev) for _ in range(NSRC)]
    srcs = [b * float(rank + 1) for b in bases]
    ranksum = float(sum(range(1, world + 1)))
    cnts = [torch.empty(n, dtype=torch.int32, device=dev) for _ in range(2)]
    outs = [torch.empty(n, dtype=torch.float32, device=dev) for _ in range(2)]
    scratch_src = torch.empty(n, dtype=torch.float32, device=dev)

    def expect(k: int, shift: float = 0.0):
        return bases[k % NSRC] * ranksum + shift * world

    def close(got, want) -> float:
        return float((got - want).abs().max() / want.abs().max())

    def verify_rounds(run_round) -> dict:
        """VERIFY rounds with inputs rewritten every round: src_k = base[k] * (rank+1) + k on every rank, so the
        expectation is base[k] * sum(rank+1) + P*k; each output is also compared with NCCL's all_reduce."""
        worst, worst_nccl, bad = 0.0, 0.0, 0
        for k in range(VERIFY):
            torch.add(srcs[k % NSRC], float(k), out=scratch_src)
            out, cnt = run_round(scratch_src, k)
            torch.cuda.synchronize()
            err = close(out, expect(k, float(k)))
            worst = max(worst, err)
            if not (err < 1e-5 and bool((cnt == world).all())):
                bad += 1
            if world > 1:
                ref = scratch_src.clone()
                dist.all_reduce(ref)
                worst_nccl = max(worst_nccl, close(out, ref))
        ok = all_ok(bad == 0 and worst_nccl < 1e-5)
        return {"rounds": VERIFY, "ok": ok, "max_rel_err": worst, "max_rel_err_vs_nccl": worst_nccl if world > 1 else None}

    extras: dict = {}
    # ------------------------------------------------------------------ headline: ordinary tensors
    def step_user(i: int) -> None:
        eng.allreduce(srcs[i % NSRC], outs[i % 2], cnts[i % 2])

    launches_before = eng.launcher.launches
    with ClockSampler(local) as clocks:
        ms = timed(step_user)
    gpu_launches = eng.launcher.launches - launches_before - W
    last = W + S - 1
    torch.cuda.synchronize()
    err_last = close(outs[last % 2], expect(last))
    ok_last = err_last < 1e-5 and bool((cnts[last % 2] == world).all())

    def run_user(src, k):
        eng.allreduce(src, outs[k % 2], cnts[k % 2])
        return outs[k % 2], cnts[k % 2]

    ver = verify_rounds(run_user)
    ok = all_ok(ok_last) and ver["ok"]
    algbw = msg_bytes / (ms * 1e-3) / 1e9
    value = algbw * world
    busbw = algbw * 2 * (world - 1) / world
    try:
        roof = roofline(world, eng.algo, msg_bytes, ms, False, os.path.join(ROOT, "MEASURED_PEAKS.json"))
    except Exception as ex:  # bookkeeping never takes the headline down
        roof = {"error": str(ex)[:200]}
    extras.update({"busbw_GBps": round(busbw, 2), "algbw_per_rank_GBps": round(algbw, 2), "roofline": roof,
                   "correct": ok, "verified": {"last_timed_round_rel_err": err_last, **ver},
                   "multicast": eng.has_multicast, "algo": eng.algo, "ctas": eng.launcher.ctas,
                   "kernel": "stream" if eng.launcher.stream_expected else "generic",
                   "timed_mode": "ordinary tensors (src/out/counts are plain torch tensors, a different source every step)"})

    # ------------------------------------------------------------------ registered buffers (zero-copy slots)
    L = eng.geo.slots
    if eng.output_view(0) is not None and not args.no_extras:
        try:  # (a secondary measurement never takes the headline line down)
            z_outs = [eng.output_view(r) for r in range(L)]
            z_ins = [eng.input_view(r) for r in range(L)]
            staged = all(v is not None for v in z_ins)
            base_round = eng.round
            for r in range(L):  # what sits in each slot during the timed loop
                if staged:
                    z_ins[(base_round + r) % L].copy_(srcs[r % NSRC])

            def step_zc(i: int) -> None:
                s = eng.round % L
                eng.allreduce(z_ins[s] if staged else srcs[i % NSRC], z_outs[s], cnts[i % 2])

            ms_z = timed(step_zc)
            # the last timed round used the content staged for its slot (or srcs[last % NSRC] without staging)
            last_slot = (eng.round - 1) % L
            k_last = ((eng.round - 1 - base_round) % L) if staged else last
            err_z = close(z_outs[last_slot], expect(k_last))

            def run_zc(src, k):
                s = eng.round % L
                if staged:
                    z_ins[s].copy_(src)  # the producer rewrites the slot every round
                eng.allreduce(z_ins[s] if staged else src, z_outs[s], cnts[k % 2])
                return z_outs[s], cnts[k % 2]

            ver_z = verify_rounds(run_zc)
            alg_z = msg_bytes / (ms_z * 1e-3) / 1e9
            try:
                roof_z = roofline(world, eng.algo, msg_bytes, ms_z, True, os.path.join(ROOT, "MEASURED_PEAKS.json"))
            except Exception as ex:
                roof_z = {"error": str(ex)[:200]}
            extras["registered_buffers"] = {
                "ms": round(ms_z, 4), "aggregate_algbw_GBps": round(alg_z * world, 2),
                "busbw_GBps": round(alg_z * 2 * (world - 1) / world, 2), "roofline": roof_z,
                "source_staged_in_slot": staged, "last_timed_round_rel_err": err_z,
                "verified": ver_z, "correct": bool(err_z < 1e-5) and ver_z["ok"],
                "note": "src and out ARE the engine's symmetric staging / sink slots (a registered buffer): no staging or flush copy"}
            ok = ok and extras["registered_buffers"]["correct"]
            del z_outs, z_ins
        except Exception as ex:
            extras["registered_buffers"] = {"error": f"{type(ex).__name__}: {str(ex)[:200]}"}

    extras["timeouts"] = eng.timeouts_seen()

    # ------------------------------------------------------------------ baselines on the same message
    if world > 1:
        buf = srcs[0].clone()
        nccl_ms = timed(lambda i: dist.all_reduce(buf))
        nccl_alg = msg_bytes / (nccl_ms * 1e-3) / 1e9
        extras["nccl"] = {"ms": round(nccl_ms, 4), "busbw_GBps": round(nccl_alg * 2 * (world - 1) / world, 2),
                          "aggregate_algbw_GBps": round(nccl_alg * world, 2), "buffer": "ordinary tensor, in place"}
        extras["vs_nccl"] = round(nccl_ms / ms, 3)
        del buf
        if not args.no_extras:
            extras["nccl_registered"] = nccl_registered_arm(n, dev, timed)
            extras["symm_mem"] = symm_mem_arm(n, dev, timed)
            rb = extras.get("registered_buffers")
            if rb:
                best = [v["ms"] for v in (extras["nccl_registered"], *extras["symm_mem"].values()) if isinstance(v, dict) and "ms" in v]
                rb["vs_nccl_ordinary"] = round(nccl_ms / rb["ms"], 3)
                if best:
                    rb["vs_best_registered_baseline"] = round(min(best) / rb["ms"], 3)
    eng.close()
    del srcs, outs, cnts, bases, scratch_src
    torch.cuda.empty_cache()

    # ------------------------------------------------------------------ end to end through the worker API
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(rank, world, local, n, mc, W, S, max_over_ranks, barrier)
        if world > 1 and e2e is not None:
            try:
                e2e["nccl_same_path"] = run_e2e_nccl(world, local, n, W, S, max_over_ranks, barrier)
                e2e["vs_nccl_same_path"] = round(e2e["nccl_same_path"]["ms_per_step"] / e2e["ms_per_step"], 3)
            except Exception as ex:
                e2e["nccl_same_path"] = {"error": str(ex)[:200]}

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
                       "cpu_placement": placement,  # rank 0's; says why when nothing was bound
                       "l2": f"inputs larger than L2: {NSRC} rotating {args.size_mb:g} MB sources, 2 outputs, 2 count buffers"},
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


def nccl_registered_arm(n, dev, timed) -> dict:
    """NCCL all_reduce on a buffer from NCCL's own allocator, registered with the communicator (what lets NCCL use
    its zero-copy NVLS path): the like-for-like comparator of the registered-buffer mode."""
    import torch
    import torch.distributed as dist

    try:
        backend = dist.group.WORLD._get_backend(dev)
        pool = torch.cuda.MemPool(backend.mem_allocator)
        with torch.cuda.use_mem_pool(pool):
            buf = torch.randn(n, device=dev)
        backend.register_mem_pool(pool)
        ms = timed(lambda i: dist.all_reduce(buf))
        world = dist.get_world_size()
        alg = n * 4 / (ms * 1e-3) / 1e9
        out = {"ms": round(ms, 4), "busbw_GBps": round(alg * 2 * (world - 1) / world, 2),
               "buffer": "ncclMemAlloc pool registered with the communicator"}
        try:
            backend.deregister_mem_pool(pool)
        except Exception:
            pass
        del buf
        return out
    except Exception as ex:
        return {"unavailable": f"{type(ex).__name__}: {str(ex)[:160]}"}


def symm_mem_arm(n, dev, timed) -> dict:
    """torch's own symmetric-memory collectives on a symm_mem buffer, in place — the closest existing sm_100
    competitors of the registered-buffer mode."""
    import torch
    import torch.distributed as dist

    out: dict = {}
    try:
        import torch.distributed._symmetric_memory as symm_mem

        group = dist.group.WORLD.group_name
        t = symm_mem.empty(n, dtype=torch.float32, device=dev)
        symm_mem.rendezvous(t, group)
        t.normal_()
        world = dist.get_world_size()
        for name, op in (("multimem_all_reduce_", lambda: torch.ops.symm_mem.multimem_all_reduce_(t, "sum", group)),
                         ("two_shot_all_reduce_", lambda: torch.ops.symm_mem.two_shot_all_reduce_(t, "sum", group))):
            try:
                op()
                torch.cuda.synchronize()
                t.normal_()
                ms = timed(lambda i: op())
                alg = n * 4 / (ms * 1e-3) / 1e9
                out[name] = {"ms": round(ms, 4), "busbw_GBps": round(alg * 2 * (world - 1) / world, 2)}
            except Exception as ex:
                out[name] = {"unavailable": f"{type(ex).__name__}: {str(ex)[:160]}"}
                try:
                    torch.cuda.synchronize()
                except Exception:
                    pass
        del t
    except Exception as ex:
        out["unavailable"] = f"{type(ex).__name__}: {str(ex)[:160]}"
    return out


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
               "peak_source": "900 GB/s per direction nominal", "fraction": round(busbw / peak, 3),
               "fraction_of_measured_770": round(busbw / (peak * 770.0 / 900.0), 3)}
    return out


def run_e2e(rank, world, local, n, mc, W, S, max_over_ranks, barrier):
    """Rounds through the public API: AllreduceMaster + DeviceWorker with a pinned-host source; every step copies its
    64 MB input host -> device and reads a digest of the result (checksum + smallest count) back to the host."""
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
    R = W + S
    digests = torch.zeros(R, 2, dtype=torch.float32).pin_memory()   # one row per round, written by async D2H copies
    d2h_done = [None] * R

    def source(req):
        return AllReduceInput(pinned[req.iteration % len(pinned)])

    def sink(out):
        # the step's result leaves the device: checksum of the reduced vector + smallest contributor count.  The copy
        # is asynchronous (the next round is already running); every row is waited for and checked at the end.
        it = out.iteration
        d = torch.stack([out.data.sum(), out.count.min().float()])
        digests[it].copy_(d, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        d2h_done[it] = ev
        if it == W - 1:
            ev.synchronize()
            marks["t0"] = time.perf_counter()
        if it == W + S - 1:
            ev.synchronize()
            marks["t1"] = time.perf_counter()

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
    for ev in d2h_done:
        if ev is not None:
            ev.synchronize()
    sec = max_over_ranks((marks["t1"] - marks["t0"]) / S)
    min_count = float(digests[:, 1].min())
    return {"value": round(world * n * 4 / sec / 1e9, 3), "unit": "GB/s", "ms_per_step": round(sec * 1e3, 4),
            "h2d_bytes_per_step": n * 4, "d2h_bytes_per_step": 8, "min_count": min_count,
            "rounds_with_full_count": int((digests[:, 1] == world).sum()), "rounds": R,
            "path": "AllreduceMaster + DeviceWorker(source=pinned host tensor, sink=device views) over the coordinator"}


def run_e2e_nccl(world, local, n, W, S, max_over_ranks, barrier):
    """The same end-to-end step with NCCL in the middle: pinned source -> H2D -> dist.all_reduce -> digest -> D2H,
    double-buffered like the worker (copy of step i+1 overlaps the collective of step i).  An anchor for the e2e number
    while the reference arm is unavailable."""
    import torch
    import torch.distributed as dist

    dev = torch.device("cuda", local)
    pinned = [torch.randn(n).pin_memory() for _ in range(3)]
    bufs = [torch.empty(n, device=dev) for _ in range(2)]
    R = W + S
    digests = torch.zeros(R, 1, dtype=torch.float32).pin_memory()
    copy_stream, comp = torch.cuda.Stream(device=dev), torch.cuda.current_stream()
    staged = [torch.cuda.Event() for _ in range(2)]
    freed = [torch.cuda.Event() for _ in range(2)]
    barrier()
    t0 = None
    for it in range(R):
        if it == W:
            torch.cuda.synchronize()
            t0 = time.perf_counter()
        s = it % 2
        with torch.cuda.stream(copy_stream):
            if it >= 2:
                copy_stream.wait_event(freed[s])
            bufs[s].copy_(pinned[it % 3], non_blocking=True)
            staged[s].record(copy_stream)
        comp.wait_event(staged[s])
        dist.all_reduce(bufs[s])
        digests[it].copy_(bufs[s].sum().reshape(1), non_blocking=True)
        freed[s].record(comp)
    torch.cuda.synchronize()
    sec = max_over_ranks((time.perf_counter() - t0) / S)
    barrier()
    return {"ms_per_step": round(sec * 1e3, 4), "value": round(world * n * 4 / sec / 1e9, 3), "unit": "GB/s",
            "path": "pinned host tensor -> H2D -> dist.all_reduce (NCCL) -> sum -> D2H, double-buffered"}


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
            "busbw_GBps": round(alg * 2 * (world - 1) / world, 2), "ctas_per_rank": cl.launcher.ctas,
            "kernel": cl.kernel_used()}


if __name__ == "__main__":
    sys.exit(main())
