"""Body of the multi-GPU tests: run under ``python -m torch.distributed.run --nproc-per-node N``.

Checks, on real NVLink peers: symmetric heap bring-up, thresholds-1 rounds vs an fp32 reference and vs
NCCL, lag-ring wrap, bf16, a muted straggler under thReduce < 1, and the DeviceWorker/AllreduceMaster path.
Exits non-zero on the first mismatch.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist


def main() -> int:
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from akka_allreduce_b200.ops.fused_allreduce import Injection
    from akka_allreduce_b200.parallel.engine import DeviceAllreduce, default_uds_dir

    def srcs_for(n, dtype, r):
        g = torch.Generator(device="cpu").manual_seed(77 + r)
        base = [torch.randn(n, generator=g).to(dtype) for _ in range(world)]
        return base

    failures = []

    def check(name, cond):
        if not cond:
            failures.append(name)
            print(f"[rank {rank}] FAIL {name}", flush=True)

    # --- thresholds 1, fp32, several shapes, ring wrap, both two-shot kernels, vs fp32 reference and NCCL
    shapes = [(1 << 20, 1 << 16, 1, True, "auto", None), (1 << 20, 1 << 16, 1, True, "twoshot", "stream"),
              (1 << 20, 1 << 16, 1, True, "twoshot", "generic"), (1 << 20, 1 << 16, 1, True, "nvls", "stream"),
              (1 << 20, 1 << 16, 1, True, "nvls", "generic"), (3 << 20, 1 << 18, 2, True, "nvls", "stream"),
              (778, 3, 3, False, "auto", None), (100003, 4096, 2, False, "auto", None),
              (100003, 4096, 2, False, "twoshot", None), (1 << 22, 1 << 18, 1, False, "auto", None),
              (256, 256, 1, False, "oneshot", None), (50000, 4096, 1, False, "auto", None)]
    for n, mc, lag, align, algo, kernel in shapes:
        try:
            eng = DeviceAllreduce(rank, world, n, mc, max_lag=lag, th_allreduce=1.0, device=local,
                                  uds_dir=default_uds_dir(), align_blocks=align, timeout_ms=20_000, algo=algo, kernel=kernel)
        except Exception as e:
            if algo == "nvls" and "multicast" in str(e):   # no NVLS on this box: every rank lands here together
                if rank == 0:
                    print(f"skip nvls n={n}: {e}", flush=True)
                continue
            raise
        out = torch.empty(n, device=dev)
        cnt = torch.empty(n, dtype=torch.int32, device=dev)
        for r in range(2 * (lag + 1) + 1):
            all_src = srcs_for(n, torch.float32, r)
            mine = all_src[rank].to(dev)
            eng.allreduce(mine, out, cnt)
            want = torch.zeros(n)
            for s in all_src:
                want += s
            nccl = mine.clone()
            dist.all_reduce(nccl)
            torch.cuda.synchronize()
            if eng.algo == "nvls":  # the switch fixes the summation order, not us: compare with a tolerance
                check(f"sum n={n} r={r} {algo}/{kernel}", torch.allclose(out.cpu(), want, rtol=1e-5, atol=1e-5))
            else:
                check(f"sum n={n} r={r} {algo}/{kernel}", torch.equal(out.cpu(), want))
            check(f"count n={n} r={r} {algo}/{kernel}", bool((cnt == world).all()))
            check(f"nccl n={n} r={r} {algo}/{kernel}", torch.allclose(out, nccl, rtol=1e-5, atol=1e-5))
        check(f"no timeouts n={n} {algo}/{kernel}", eng.timeouts_seen() == 0)
        # zero-copy views (registered buffers): the producer rewrites the staging slot every round
        if eng.output_view(0) is not None:
            L = eng.geo.slots
            for r in range(2 * L + 1):
                rr = eng.round
                all_src = srcs_for(n, torch.float32, 100 + r)
                iv, ov = eng.input_view(rr), eng.output_view(rr)
                mine = all_src[rank].to(dev)
                if iv is not None:
                    iv.copy_(mine)
                eng.allreduce(iv if iv is not None else mine, ov, cnt)
                want = torch.zeros(n)
                for s in all_src:
                    want += s
                torch.cuda.synchronize()
                check(f"zero-copy sum n={n} r={r} {algo}/{kernel}", torch.allclose(ov.cpu(), want, rtol=1e-5, atol=1e-5))
                check(f"zero-copy count n={n} r={r} {algo}/{kernel}", bool((cnt == world).all()))
            check(f"no timeouts (zero-copy) n={n} {algo}/{kernel}", eng.timeouts_seen() == 0)
        if rank == 0:
            print(f"ok fp32 n={n} mc={mc} lag={lag} multicast={eng.has_multicast} algo={eng.algo} "
                  f"kernel={'stream' if eng.launcher.stream_expected else 'generic'}", flush=True)
        eng.close()

    # --- the two halves of the round on their own: reduce_scatter, all_gather, and the two chained = allreduce
    n = 1 << 20
    for algo in ("twoshot", "nvls"):
        try:
            eng = DeviceAllreduce(rank, world, n, 1 << 16, max_lag=1, device=local, uds_dir=default_uds_dir(),
                                  timeout_ms=20_000, algo=algo, kernel="stream")
        except Exception as e:
            if algo == "nvls" and "multicast" in str(e):
                continue
            raise
        lo, hi = eng.block_range()
        blk = torch.empty(hi - lo, device=dev)
        bcnt = torch.empty(hi - lo, dtype=torch.int32, device=dev)
        out = torch.empty(n, device=dev)
        cnt = torch.empty(n, dtype=torch.int32, device=dev)
        for r in range(4):
            all_src = srcs_for(n, torch.float32, 40 + r)
            want = torch.zeros(n)
            for s in all_src:
                want += s
            eng.reduce_scatter(all_src[rank].to(dev), blk, bcnt)
            torch.cuda.synchronize()
            ok = torch.allclose(blk.cpu(), want[lo:hi], rtol=1e-5, atol=1e-5) and bool((bcnt == world).all())
            check(f"reduce_scatter {algo} r={r}", ok)
            eng.all_gather(blk, out, cnt)
            torch.cuda.synchronize()
            check(f"all_gather {algo} r={r}", torch.allclose(out.cpu(), want, rtol=1e-5, atol=1e-5) and bool((cnt == 1).all()))
        eng.allreduce(all_src[rank].to(dev), out, cnt)   # the ring is still good for whole rounds
        torch.cuda.synchronize()
        check(f"allreduce after halves {algo}", torch.allclose(out.cpu(), want, rtol=1e-5, atol=1e-5) and bool((cnt == world).all()))
        check(f"no timeouts halves {algo}", eng.timeouts_seen() == 0)
        if rank == 0:
            print(f"ok reduce_scatter / all_gather algo={eng.algo}", flush=True)
        eng.close()

    # --- bf16 wire
    n = 1 << 20
    for kernel in ("stream", "generic"):
        eng = DeviceAllreduce(rank, world, n, 1 << 16, dtype=torch.bfloat16, device=local, uds_dir=default_uds_dir(),
                              timeout_ms=20_000, kernel=kernel)
        all_src = srcs_for(n, torch.bfloat16, 5)
        out = torch.empty(n, dtype=torch.bfloat16, device=dev)
        eng.allreduce(all_src[rank].to(dev), out, None)
        torch.cuda.synchronize()
        want = sum((s.float() for s in all_src[1:]), all_src[0].float())
        check(f"bf16 {kernel}", torch.allclose(out.float().cpu(), want, rtol=2 ** -7, atol=1e-2))
        eng.close()

    # --- thReduce < 1 with a rank that never scatters (needs >= 3 ranks to be meaningful)
    if world >= 3:
        th = (world - 1) / world
        for kernel, mcb in (("stream", 0), ("generic", 0), ("stream", 1), ("generic", 1)):
            os.environ["AAR_MC_BCAST"] = str(mcb)   # 1: ReduceBlock broadcast with one multimem.st (needs NVLS)
            eng = DeviceAllreduce(rank, world, n, 1 << 16, th_reduce=th, th_complete=1.0, device=local,
                                  uds_dir=default_uds_dir(), timeout_ms=20_000, kernel=kernel)
            out = torch.empty(n, device=dev)
            cnt = torch.empty(n, dtype=torch.int32, device=dev)
            for r in range(3):
                all_src = srcs_for(n, torch.float32, 9 + r)
                eng.allreduce(all_src[rank].to(dev), out, cnt, inj=Injection(mute=1 << (world - 1)))
                torch.cuda.synchronize()
                want = torch.zeros(n)
                for s in all_src[:-1]:
                    want += s
                check(f"muted sum {kernel} mc_bcast={mcb} r={r}", torch.equal(out.cpu(), want))
                check(f"muted count {kernel} mc_bcast={mcb} r={r}", bool((cnt == world - 1).all()))
            eng.close()
        os.environ.pop("AAR_MC_BCAST", None)

    # --- worker API over the coordinator: AllreduceMaster on rank 0 + one DeviceWorker per rank
    from akka_allreduce_b200 import AllReduceInput, AllreduceConfig
    from akka_allreduce_b200.master import AllreduceMaster
    from akka_allreduce_b200.parallel.device_worker import DeviceWorker

    for gate in ("device", "host"):
        cfg = AllreduceConfig()
        cfg.worker.totalSize, cfg.worker.maxLag = world, 2
        cfg.data.dataSize, cfg.data.maxChunkSize, cfg.data.maxRound = 5000, 64, 30
        cfg.thresholds.thAllreduce = cfg.thresholds.thReduce = cfg.thresholds.thComplete = 1.0
        cfg.device.backend, cfg.device.gate, cfg.device.timeoutMs = "cuda", gate, 20_000
        cfg.cluster.port = int(os.environ["MASTER_PORT"]) + (31 if gate == "device" else 47)
        master = AllreduceMaster(cfg) if rank == 0 else None
        if master:
            master.start()
        ramp = torch.arange(5000, dtype=torch.float32)
        seen = {}

        def source(req):
            return AllReduceInput(ramp + req.iteration)

        def sink(o):
            seen[o.iteration] = bool((o.data.cpu() == world * (ramp + o.iteration)).all()) and bool((o.count == world).all())

        w = DeviceWorker(source, sink, cfg, master_port=cfg.cluster.port, device=local)
        dist.barrier()
        w.start()
        ok = w.wait(120)
        w.stop()
        if master:
            check(f"master finished ({gate})", master.wait(30))
            master.stop()
        check(f"worker api rounds ({gate})", ok and sorted(seen) == list(range(31)) and all(seen.values()))
        dist.barrier()
        if rank == 0:
            print(f"ok worker api gate={gate}", flush=True)

    t = torch.tensor([len(failures)], device=dev)
    dist.all_reduce(t)
    dist.barrier()
    dist.destroy_process_group()
    return 1 if int(t.item()) else 0


if __name__ == "__main__":
    sys.exit(main())
