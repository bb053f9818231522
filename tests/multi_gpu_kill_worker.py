"""A GPU worker dies in the middle of a job; the others carry on at full speed (run under torchrun, >= 2 GPUs).

Reference behaviour: the master drops a terminated worker from its map (AllreduceMaster.scala:46-52), workers skip
peers they do not know (AllreduceWorker.scala:213-216) and the thresholds absorb the missing contributions.  On
the GPU the coordinator's removal reaches the kernels as a dead-peer mask (refreshed InitWorkers ->
DeviceWorker -> DeviceAllreduce.set_dead_peers): nothing is pushed to the dead rank and nothing waits for it.

All data thresholds are 1 here on purpose: without the mask every round after the death would sit out the full
in-kernel timeout (2 s); with it the survivors finish the remaining rounds in milliseconds, with count P-1 on the
blocks the survivors own and zeros / count 0 on the dead rank's block.
No NCCL collective is issued after the death (it would hang): every rank reports through its exit code.
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch


def main() -> int:
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    from akka_allreduce_b200 import AllReduceInput, AllreduceConfig
    from akka_allreduce_b200.master import AllreduceMaster
    from akka_allreduce_b200.parallel.device_worker import DeviceWorker

    n, rounds, die_at, victim = 1 << 20, 400, 60, world - 1
    cfg = AllreduceConfig()
    cfg.worker.totalSize, cfg.worker.maxLag = world, 2
    cfg.data.dataSize, cfg.data.maxChunkSize, cfg.data.maxRound = n, 1 << 16, rounds - 1
    cfg.thresholds.thReduce = cfg.thresholds.thComplete = 1.0
    cfg.thresholds.thAllreduce = 0.5                       # the gate must stay satisfiable without the victim
    cfg.device.backend, cfg.device.gate, cfg.device.timeoutMs = "cuda", "device", 2000
    cfg.cluster.heartbeat, cfg.cluster.downAfter = 0.2, 2.0
    cfg.cluster.port = int(os.environ["MASTER_PORT"]) + 61
    master = AllreduceMaster(cfg) if rank == 0 else None
    if master:
        master.start()
    time.sleep(1.0 if rank else 0.0)
    ones = torch.ones(n, device="cuda")
    seen = {}

    def source(req):
        return AllReduceInput(ones)

    def sink(o):
        # sum of ones == contributors: every element must equal its own count
        ok = bool((o.data == o.count.float()).all())
        seen[o.iteration] = (time.perf_counter(), ok, int(o.count.min()), int(o.count.max()))
        if rank == victim and o.iteration == die_at:
            torch.cuda.synchronize()
            os._exit(0)                                    # no goodbye: the coordinator sees the connection drop

    w = DeviceWorker(source, sink, cfg, master_port=cfg.cluster.port, device=local)
    w.start()
    ok = w.wait(180)
    w.stop()
    if master:
        master.stop()
    if not ok or len(seen) != rounds:
        print(f"[rank {rank}] FAIL finished={ok} rounds seen {len(seen)}/{rounds}", flush=True)
        return 1
    bad = [r for r, v in seen.items() if not v[1]]
    if bad:
        print(f"[rank {rank}] FAIL sums do not match counts in rounds {bad[:8]}", flush=True)
        return 1
    # the tail of the job ran without the victim: count P-1 on the survivors' blocks, 0 on the victim's (nobody
    # reduces it) — or P-1 everywhere when the algorithm has no owners (one-shot / direct: every rank sums what it holds)
    tail = [seen[r] for r in range(rounds - 50, rounds)]
    if not all(v[2] in (0, world - 1) and v[3] == world - 1 for v in tail):
        print(f"[rank {rank}] FAIL tail counts {[(v[2], v[3]) for v in tail[:5]]}", flush=True)
        return 1
    tail_s = (seen[rounds - 1][0] - seen[rounds - 51][0]) / 50
    head_s = (seen[die_at - 5][0] - seen[5][0]) / (die_at - 10)
    slow = [r for r in range(1, rounds) if seen[r][0] - seen[r - 1][0] > 0.5]
    print(f"[rank {rank}] ok: {head_s * 1e3:.3f} ms/round before the death, {tail_s * 1e3:.3f} ms/round after; "
          f"rounds that took > 0.5 s: {slow}; timeouts flag {w.engine.timeouts_seen() if w.engine else None}", flush=True)
    # at most a handful of rounds may have sat out the 2 s timeout (between the death and the coordinator's notice)
    if len(slow) > 4 or tail_s > max(20 * head_s, 0.02):
        print(f"[rank {rank}] FAIL the survivors kept waiting for the dead rank", flush=True)
        return 1
    return 0


if __name__ == "__main__":
    rc = main()
    sys.stdout.flush()
    os._exit(rc)   # skip interpreter teardown: peers' mappings of a dead rank's memory are released by the driver
