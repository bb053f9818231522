#!/usr/bin/env python
"""NVLink byte counters around K rounds: does the fused round put its ALGORITHMIC bytes on the links, no more?

    torchrun --nproc-per-node 8 benchmarks/nvlink_bytes.py --size-mb 64 --rounds 200 [--algo nvls] [--zero-copy]

Every rank reads its GPU's NVLink data counters (NVML field values NVLINK_THROUGHPUT_DATA_TX / _RX, summed over the
links; KiB) before and after the rounds and compares the difference with what the algorithm has to move per rank
and direction:

    two-shot push   2 N (P-1)/P          NVLS   N + N/P  (payload; the multicast pulls are counted where they land)
    direct (P = 2)  N                    one-shot push  (P-1) N,  pull  P N in / N out per peer

Prints one JSON line with measured / algorithmic ratios per rank.  Counters that the platform does not expose come
back as an error string instead of a number.  (Written in round 2 after the GPU budget was spent: not yet run.)
"""
from __future__ import annotations

import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist


def nvlink_kib(index: int):
    """(tx_kib, rx_kib) summed over all links of GPU `index`, or an error string."""
    try:
        import pynvml as nv

        nv.nvmlInit()
        try:
            p = torch.cuda.get_device_properties(index)
            bus = f"{p.pci_domain_id:08x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
            h = nv.nvmlDeviceGetHandleByPciBusId(bus.encode())
            # scope id UINT_MAX = all links of the device
            vals = nv.nvmlDeviceGetFieldValues(h, [(nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, 0xFFFFFFFF),
                                                   (nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX, 0xFFFFFFFF)])
            out = []
            for v in vals:
                if v.nvmlReturn != 0:
                    return f"NVML field {v.fieldId}: return code {v.nvmlReturn}"
                out.append(int(v.value.ullVal))
            return tuple(out)
        finally:
            nv.nvmlShutdown()
    except Exception as ex:  # no NVML, no permission, field not supported
        return f"{type(ex).__name__}: {str(ex)[:120]}"


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--size-mb", type=float, default=64.0)
    ap.add_argument("--chunk-mb", type=float, default=1.0)
    ap.add_argument("--rounds", type=int, default=200)
    ap.add_argument("--algo", default="auto")
    ap.add_argument("--zero-copy", action="store_true")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from akka_allreduce_b200.parallel.engine import DeviceAllreduce, default_uds_dir

    n = int(a.size_mb * (1 << 20)) // 4
    eng = DeviceAllreduce(rank, world, n, max(1, int(a.chunk_mb * (1 << 20)) // 4), max_lag=1, th_allreduce=1.0,
                          device=local, uds_dir=default_uds_dir(), timeout_ms=10_000, algo=a.algo)
    src, out = torch.randn(n, device=dev), torch.empty(n, device=dev)
    cnt = torch.empty(n, dtype=torch.int32, device=dev)
    zc = a.zero_copy and eng.output_view(0) is not None

    def one_round():
        r = eng.round
        iv = eng.input_view(r) if zc else None
        eng.allreduce(iv if iv is not None else src, eng.output_view(r) if zc else out, cnt)

    for _ in range(5):
        one_round()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    before = nvlink_kib(local)
    for _ in range(a.rounds):
        one_round()
    torch.cuda.synchronize()
    dist.barrier()           # (NCCL's own barrier traffic is a few bytes)
    torch.cuda.synchronize()
    after = nvlink_kib(local)
    N, P = n * 4, world
    algo_bytes = {"twoshot": 2 * N * (P - 1) / P, "nvls": N + N / P, "direct": float(N),
                  "oneshot": float((P - 1) * N)}[eng.algo] * a.rounds
    if isinstance(before, tuple) and isinstance(after, tuple):
        tx, rx = (after[0] - before[0]) * 1024.0, (after[1] - before[1]) * 1024.0
        mine = [tx, rx, tx / algo_bytes if algo_bytes else 0.0, rx / algo_bytes if algo_bytes else 0.0]
    else:
        mine = [-1.0, -1.0, -1.0, -1.0]
    t = torch.tensor(mine, dtype=torch.float64, device=dev)
    allt = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(allt, t)
    if rank == 0:
        print(json.dumps({"nvlink_bytes": True, "workers": world, "bytes": N, "rounds": a.rounds, "algo": eng.algo,
                          "zero_copy": bool(zc), "algorithmic_bytes_per_rank_and_direction": algo_bytes,
                          "tx_over_algorithmic_per_rank": [round(float(x[2]), 4) for x in allt],
                          "rx_over_algorithmic_per_rank": [round(float(x[3]), 4) for x in allt],
                          "counter_error": None if isinstance(before, tuple) else before}), flush=True)
    eng.close()
    dist.barrier()
    dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
