#!/usr/bin/env python
"""How long do fences and peer flag stores take on this box?  (torchrun --nproc-per-node 2)

For each case every rank launches G probe CTAs at the same time; each CTA writes a burst, then times
fence.gpu, fence.sys, a relaxed flag store to the peer, a second fence.sys (= the flag's ack) and a
st.release.sys.  Prints median / max over the CTAs in microseconds.
"""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from akka_allreduce_b200.ops import cuda as K
from akka_allreduce_b200.parallel.engine import default_uds_dir


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    lib = K.lib()
    lib.aar_probe_fence.restype = C.c_int
    lib.aar_probe_fence.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p]
    heap = C.c_void_p()
    size = 256 << 20
    K.check(lib.aar_heap_create(C.byref(heap), rank, world, local, size, default_uds_dir().encode(), 1), "heap")
    bases = (C.c_void_p * world)()
    mc = C.c_void_p()
    lib.aar_heap_ptrs(heap, bases, C.byref(mc))
    me, peer = int(bases[rank]), int(bases[(rank + 1) % world])
    plain = torch.zeros(size // 2, dtype=torch.uint8, device=dev)
    rows = []
    for ctas in (1, 32, 96):
        for burst_kb in (0, 64, 512):
            for burst_where in ("local-sym", "peer-sym", "local-plain"):
                for flag_where in ("peer", "local"):
                    if burst_kb == 0 and burst_where != "local-sym":
                        continue
                    stamps = torch.zeros(ctas, 8, dtype=torch.int64, device=dev)
                    dst = {"local-sym": me + (64 << 20), "peer-sym": peer + (64 << 20), "local-plain": plain.data_ptr()}[burst_where]
                    fbase = peer if flag_where == "peer" else me
                    for it in range(3):
                        dist.barrier()
                        torch.cuda.synchronize()
                        K.check(lib.aar_probe_fence(dst, burst_kb << 10, fbase, fbase + 4096, it + 1, ctas, stamps.data_ptr(),
                                                    torch.cuda.current_stream().cuda_stream), "probe")
                        torch.cuda.synchronize()
                    t = stamps.cpu().double() / 1e3
                    d = {"burst_us": t[:, 1] - t[:, 0], "fence_gpu_us": t[:, 2] - t[:, 1], "fence_sys_us": t[:, 3] - t[:, 2],
                         "flag_st_us": t[:, 4] - t[:, 3], "flag_ack_fence_sys_us": t[:, 5] - t[:, 4], "st_release_sys_us": t[:, 6] - t[:, 5]}
                    row = {"ctas": ctas, "burst_kb_per_cta": burst_kb, "burst_to": burst_where, "flag_to": flag_where}
                    for k, v in d.items():
                        row[k] = [round(float(v.median()), 2), round(float(v.max()), 2)]
                    rows.append(row)
                    if rank == 0:
                        print(json.dumps(row), flush=True)
    dist.barrier()
    lib.aar_heap_destroy(heap)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
