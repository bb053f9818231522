"""DDP communication hook backed by the fused allreduce kernel.

Data-parallel gradient sync is the use the reference's primitive is meant for (source = gradients,
sink = summed gradients, ``count`` lets the sink normalise partial sums: SURVEY.md 2.3); the
reference has no training loop of its own.  Here every DDP gradient bucket becomes one allreduce
round: the hook launches the kernel on a side stream as soon as the bucket is ready — so the
transfer overlaps the rest of the backward pass — and returns a CUDA-aware future.  Gradients are
averaged in the kernel's reduce phase (division by the contributor count), so with thresholds < 1 a
straggler's missing contribution does not bias the mean.

    model = DDP(model, device_ids=[local_rank])
    hook = AllreduceCommHook(rank, world, local_rank)
    model.register_comm_hook(None, hook)
"""
from typing import Dict, Optional, Tuple

import torch
import torch.distributed as dist

from .engine import DeviceAllreduce, default_uds_dir


class AllreduceCommHook:
    def __init__(self, rank: int, world: int, device: int, max_chunk_bytes: int = 1 << 20, max_lag: int = 1,
                 th_reduce: float = 1.0, th_complete: float = 1.0, uds_dir: Optional[str] = None,
                 timeout_ms: float = 30_000.0, algo: str = "auto", ctas: Optional[int] = None,
                 capacity_bytes: Optional[int] = None):
        self.rank, self.world, self.device_index = rank, world, device
        self.__name__ = self.__qualname__ = "akka_allreduce_b200_fused_allreduce_hook"  # DDP logs the hook's name
        self.device = torch.device("cuda", device)
        self.max_chunk_bytes, self.max_lag = max_chunk_bytes, max_lag
        self.th_reduce, self.th_complete = th_reduce, th_complete
        self.uds_dir = uds_dir or default_uds_dir("ddp")
        self.timeout_ms, self.algo, self.ctas = timeout_ms, algo, ctas
        self.stream = torch.cuda.Stream(device=self.device)
        # One symmetric heap for ALL bucket sizes up to `capacity_bytes` (per dtype), instead of one heap per distinct
        # size: buckets run as rounds of different lengths on the region of the largest (DeviceAllreduce.allreduce_part).
        # Pick it >= DDP's bucket_cap_mb (and >= the largest single parameter).  None = one engine per size.
        self.capacity_bytes = capacity_bytes
        self._shared: Dict[torch.dtype, DeviceAllreduce] = {}
        self._engines: Dict[Tuple[int, torch.dtype], DeviceAllreduce] = {}
        self._outs: Dict[Tuple[int, torch.dtype], torch.Tensor] = {}  # only used with thresholds below 1
        self.buckets_reduced = 0
        self.kernel_launches = 0

    def _shared_engine(self, numel: int, dtype: torch.dtype) -> Optional[DeviceAllreduce]:
        """The capacity engine of this dtype if the bucket fits (created collectively on first use)."""
        if self.capacity_bytes is None:
            return None
        esz = torch.empty((), dtype=dtype).element_size()
        cap = self.capacity_bytes // esz // 128 * 128
        if numel > cap or cap <= 0:
            return None
        eng = self._shared.get(dtype)
        if eng is None:
            # two-shot / NVLS only (picked from the capacity): the one-shot and direct forms keep whole messages in rows
            # of their own and are not laid out for shorter rounds
            algo = self.algo if self.algo in ("twoshot", "nvls") else ("nvls" if self.world >= 5 else "twoshot")
            try:
                eng = DeviceAllreduce(self.rank, self.world, cap, max(1, self.max_chunk_bytes // esz), self.max_lag,
                                      dtype=dtype, th_reduce=self.th_reduce, th_complete=self.th_complete,
                                      device=self.device_index, uds_dir=self.uds_dir, align_blocks=True,
                                      timeout_ms=self.timeout_ms, normalize=True, algo=algo, ctas=self.ctas)
            except Exception:
                if algo != "nvls":
                    raise
                eng = DeviceAllreduce(self.rank, self.world, cap, max(1, self.max_chunk_bytes // esz), self.max_lag,
                                      dtype=dtype, th_reduce=self.th_reduce, th_complete=self.th_complete,
                                      device=self.device_index, uds_dir=self.uds_dir, align_blocks=True,
                                      timeout_ms=self.timeout_ms, normalize=True, algo="twoshot", ctas=self.ctas)
            self._shared[dtype] = eng
        return eng

    def _engine(self, numel: int, dtype: torch.dtype) -> DeviceAllreduce:
        key = (numel, dtype)
        eng = self._engines.get(key)
        if eng is None:
            # Every rank meets the same bucket sizes in the same order, so engines (and their symmetric
            # regions) are created collectively on first use.  Blocks are aligned so that any bucket
            # length runs on the 16-byte paths.
            esz = torch.empty((), dtype=dtype).element_size()
            eng = DeviceAllreduce(self.rank, self.world, numel, max(1, self.max_chunk_bytes // esz), self.max_lag,
                                  dtype=dtype, th_reduce=self.th_reduce, th_complete=self.th_complete,
                                  device=self.device_index, uds_dir=self.uds_dir, align_blocks=True,
                                  timeout_ms=self.timeout_ms, normalize=True, algo=self.algo, ctas=self.ctas)
            self._engines[key] = eng
        return eng

    def __call__(self, state, bucket: dist.GradBucket) -> torch.futures.Future[torch.Tensor]:
        buf = bucket.buffer()
        shared = self._shared_engine(buf.numel(), buf.dtype)
        eng = shared if shared is not None else self._engine(buf.numel(), buf.dtype)
        run = eng.allreduce_part if shared is not None else eng.allreduce
        fut: torch.futures.Future[torch.Tensor] = torch.futures.Future(devices=[self.device])
        self.stream.wait_stream(torch.cuda.current_stream(self.device))  # the bucket's gradients are final
        with torch.cuda.stream(self.stream):
            before = eng.launcher.launches
            if eng.partial_thresholds:
                # a round that may finish without this rank must not write over gradients still to be scattered
                out = self._outs.get((buf.numel(), buf.dtype))
                if out is None:
                    out = self._outs[(buf.numel(), buf.dtype)] = torch.empty_like(buf)
                run(buf, out, None, stream=self.stream)
                buf.copy_(out, non_blocking=True)
            else:
                run(buf, buf, None, stream=self.stream)  # in place: mean over the contributors
            self.kernel_launches += eng.launcher.launches - before
            buf.record_stream(self.stream)
            fut.set_result(buf)  # captures the side stream's position
        self.buckets_reduced += 1
        return fut

    def close(self) -> None:
        for eng in list(self._engines.values()) + list(self._shared.values()):
            eng.close()
        self._engines.clear()
        self._shared.clear()
        self._outs.clear()


class ShmAllreduceCommHook:
    """The same hook for CPU data parallelism on one host: gradient buckets are averaged by the device protocol in
    POSIX shared memory (``host.ShmCluster``) instead of the fused kernel over NVLink.

    DDP itself still needs a process group for its own bookkeeping (``gloo``); the gradient traffic does not go
    through it.  One shared segment per bucket size, created by rank 0 and attached by the others on first use (every
    rank meets the same bucket sizes in the same order).  The call blocks until this rank's round is complete — there
    is no second stream to overlap with on a CPU — and returns an already-completed future.  fp32 on the wire.

        model = DDP(model)                                   # CPU tensors, gloo process group
        model.register_comm_hook(None, ShmAllreduceCommHook(rank, world, name="/job42"))
    """

    def __init__(self, rank: int, world: int, name: str, max_chunk_bytes: int = 1 << 18, max_lag: int = 1,
                 th_reduce: float = 1.0, th_complete: float = 1.0, timeout_ms: float = 30_000.0):
        self.rank, self.world, self.name = rank, world, name
        self.__name__ = self.__qualname__ = "akka_allreduce_b200_shm_allreduce_hook"
        self.max_chunk_bytes, self.max_lag = max_chunk_bytes, max_lag
        self.th_reduce, self.th_complete, self.timeout_ms = th_reduce, th_complete, timeout_ms
        self._clusters: Dict[int, Tuple[object, int]] = {}   # numel -> (cluster, next round)
        self.buckets_reduced = 0

    def _cluster(self, numel: int):
        entry = self._clusters.get(numel)
        if entry is None:
            from .. import host

            c = host.ShmConfig()
            c.world, c.n, c.max_chunk, c.max_lag = self.world, numel, max(1, self.max_chunk_bytes // 4), self.max_lag
            c.th_reduce, c.th_complete, c.th_allreduce = self.th_reduce, self.th_complete, 1.0
            c.timeout_ms, c.normalize = self.timeout_ms, True       # mean over the contributors, like the kernel
            entry = (host.ShmCluster.attach(c, f"{self.name}_{numel}", self.rank == 0), 0)
        return entry

    def __call__(self, state, bucket: dist.GradBucket) -> torch.futures.Future[torch.Tensor]:
        buf = bucket.buffer()
        cluster, round_ = self._cluster(buf.numel())
        src = buf.detach().to(torch.float32).contiguous().numpy()
        out, _counts, _info = cluster.rank_round(self.rank, round_, src)
        self._clusters[buf.numel()] = (cluster, round_ + 1)
        buf.copy_(torch.from_numpy(out).to(buf.dtype))
        self.buckets_reduced += 1
        fut: torch.futures.Future[torch.Tensor] = torch.futures.Future()
        fut.set_result(buf)
        return fut

    def close(self) -> None:
        self._clusters.clear()   # unmaps the segments (rank 0 unlinks them)
