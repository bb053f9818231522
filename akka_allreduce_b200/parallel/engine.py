"""DeviceAllreduce: one rank of the fused allreduce on its own GPU.

Owns the rank's slice of the symmetric heap (``csrc/cuda/sym_heap.cu``), its scratch and mapped
host flags, and launches one fused kernel per round.  The peers are plain device pointers into
the other ranks' regions, reached over NVLink; no NCCL call is involved anywhere on this path.

Rendezvous needs nothing but a directory every rank can see (unix sockets carry the VMM file
descriptors), so the engine can be brought up by ``torchrun`` ranks, by workers that got their
ids from the coordinator, or by hand.
"""
from __future__ import annotations

import ctypes as C
import os
import tempfile

import torch

from ..ops import cuda as K
from ..ops.fused_allreduce import Injection, RoundLauncher
from ..ops.geometry import Geometry, default_ctas


def default_uds_dir(tag: str | int | None = None) -> str:
    tag = tag if tag is not None else os.environ.get("MASTER_PORT", "0")
    d = os.path.join(tempfile.gettempdir(), f"aar_b200_{os.getuid()}_{tag}")
    os.makedirs(d, exist_ok=True)
    return d


class DeviceAllreduce:
    def __init__(self, rank: int, world: int, n: int, max_chunk: int, max_lag: int = 1,
                 dtype: torch.dtype = torch.float32, wire_dtype: torch.dtype | None = None,
                 th_reduce: float = 1.0, th_complete: float = 1.0, th_allreduce: float | None = None,
                 device: int | None = None, uds_dir: str | None = None, multicast: bool = True,
                 align_blocks: bool = True, ctas: int | None = None, timeout_ms: float = 2000.0,
                 scale: float = 1.0, normalize: bool = False, algo: str = "auto"):
        self.rank, self.world = rank, world
        self.device_index = torch.cuda.current_device() if device is None else device
        self.device = torch.device("cuda", self.device_index)
        torch.cuda.set_device(self.device)
        self.geo = Geometry.make(n, max_chunk, world, max_lag, align_blocks)
        wire = wire_dtype or dtype
        elem = torch.empty((), dtype=dtype).element_size()
        self.launcher = RoundLauncher(self.geo, dtype, wire, th_reduce, th_complete, th_allreduce,
                                      ctas or default_ctas(n * elem), timeout_ms, scale, normalize)
        lib = self.launcher.lib
        if algo not in ("auto", "twoshot", "nvls"):
            raise ValueError(f"unknown algo {algo!r}")
        self._heap = C.c_void_p()
        K.check(lib.aar_heap_create(C.byref(self._heap), rank, world, self.device_index, self.launcher.sym_bytes,
                                    (uds_dir or default_uds_dir()).encode(), int(multicast)), "aar_heap_create")
        bases = (C.c_void_p * world)()
        mc = C.c_void_p()
        K.check(lib.aar_heap_ptrs(self._heap, bases, C.byref(mc)), "aar_heap_ptrs")
        self.peer_ptrs = [int(bases[i]) for i in range(world)]
        self.mc_ptr = mc.value  # None when the driver exposes no multicast
        # NVLS reduces inside the switch over ALL ranks, so it is only eligible when no contribution may be
        # left out (thReduce = 1); chunks that time out fall back to unicast pulls inside the same kernel.
        nvls_ok = bool(self.mc_ptr) and world > 1 and self.launcher.req_reduce >= world
        if algo == "nvls" and not nvls_ok:
            raise K.AarError("algo=nvls needs a multicast alias (NVLS), world > 1 and thReduce = 1")
        self.algo = "nvls" if (algo == "nvls" or (algo == "auto" and nvls_ok and os.environ.get("AAR_NO_NVLS") is None)) else "twoshot"
        self.launcher.algo = K.ALGO_NAMES[self.algo]
        self.scratch = torch.zeros(self.launcher.scratch_bytes, dtype=torch.uint8, device=self.device)
        self.host_flags = torch.zeros(4, dtype=torch.int32).pin_memory()
        off = self.launcher.force_offset
        self._force_word = self.scratch[off:off + 4].view(torch.int32)      # device word the kernels poll
        self._force_src = torch.zeros(1, dtype=torch.int32).pin_memory()
        self._ctl_stream = torch.cuda.Stream(device=self.device)
        self.round = 0
        self._launch = a = K.AarLaunch()
        self.launcher.prepare(a)
        a.rank, a.emu = rank, 0
        a.scratch[0] = self.scratch.data_ptr()
        for i in range(world):
            a.sym[i] = self.peer_ptrs[i]
        a.mc_base = self.mc_ptr
        a.host_flags = self.host_flags.data_ptr()

    @property
    def has_multicast(self) -> bool:
        return bool(self.mc_ptr)

    def allreduce(self, src: torch.Tensor, out: torch.Tensor, counts: torch.Tensor | None = None,
                  round_: int | None = None, stream: torch.cuda.Stream | None = None,
                  inj: Injection | None = None) -> None:
        """Enqueues one round on ``stream``: ``out`` <- sum over the contributing ranks of ``src``,
        ``counts`` <- contributors per element.  Rounds must be issued in the same order on every rank."""
        g = self.geo
        if src.numel() != g.n or out.numel() != g.n:
            raise ValueError(f"Input data size {src.numel()} is different from initialization time {g.n}!")
        if src.dtype != self.launcher.dtype or out.dtype != self.launcher.dtype:
            raise ValueError("dtype differs from the engine's")
        r = self.round if round_ is None else round_
        a = self._launch
        self.launcher.fill(a, r, inj)
        a.src[0] = src.data_ptr()
        a.out[0] = out.data_ptr()
        a.counts[0] = counts.data_ptr() if counts is not None else None
        self.launcher.launch(a, stream)
        if round_ is None:
            self.round += 1

    def enable_trace(self) -> torch.Tensor:
        """Per-CTA phase timestamps of the LAST launch: int64 [ctas+1, 8] of globaltimer ns
        (start, gate passed, S done, R done, C done, ns waiting in R, ns waiting in C, 0)."""
        self.trace = torch.zeros(self.launcher.ctas + 1, 16, dtype=torch.int64, device=self.device)
        self._launch.trace = self.trace.data_ptr()
        return self.trace

    def trace_summary(self) -> dict:
        torch.cuda.synchronize(self.device)
        t = self.trace[1:].cpu().double()  # work CTAs only (row 0 is the control CTA)
        t0 = float(t[:, 0].min())
        us = lambda x: round(float(x) / 1e3, 1)
        return {"gate_us": us((t[:, 1] - t0).max()), "S_done_us": [us((t[:, 2] - t0).min()), us((t[:, 2] - t0).max())],
                "R_done_us": [us((t[:, 3] - t0).min()), us((t[:, 3] - t0).max())],
                "C_done_us": [us((t[:, 4] - t0).min()), us((t[:, 4] - t0).max())],
                "wait_in_R_us": [us(t[:, 5].min()), us(t[:, 5].max())], "wait_in_C_us": [us(t[:, 6].min()), us(t[:, 6].max())],
                "publish_in_S_us": [us(t[:, 7].min()), us(t[:, 7].max())],
                "S_done_sorted_us": [us(x) for x in sorted((t[:, 2] - t0).tolist())[:: max(1, len(t) // 12)]]}

    def force_complete(self, round_: int) -> None:
        """Host side of the catch-up path: any wait of rounds <= ``round_`` closes immediately.  The tag
        travels to the device as a 4-byte async copy on a side stream, so it also reaches kernels that
        are already running."""
        if int(self._force_src[0]) >= round_ + 1:
            return
        self._ctl_stream.synchronize()
        self._force_src[0] = round_ + 1
        with torch.cuda.stream(self._ctl_stream):
            self._force_word.copy_(self._force_src, non_blocking=True)

    def last_done_round(self) -> int:
        return int(self.host_flags[1]) - 1

    def timeouts_seen(self) -> int:
        return int(self.host_flags[2])

    def close(self) -> None:
        if self._heap:
            torch.cuda.synchronize(self.device)
            self.launcher.lib.aar_heap_destroy(self._heap)
            self._heap = C.c_void_p()

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass
