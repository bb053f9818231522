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
from ..ops.geometry import Geometry


def oneshot_eligible(geo: Geometry, th_reduce: float, th_complete: float, max_lag: int, th_allreduce) -> bool:
    full = geo.req_reduce(th_reduce) >= geo.world and geo.req_complete(th_complete) >= geo.total_chunks
    return full and (max_lag >= 1 or th_allreduce is not None)


def default_uds_dir(tag: str | int | None = None) -> str:
    tag = tag if tag is not None else os.environ.get("MASTER_PORT", "0")
    d = os.path.join(tempfile.gettempdir(), f"aar_b200_{os.getuid()}_{tag}")
    os.makedirs(d, exist_ok=True)
    return d


class _DevicePointer:
    """Minimal __cuda_array_interface__ holder so torch can wrap a raw device address."""

    _TYPESTR = {torch.float32: "<f4", torch.bfloat16: "<u2", torch.int32: "<i4"}

    def __init__(self, ptr: int, numel: int, dtype: torch.dtype):
        self.dtype = dtype
        self.__cuda_array_interface__ = {"shape": (numel,), "typestr": self._TYPESTR[dtype], "data": (ptr, False),
                                         "version": 3, "strides": None}


class DeviceAllreduce:
    def __init__(self, rank: int, world: int, n: int, max_chunk: int, max_lag: int = 1,
                 dtype: torch.dtype = torch.float32, wire_dtype: torch.dtype | None = None,
                 th_reduce: float = 1.0, th_complete: float = 1.0, th_allreduce: float | None = None,
                 device: int | None = None, uds_dir: str | None = None, multicast: bool = True,
                 align_blocks: bool = True, ctas: int | None = None, timeout_ms: float = 2000.0,
                 scale: float = 1.0, normalize: bool = False, algo: str = "auto", kernel: str | None = None):
        self.rank, self.world = rank, world
        self.device_index = torch.cuda.current_device() if device is None else device
        self.device = torch.device("cuda", self.device_index)
        torch.cuda.set_device(self.device)
        self.geo = Geometry.make(n, max_chunk, world, max_lag, align_blocks)
        wire = wire_dtype or dtype
        elem = torch.empty((), dtype=dtype).element_size()
        if algo not in ("auto", "twoshot", "nvls", "oneshot", "direct"):
            raise ValueError(f"unknown algo {algo!r}")
        # One-shot (every rank sums all P copies itself) is the latency path for small messages; it needs
        # thresholds of 1 (ranks reduce independently, so they must all see the same contributors).
        oneshot_ok = oneshot_eligible(self.geo, th_reduce, th_complete, max_lag, th_allreduce)
        # crossover measured on B200s (profiles/raw/sweep_8gpu_oneshot_*): with the pull form (one multimem.ld_reduce
        # per 16 bytes of the P staged copies) one hop beats the NVLS two-shot round up to 2 MB on 8 GPUs (44 vs 54 us
        # at 2 MB, 69 vs 55 us at 4 MB) and on 2-4 GPUs
        small = n * elem <= int(os.environ.get("AAR_ONESHOT_BYTES", 2 << 20))
        if algo in ("oneshot", "direct") and not oneshot_ok:
            raise K.AarError(f"algo={algo} needs thReduce = thComplete = 1 and maxLag >= 1 (or the thAllreduce gate)")
        use_oneshot = algo == "oneshot" or (algo == "auto" and oneshot_ok and small and world > 1)
        # Direct: the one-shot idea streamed chunk by chunk for LARGE messages on two ranks — same bytes on the link
        # as two-shot (N each way), one hop instead of two.  Costs a second message-sized row per ring slot.
        align = 16 // elem
        direct_ok = (oneshot_ok and world == 2 and wire == dtype and scale == 1.0 and n % align == 0
                     and (max_chunk % align == 0 or max_chunk >= n) and self.geo.step % align == 0)
        if algo == "direct" and not direct_ok:
            raise K.AarError("algo=direct needs two ranks, thresholds 1, dtype == wire dtype and 16-byte aligned sizes")
        use_direct = algo == "direct" or (algo == "auto" and direct_ok and not use_oneshot
                                           and os.environ.get("AAR_NO_DIRECT") is None)
        self.launcher = RoundLauncher(self.geo, dtype, wire, th_reduce, th_complete, th_allreduce,
                                      ctas, timeout_ms, scale, normalize, os_cap=n if (use_oneshot or use_direct) else 0,
                                      kernel=kernel)
        lib = self.launcher.lib
        # every CTA of a launch must be resident: the generic kernel fits one per SM, the stream kernel two
        cap = K.check(lib.aar_max_ctas(K.KERNEL_STREAM if self.launcher.stream_expected else K.KERNEL_GENERIC, 1,
                                       self.device_index), "aar_max_ctas")
        self.launcher.ctas = max(1, min(self.launcher.ctas, cap))
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
        # NVLS moves N + N/P bytes per direction against 2N(P-1)/P for push: nothing to gain at 2 ranks, and at
        # 3-4 ranks the push kernel was measured faster from 16 MB up (profiles/sweep_4gpu_*).
        nvls_ok = bool(self.mc_ptr) and self.launcher.req_reduce >= world and (
            world >= 5 or (world >= 3 and n * elem < (16 << 20)))
        if algo == "nvls" and not (bool(self.mc_ptr) and world > 1 and self.launcher.req_reduce >= world):
            raise K.AarError("algo=nvls needs a multicast alias (NVLS), world > 1 and thReduce = 1")
        if use_oneshot:
            self.algo = "oneshot"
            if self.launcher.oneshot_pull < 0:  # stage locally, let the switch sum the P staged copies on the way in
                self.launcher.oneshot_pull = 1 if self.mc_ptr else 0
        elif use_direct:
            self.algo = "direct"
        elif algo == "nvls" or (algo == "auto" and nvls_ok and os.environ.get("AAR_NO_NVLS") is None):
            self.algo = "nvls"
        else:
            self.algo = "twoshot"
        self.launcher.algo = K.ALGO_NAMES[self.algo]
        self.scratch = torch.zeros(self.launcher.scratch_bytes, dtype=torch.uint8, device=self.device)
        self.host_flags = torch.zeros(4, dtype=torch.int32).pin_memory()
        off = self.launcher.force_offset
        self._force_word = self.scratch[off:off + 4].view(torch.int32)      # device word the kernels poll
        self._force_src = torch.zeros(1, dtype=torch.int32).pin_memory()
        off = self.launcher.dead_offset
        self._dead_word = self.scratch[off:off + 4].view(torch.int32)       # bit q: peer q is known dead
        self._dead_src = torch.zeros(1, dtype=torch.int32).pin_memory()
        self._ctl_stream = torch.cuda.Stream(device=self.device)
        self.round = 0
        self._inplace_src: torch.Tensor | None = None   # direct algorithm, in-place calls only
        self._launch = a = K.AarLaunch()
        self.launcher.prepare(a)
        a.rank, a.emu = rank, 0
        a.scratch[0] = self.scratch.data_ptr()
        for i in range(world):
            a.sym[i] = self.peer_ptrs[i]
        a.mc_base = self.mc_ptr
        a.host_flags = self.host_flags.data_ptr()

    # ---- zero-copy buffers ---------------------------------------------------------------------
    def _view(self, byte_offset: int, numel: int) -> torch.Tensor:
        dt = self.launcher.wire_dtype
        holder = _DevicePointer(self.peer_ptrs[self.rank] + byte_offset, numel, dt)
        t = torch.as_tensor(holder, device=self.device)
        return t.view(torch.bfloat16) if dt == torch.bfloat16 else t

    @property
    def zero_copy_ok(self) -> bool:
        """True when a round's staging/sink slot has the flat layout of the message (blocks unpadded) and
        no conversion is needed, so user tensors can alias the symmetric slots."""
        g, l = self.geo, self.launcher
        return g.step % 128 == 0 and l.dtype == l.wire_dtype and l.scale == 1.0 and self.algo != "oneshot"

    @property
    def partial_thresholds(self) -> bool:
        """True when a round may finish without some contribution (thReduce < 1 or thComplete < 1)."""
        return self.launcher.req_reduce < self.geo.world or self.launcher.req_complete < self.geo.total_chunks

    def input_view(self, round_: int) -> torch.Tensor | None:
        """Tensor aliasing the staging slot of ``round_`` (NVLS mode only: there the scatter stages the
        message locally and the switch pulls it).  A source written here needs no copy at all.
        Only handed out when both data thresholds are 1: the caller writes the slot outside the launch, i.e. past
        the kernel's back-pressure, and the only thing that keeps a slow owner from still reading round r - L out of
        it is that this rank could not have finished r - L without that owner's chunks (thComplete = 1)."""
        # (push mode stages nothing locally — except with a single rank, whose own scat row IS the staging slot)
        if not (self.zero_copy_ok and (self.algo == "nvls" or self.world == 1)):
            return None
        if self.partial_thresholds:
            return None
        l, g = self.launcher, self.geo
        esz = torch.empty((), dtype=l.wire_dtype).element_size()
        slot = round_ % g.slots
        return self._view(l.off_scat + slot * g.world * g.step * esz, g.n)

    def output_view(self, round_: int) -> torch.Tensor | None:
        """Tensor aliasing the sink slot of ``round_``: the reduced message lands here anyway, so using it
        as ``out`` skips the flush copy.  Valid until the slot is recycled ``maxLag + 1`` rounds later.
        Only handed out when both data thresholds are 1.  Below thComplete = 1 a straggler's late push could
        overwrite zeros the caller was told about; below thReduce = 1 the owners can reduce round r + maxLag + 1
        WITHOUT this rank and write its sink slot while the caller is still reading round r from it (with
        thReduce = 1 they cannot start before this rank has scattered that round, i.e. given the slot back).
        Found by the protocol search: ModelConfig(output_view=True, req_reduce < P)."""
        l, g = self.launcher, self.geo
        if not self.zero_copy_ok or l.req_complete < g.total_chunks or l.req_reduce < g.world:
            return None
        esz = torch.empty((), dtype=l.wire_dtype).element_size()
        slot = round_ % g.slots
        return self._view(l.off_sink + slot * g.world * g.step * esz, g.n)

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
        if self.partial_thresholds and self._overlap(src, out):
            # With thresholds of 1 a block of `out` is written only after its owner has received THIS rank's slice
            # of it, so the launch may run in place.  Below 1 the owner can finish without this rank: the result
            # (or the zeros of a missing chunk) could land on source elements that are still to be scattered.
            raise ValueError("allreduce in place (out overlaps src) needs thReduce = thComplete = 1; "
                             "pass a separate output tensor")
        if self.algo == "direct" and self._overlap(src, out):
            # In a two-shot round a block of `out` is written only after its owner holds this rank's slice of it; in
            # a direct round the sum warps write chunk m as soon as the PEER's chunk m has landed, whether or not
            # this rank's own push of chunk m has read the source yet.  In place (what the DDP hook does) therefore
            # goes through a private copy of the source, on the launch stream.
            s = stream if stream is not None else torch.cuda.current_stream(self.device)
            if self._inplace_src is None:
                self._inplace_src = torch.empty(g.n, dtype=self.launcher.dtype, device=self.device)
            with torch.cuda.stream(s):
                self._inplace_src.copy_(src.reshape(-1), non_blocking=True)
            src = self._inplace_src
        r = self.round if round_ is None else round_
        a = self._launch
        self.launcher.fill(a, r, inj)
        a.src[0] = src.data_ptr()
        a.out[0] = out.data_ptr()
        a.counts[0] = counts.data_ptr() if counts is not None else None
        self.launcher.launch(a, stream)
        if round_ is None:
            self.round += 1

    def allreduce_part(self, src: torch.Tensor, out: torch.Tensor, counts: torch.Tensor | None = None,
                       stream: torch.cuda.Stream | None = None) -> None:
        """One round over a SHORTER message on this engine's region: ``src.numel() <= n``.  The region, the flags and
        the ring keep the layout of the capacity ``n`` (row pitch, flags per block); only the partition follows the
        actual length (``AarLaunch.n_cap``).  Rounds of different lengths are ordinary consecutive rounds of the same
        ring — same tags, same slots, same back-pressure — which is how one heap serves DDP buckets of several sizes
        (``AllreduceCommHook(capacity_bytes=...)``).  Every rank must pass the same length for the same round.
        Needs ``align_blocks=True`` and a two-shot or NVLS engine."""
        g, l = self.geo, self.launcher
        m = src.numel()
        if m == g.n:
            return self.allreduce(src, out, counts, stream=stream)
        if g.reference_partition or self.algo not in ("twoshot", "nvls"):
            raise K.AarError("allreduce_part needs an engine with align_blocks=True and algo twoshot or nvls")
        if not 0 < m < g.n or out.numel() != m or (counts is not None and counts.numel() != m):
            raise ValueError(f"allreduce_part: 0 < src.numel() <= {g.n}, out and counts of the same length")
        if src.dtype != l.dtype or out.dtype != l.dtype:
            raise ValueError("dtype differs from the engine's")
        if self.partial_thresholds and self._overlap(src, out):
            raise ValueError("allreduce in place (out overlaps src) needs thReduce = thComplete = 1")
        sub = Geometry.make(m, g.max_chunk, g.world, g.slots - 1, True)
        a = self._launch
        l.fill(a, self.round, None)
        a.n, a.block_step, a.n_cap, a.block_step_cap = m, sub.step, g.n, g.step
        a.req_complete = sub.req_complete(l.th_complete)
        a.src[0], a.out[0] = src.data_ptr(), out.data_ptr()
        a.counts[0] = counts.data_ptr() if counts is not None else None
        try:
            l.launch(a, stream)
        finally:
            a.n, a.block_step, a.n_cap, a.block_step_cap = g.n, g.block_step_arg, 0, 0
            a.req_complete = l.req_complete
        self.round += 1

    @staticmethod
    def _overlap(x: torch.Tensor, y: torch.Tensor) -> bool:
        return (x.data_ptr() < y.data_ptr() + y.numel() * y.element_size()
                and y.data_ptr() < x.data_ptr() + x.numel() * x.element_size())

    # ---- the two halves of the round on their own (the scatter->reduce half is a reduce-scatter,
    # the broadcast half an all-gather; AllreduceWorker.scala:212-238 and :252-268) ---------------------------
    def block_range(self, rank: int | None = None) -> tuple[int, int]:
        """Element range of the block ``rank`` (default: this rank) owns."""
        r = self.rank if rank is None else rank
        return self.geo.start(r), self.geo.end(r)

    def _half_round(self, op: int, src: torch.Tensor, out: torch.Tensor, counts: torch.Tensor | None,
                    stream: torch.cuda.Stream | None) -> None:
        if self.algo in ("oneshot", "direct") or not self.launcher.stream_expected:
            raise K.AarError("reduce_scatter / all_gather need the stream kernel: aligned blocks (align_blocks=True), "
                             "dtype == wire dtype, and a two-shot or NVLS engine (algo != oneshot)")
        a = self._launch
        self.launcher.fill(a, self.round, None)
        a.op = op
        a.src[0], a.out[0] = src.data_ptr(), out.data_ptr()
        a.counts[0] = counts.data_ptr() if counts is not None else None
        try:
            self.launcher.launch(a, stream)
        finally:
            a.op = K.OP_ALLREDUCE
        self.round += 1

    def reduce_scatter(self, src: torch.Tensor, out_block: torch.Tensor, counts_block: torch.Tensor | None = None,
                       stream: torch.cuda.Stream | None = None) -> None:
        """``out_block`` <- this rank's block of the sum of every rank's ``src`` (the whole message);
        ``counts_block`` <- contributors per element of that block.  Uses one round of the ring."""
        lo, hi = self.block_range()
        if src.numel() != self.geo.n or out_block.numel() != hi - lo or (counts_block is not None and counts_block.numel() != hi - lo):
            raise ValueError(f"reduce_scatter: src must hold {self.geo.n} elements, out_block / counts_block {hi - lo}")
        if src.dtype != self.launcher.dtype or out_block.dtype != self.launcher.dtype:
            raise ValueError("dtype differs from the engine's")
        self._half_round(K.OP_REDUCE_SCATTER, src, out_block, counts_block, stream)

    def all_gather(self, src_block: torch.Tensor, out: torch.Tensor, counts: torch.Tensor | None = None,
                   stream: torch.cuda.Stream | None = None) -> None:
        """``out`` <- every rank's ``src_block`` at its owner's block range; ``counts`` <- 1 where a block arrived,
        0 (and zeros in ``out``) where it did not before the round closed (thComplete / timeout)."""
        lo, hi = self.block_range()
        if src_block.numel() != hi - lo or out.numel() != self.geo.n or (counts is not None and counts.numel() != self.geo.n):
            raise ValueError(f"all_gather: src_block must hold {hi - lo} elements, out / counts {self.geo.n}")
        if src_block.dtype != self.launcher.dtype or out.dtype != self.launcher.dtype:
            raise ValueError("dtype differs from the engine's")
        self._half_round(K.OP_ALL_GATHER, src_block, out, counts, stream)

    def enable_trace(self) -> torch.Tensor:
        """Per-CTA phase timestamps of the LAST launch: int64 [ctas+1, 8] of globaltimer ns
        (start, gate passed, S done, R done, C done, ns waiting in R, ns waiting in C, 0)."""
        self.trace = torch.zeros(self.launcher.ctas + 1, 16, dtype=torch.int64, device=self.device)
        self._launch.trace = self.trace.data_ptr()
        return self.trace

    def trace_summary(self) -> dict:
        """Phase timeline of the last traced launch, in microseconds from the first CTA's start, per role
        (role 0 = CTA runs all phases, 1 = scatter, 2 = reduce+broadcast, 3 = complete)."""
        torch.cuda.synchronize(self.device)
        t = self.trace[1:].cpu().double()  # work CTAs only (row 0 is the control CTA)
        t0 = float(t[:, 0].min())
        us = lambda x: round(float(x) / 1e3, 1)
        out = {"gate_us": us((t[:, 1] - t0).max()), "end_us": us((t[:, [4, 10]].max(dim=1).values - t0).max())}
        for role, name, col in ((0, "all", 4), (1, "S", 2), (2, "R", 3), (3, "C", 4), (4, "stream", 4)):
            rows = t[t[:, 8] == role]
            if len(rows) == 0:
                continue
            entry = {"ctas": len(rows), "done_us": [us((rows[:, col] - t0).min()), us((rows[:, col] - t0).max())]}
            if role in (0, 2):
                entry["wait_us"] = [us(rows[:, 5].min()), us(rows[:, 5].max())]
            if role in (0, 3):
                entry["wait_c_us"] = [us(rows[:, 6].min()), us(rows[:, 6].max())]
            if role == 4:  # stream kernel: one time line per warp role (first warp of the role in every CTA)
                def span(col):
                    v = rows[:, col][rows[:, col] > 0]
                    return [us((v - t0).min()), us((v - t0).max())] if len(v) else None
                entry = {"ctas": len(rows), "S_done_us": span(2), "R_done_us": span(3), "C_done_us": span(4),
                         "first_reduce_us": span(9), "cta_done_us": span(10),
                         "wait_r_us": [us(rows[:, 5].min()), us(rows[:, 5].max())],
                         "wait_c_us": [us(rows[:, 6].min()), us(rows[:, 6].max())]}
            if role == 0:
                entry["S_done_us"] = [us((rows[:, 2] - t0).min()), us((rows[:, 2] - t0).max())]
                entry["R_done_us"] = [us((rows[:, 3] - t0).min()), us((rows[:, 3] - t0).max())]
            out[name] = entry
        return out

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

    def set_dead_peers(self, mask: int) -> None:
        """Peers the coordinator reported gone (bit q = rank q): the kernels stop pushing to them and stop waiting
        for them — they count as absent contributors (AllreduceMaster.scala:46-52, AllreduceWorker.scala:213-216).
        Travels like the force tag: a 4-byte async copy that also reaches a kernel that is already waiting."""
        mask &= ~(1 << self.rank)
        if int(self._dead_src[0]) == mask:
            return
        self._ctl_stream.synchronize()
        self._dead_src[0] = mask
        with torch.cuda.stream(self._ctl_stream):
            self._dead_word.copy_(self._dead_src, non_blocking=True)
        # (the NVLS kernel needs no switch: a dead peer never raises a flag, so every chunk takes the unicast path)

    @property
    def dead_peers(self) -> int:
        return int(self._dead_src[0])

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
