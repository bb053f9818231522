"""Launchers for the fused scatter -> reduce -> broadcast kernel.

``RoundLauncher`` owns everything a rank needs to run rounds on its GPU — scratch, mapped host
flags, the geometry — and fills the ``AarLaunch`` struct.  Two front ends use it:

* ``EmulatedCluster`` — all ``world`` ranks on ONE GPU in one cooperative launch (rank =
  ``blockIdx.y``).  Ranks that wait on each other must never be separate launches on one device,
  so this is how the full multi-rank protocol (thresholds, stragglers, lag ring) is exercised
  on a single B200, in tests and in the 1-GPU benchmark.
* ``akka_allreduce_b200.parallel.engine.DeviceAllreduce`` — one rank per process/GPU over the
  symmetric heap (NVLink peer pointers).
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field

import torch

from . import cuda as K
from .geometry import Geometry, default_ctas, gate_count


@dataclass
class Injection:
    """Straggler injection for one round (tests and the straggler benchmark)."""

    mute: int = 0                                   # bit r: rank r scatters nothing this round
    skip: dict = field(default_factory=dict)        # rank -> bitmask of destinations it skips
    delay_ns: dict = field(default_factory=dict)    # rank -> idle time before it starts


class RoundLauncher:
    def __init__(self, geo: Geometry, dtype: torch.dtype, wire_dtype: torch.dtype, th_reduce: float,
                 th_complete: float, th_allreduce: float | None, ctas: int | None, timeout_ms: float,
                 scale: float = 1.0, normalize: bool = False, os_cap: int = 0, algo: int = K.ALGO_TWOSHOT,
                 kernel: str | None = None):
        self.lib = K.lib()
        self.geo = geo
        self.dtype, self.wire_dtype = dtype, wire_dtype
        self.in_code, self.wire_code = K.dtype_code(dtype), K.dtype_code(wire_dtype)
        self.os_cap = os_cap
        self.algo = algo
        self.timeout_ns = int(timeout_ms * 1e6)
        self.scale, self.normalize = scale, normalize
        self.th_reduce, self.th_complete = th_reduce, th_complete
        self.req_reduce = geo.req_reduce(th_reduce)
        self.req_complete = geo.req_complete(th_complete)
        self.gate = 0 if th_allreduce is None else gate_count(th_allreduce, geo.world)
        # With every threshold at 1 a rank cannot finish round r without everybody's contribution to r, so
        # nobody can be more than one round ahead and the thAllreduce gate is implied by the data flow itself:
        # with at least two ring slots the in-kernel wait on the peers' completion flags is skipped.
        if (self.gate >= geo.world and self.req_reduce >= geo.world and self.req_complete >= geo.total_chunks
                and geo.slots >= 2 and os.environ.get("AAR_KEEP_GATE") is None):
            self.gate_implied, self.gate = True, 0
        else:
            self.gate_implied = False
        self.sym_bytes = self.lib.aar_sym_bytes(geo.n, geo.max_chunk, geo.world, geo.slots, self.wire_code, os_cap,
                                                geo.block_step_arg)
        self.flag_bytes = self.lib.aar_sym_flag_bytes(geo.n, geo.max_chunk, geo.world, geo.slots, self.wire_code,
                                                      os_cap, geo.block_step_arg)
        layout_args = (geo.n, geo.max_chunk, geo.world, geo.slots, self.wire_code, os_cap, geo.block_step_arg)
        self.off_scat = self.lib.aar_sym_scat_offset(*layout_args)
        self.off_sink = self.lib.aar_sym_sink_offset(*layout_args)
        self.scratch_bytes = self.lib.aar_scratch_bytes(geo.n, geo.max_chunk, geo.world, geo.block_step_arg)
        self.force_offset = self.lib.aar_scratch_force_offset(geo.n, geo.max_chunk, geo.world, geo.block_step_arg)
        self.dead_offset = self.lib.aar_scratch_dead_offset(geo.n, geo.max_chunk, geo.world, geo.block_step_arg)
        # Which two-shot kernel: "auto" = the stream kernel (warp tiles, two CTAs per SM) whenever the launch is
        # 16-byte aligned and needs no dtype conversion, else the generic one.  AAR_KERNEL=generic|stream forces it.
        self.kernel = K.KERNEL_NAMES[kernel or os.environ.get("AAR_KERNEL", "auto")]
        # ReduceBlock broadcast through the multicast alias (one multimem.st per 16 bytes instead of P stores)
        # on the two-shot push path; NVLS always broadcasts that way.  AAR_MC_BCAST=1.
        self.mc_bcast = int(os.environ.get("AAR_MC_BCAST", 0))
        self.launches = 0
        # the stream kernel needs 16-byte aligned block starts and chunk starts (and aligned tensors, checked per launch)
        align = 16 // min(torch.empty((), dtype=dtype).element_size(), torch.empty((), dtype=wire_dtype).element_size())
        self.stream_expected = (self.kernel != K.KERNEL_GENERIC and dtype == wire_dtype and geo.step % align == 0
                                and geo.n % align == 0 and (geo.max_chunk % align == 0 or geo.max_chunk >= geo.step))
        # work CTAs per rank; the library clamps it to what the device keeps resident (aar_max_ctas)
        self.ctas = ctas or default_ctas(geo.n * torch.empty((), dtype=dtype).element_size(), self.stream_expected)
        # one-shot flavour: 0 push (default), 1 pull with in-switch reduce, 2 pull with unicast loads (opt-in:
        # written after the round's multi-GPU budget was spent, validated in emulation only — see DESIGN.md 7)
        self.oneshot_pull = int(os.environ.get("AAR_ONESHOT_PULL", -1))   # -1: the engine picks (pull with NVLS)
        self._prepared: set[int] = set()   # launch structs (by id) whose static fields are filled
        self._injected: set[int] = set()   # launch structs that still carry last round's straggler injection

    def prepare(self, a: K.AarLaunch) -> None:
        """Fills everything that does not change from round to round (done once per launch struct)."""
        g = self.geo
        a.world, a.algo = g.world, self.algo
        a.in_dtype, a.wire_dtype = self.in_code, self.wire_code
        a.ctas, a.use_tma = self.ctas, 0 if os.environ.get("AAR_NO_TMA") else 1
        a.fast_fences = 1 if os.environ.get("AAR_FAST_FENCES") else 0
        # Interleaving completion items with reduce items (flush of delivered chunks between reduce items):
        # measured on 8 GPUs 1 GB 2.70 ms vs 2.93 ms without, 256 MB 0.70 vs 0.68, 64 MB equal; neutral on 2 GPUs.
        # On for messages of 512 MB and more.  AAR_OVERLAP=1/0 forces it.
        ov = os.environ.get("AAR_OVERLAP")
        a.overlap_flush = int(ov) if ov in ("0", "1") else (1 if g.n * (2 if self.wire_dtype == torch.bfloat16 else 4) >= (512 << 20) else 0)
        a.oneshot_pull = max(0, self.oneshot_pull)
        a.kernel, a.mc_bcast = self.kernel, self.mc_bcast
        a.self_publish = int(os.environ.get("AAR_SELF_PUBLISH", 0))
        a.push_chunks = int(os.environ.get("AAR_PUSH_CHUNKS", 0))
        a.role_s = int(os.environ.get("AAR_ROLE_S", 0))
        a.role_c = int(os.environ.get("AAR_ROLE_C", 0))
        a.nvls_unroll = int(os.environ.get("AAR_NVLS_UNROLL", 0))
        # largest work item: 256 KB, 64 KB for messages of 512 MB and more (measured on 8 GPUs at 1 GB: 64 KB
        # 2.93 ms, 256 KB 3.01 ms, 1 MB 3.11 ms — smaller windows keep the pages in flight within TLB reach)
        big = g.n * (2 if self.wire_dtype == torch.bfloat16 else 4) >= (512 << 20)
        a.part_kb = int(os.environ.get("AAR_PART_KB", 64 if (big and not self.stream_expected) else 0))
        a.part_sc_kb = int(os.environ.get("AAR_PART_SC_KB", 0))
        a.n, a.max_chunk, a.os_cap, a.block_step = g.n, g.max_chunk, self.os_cap, g.block_step_arg
        a.n_cap = a.block_step_cap = 0
        a.L = g.slots
        a.req_reduce, a.req_complete, a.gate_count = self.req_reduce, self.req_complete, self.gate
        a.normalize, a.scale = int(self.normalize), self.scale
        a.timeout_ns = self.timeout_ns
        a.mute_mask = 0
        for r in range(K.MAX_RANKS):
            a.skip_dest_mask[r] = 0
            a.delay_ns[r] = 0
        self._prepared.add(id(a))
        self._injected.discard(id(a))

    def fill(self, a: K.AarLaunch, round_: int, inj: Injection | None) -> None:
        """Per-round fields only: ring slot, round tag and (rarely) the straggler injection."""
        if id(a) not in self._prepared:
            self.prepare(a)
        a.slot, a.tag = round_ % self.geo.slots, round_ + 1
        if inj is not None or id(a) in self._injected:  # set this round's injection / clear last round's
            a.mute_mask = 0
            for r in range(K.MAX_RANKS):
                a.skip_dest_mask[r] = 0
                a.delay_ns[r] = 0
            self._injected.discard(id(a))
            if inj is not None:
                self._injected.add(id(a))
                a.mute_mask = inj.mute
                for r, m in inj.skip.items():
                    a.skip_dest_mask[r] = m
                for r, d in inj.delay_ns.items():
                    a.delay_ns[r] = int(d)

    def launch(self, a: K.AarLaunch, stream: torch.cuda.Stream | None = None) -> int:
        s = stream if stream is not None else torch.cuda.current_stream()
        n = K.check(self.lib.aar_allreduce_round(C.byref(a), C.c_void_p(s.cuda_stream)), "aar_allreduce_round")
        self.launches += n
        return n


class EmulatedCluster:
    """``world`` virtual ranks on one GPU; every round is ONE cooperative kernel launch."""

    def __init__(self, world: int, n: int, max_chunk: int, max_lag: int = 1, dtype=torch.float32,
                 wire_dtype=None, th_reduce: float = 1.0, th_complete: float = 1.0,
                 th_allreduce: float | None = None, ctas: int | None = None, align_blocks: bool = False,
                 timeout_ms: float = 2000.0, scale: float = 1.0, normalize: bool = False,
                 device: str | torch.device = "cuda:0", algo: str = "twoshot", kernel: str | None = None):
        self.device = torch.device(device)
        torch.cuda.set_device(self.device)
        self.geo = Geometry.make(n, max_chunk, world, max_lag, align_blocks)
        wire = wire_dtype or dtype
        lib = K.lib()
        if algo not in ("twoshot", "oneshot", "direct"):
            raise ValueError("the emulated cluster runs the two-shot, one-shot and direct kernels (NVLS needs real peers)")
        self.launcher = RoundLauncher(self.geo, dtype, wire, th_reduce, th_complete, th_allreduce,
                                      ctas, timeout_ms, scale, normalize,
                                      os_cap=n if algo in ("oneshot", "direct") else 0, algo=K.ALGO_NAMES[algo], kernel=kernel)
        L = self.launcher
        # all ranks share this GPU: (ctas + 1) x world CTAs must be resident at once
        cap = K.check(lib.aar_max_ctas(K.KERNEL_STREAM if L.stream_expected else K.KERNEL_GENERIC, world,
                                       self.device.index or 0), "aar_max_ctas")
        L.ctas = max(1, min(L.ctas, cap))
        self.world = world
        self.sym = torch.zeros(world, L.sym_bytes, dtype=torch.uint8, device=self.device)
        self.scratch = torch.zeros(world, L.scratch_bytes, dtype=torch.uint8, device=self.device)
        self.host_flags = torch.zeros(world * 4, dtype=torch.int32).pin_memory()
        self.outs = [torch.empty(n, dtype=dtype, device=self.device) for _ in range(world)]
        self.counts = [torch.empty(n, dtype=torch.int32, device=self.device) for _ in range(world)]
        self.round = 0
        self._forced = False
        self._a = K.AarLaunch()
        self.launcher.prepare(self._a)
        a = self._a
        a.rank, a.emu = 0, 1
        for i in range(world):
            a.sym[i] = self.sym[i].data_ptr()
            a.scratch[i] = self.scratch[i].data_ptr()
        a.mc_base = None
        a.host_flags = self.host_flags.data_ptr()

    def run_round(self, srcs, inj: Injection | None = None, outs=None, counts=None, round_: int | None = None,
                  force: bool = False, op: int = K.OP_ALLREDUCE):
        """One allreduce round over ``srcs`` (one tensor per virtual rank). Returns (outs, counts)."""
        r = self.round if round_ is None else round_
        outs = outs or self.outs
        counts = counts or self.counts
        a = self._a
        self.launcher.fill(a, r, inj)
        a.op = op
        for i in range(self.world):
            want = self.geo.size(i) if op == K.OP_ALL_GATHER else self.geo.n
            assert srcs[i].numel() == want and srcs[i].dtype == self.launcher.dtype and srcs[i].is_contiguous()
            a.src[i] = srcs[i].data_ptr()
            a.out[i] = outs[i].data_ptr()
            a.counts[i] = counts[i].data_ptr()
        if force or self._forced:  # the force tag is a device word inside each rank's scratch
            off = self.launcher.force_offset
            self.scratch[:, off:off + 4].view(torch.int32).fill_((r + 1) if force else 0)
            self._forced = force
        try:
            self.launcher.launch(a)
        finally:
            a.op = K.OP_ALLREDUCE
        if round_ is None:
            self.round += 1
        return outs, counts

    def run_part(self, srcs, inj: Injection | None = None):
        """One round over a SHORTER message (every ``srcs[i]`` holds m <= n elements) on this cluster's regions: the
        capacity layout of ``AarLaunch.n_cap`` (strides from n, partition from m).  Needs align_blocks=True.
        Returns (outs, counts) of m elements each."""
        g, L = self.geo, self.launcher
        m = srcs[0].numel()
        assert not g.reference_partition and 0 < m <= g.n
        if m == g.n:
            return self.run_round(srcs, inj)
        sub = Geometry.make(m, g.max_chunk, g.world, g.slots - 1, True)
        outs = [torch.empty(m, dtype=L.dtype, device=self.device) for _ in range(self.world)]
        counts = [torch.empty(m, dtype=torch.int32, device=self.device) for _ in range(self.world)]
        a = self._a
        L.fill(a, self.round, inj)
        a.n, a.block_step, a.n_cap, a.block_step_cap = m, sub.step, g.n, g.step
        a.req_complete = sub.req_complete(L.th_complete)
        for i in range(self.world):
            assert srcs[i].numel() == m and srcs[i].dtype == L.dtype and srcs[i].is_contiguous()
            a.src[i], a.out[i], a.counts[i] = srcs[i].data_ptr(), outs[i].data_ptr(), counts[i].data_ptr()
        try:
            L.launch(a)
        finally:
            a.n, a.block_step, a.n_cap, a.block_step_cap = g.n, g.block_step_arg, 0, 0
            a.req_complete = L.req_complete
        self.round += 1
        return outs, counts

    def set_dead_peers(self, masks: dict) -> None:
        """rank -> bit mask of the peers that rank is told are dead (the host-written word of its scratch)."""
        off = self.launcher.dead_offset
        for r in range(self.world):
            self.scratch[r, off:off + 4].view(torch.int32).fill_(int(masks.get(r, 0)))

    def kernel_used(self) -> str:
        """Which kernel the library picks for this cluster's launches: generic | stream | oneshot."""
        return {1: "generic", 2: "stream", 3: "oneshot"}[K.check(self.lib_kernel_query(), "aar_round_kernel")]

    def lib_kernel_query(self) -> int:
        a = self._a
        self.launcher.fill(a, self.round, None)
        for i in range(self.world):
            a.src[i] = a.src[i] or self.outs[i].data_ptr()
            a.out[i] = self.outs[i].data_ptr()
            a.counts[i] = self.counts[i].data_ptr()
        return self.launcher.lib.aar_round_kernel(C.byref(a))

    def timeouts_seen(self) -> list[int]:
        torch.cuda.synchronize(self.device)
        return [int(self.host_flags[4 * i + 2]) for i in range(self.world)]

    def done_tags(self) -> list[int]:
        torch.cuda.synchronize(self.device)
        return [int(self.host_flags[4 * i + 1]) for i in range(self.world)]
