"""DeviceWorker: the AllreduceWorker of the CUDA backend.

Same contract as the reference worker (AllreduceWorker.scala:7-8): built from a data-source and a
data-sink callback, registers with the master, runs the rounds the master starts, hands
``AllReduceOutput(data, count, iteration)`` to the sink and reports ``CompleteAllreduce``.  What
differs is where the work happens: scatter, threshold counting, reduction, broadcast and the
assembly of output + counts are ONE fused kernel per round over NVLink peer memory
(``csrc/cuda/fused_allreduce.cu``); this class only keeps the reference's round bookkeeping
(``round``, ``maxRound``, lag ring of ``maxLag + 1`` output buffers) and feeds the kernel.

Round gating (``device.gate``):

* ``device`` — the ``thAllreduce`` gate is evaluated inside the kernel from the completion flags
  every rank publishes in symmetric memory; the worker free-runs rounds ``0..maxRound`` after the
  master's first ``StartAllreduce`` and reports completions to the master asynchronously.
* ``host``   — every round waits for the master's ``StartAllreduce(r)`` over the coordinator, and a
  start that leaves this worker more than ``maxLag`` behind force-completes the old rounds (the
  reference's catch-up, AllreduceWorker.scala:100-106).
"""
from __future__ import annotations

import os
import threading
import time

import torch

from .. import host
from ..api import AllReduceInput, AllReduceInputRequest, AllReduceOutput, DataSink, DataSource
from ..config import AllreduceConfig
from ..utils.affinity import bind_to_gpu
from ..utils.logging import get_logger
from .engine import DeviceAllreduce, default_uds_dir

_DTYPES = {"float32": torch.float32, "bfloat16": torch.bfloat16}


class SinkStats:
    """What the built-in throughput/assert sink has seen (mirror of the CPU backend's ThroughputStats)."""

    def __init__(self) -> None:
        self.checkpoints = 0
        self.failures = 0
        self.last_mbytes_per_s = 0.0


class DeviceWorker:
    def __init__(self, data_source: DataSource, data_sink: DataSink, config: AllreduceConfig,
                 master_host: str | None = None, master_port: int | None = None, device: int | None = None,
                 uds_dir: str | None = None, align_blocks: bool = False):
        self.source, self.sink, self.cfg = data_source, data_sink, config
        self.master_host = master_host or config.cluster.host
        self.master_port = master_port or config.cluster.port
        self.device_index = device
        self.uds_dir = uds_dir
        self.align_blocks = align_blocks
        self.log = get_logger("worker", config.cluster.loglevel)
        opts = host.WorkerNodeOptions()
        opts.heartbeat_s = config.cluster.heartbeat
        self._node = host.WorkerNode(self.master_host, self.master_port, None, None, opts)
        self._node.set_handler(self._on_message)
        self._cv = threading.Condition()
        self._init = None
        self._start_round = -1
        self._dead_mask = 0
        self._stop = False
        self._thread = None
        self._error = None
        self.engine: DeviceAllreduce | None = None
        self.id = -1
        self.round = 0            # lowest round not yet flushed to the sink
        self.rounds_completed = 0
        self.h2d_bytes = 0
        self.finished = threading.Event()
        self._nvtx = os.environ.get("AAR_NVTX") is not None  # NVTX ranges per round phase for timeline tools

    # ---- coordinator messages (IO thread) ---------------------------------------------------------
    def _on_message(self, msg) -> None:
        with self._cv:
            if isinstance(msg, host.InitWorkers):
                if self._init is None:
                    self._init = msg
                # later InitWorkers only refresh the peer map (AllreduceWorker.scala:87-89): ids that are no longer
                # in it belong to workers the master has removed (AllreduceMaster.scala:46-52).  On the GPU that
                # becomes the kernels' dead-peer mask: no pushes to them, no waiting for them.  Membership of the
                # symmetric heap is fixed, so a removed id stays dead for the rest of the job.
                alive = set(msg.workers.keys())
                for q in range(msg.worker_num):
                    if q not in alive and q != msg.dest_id:
                        self._dead_mask |= 1 << q
                if self.engine is not None and self._dead_mask:
                    self.engine.set_dead_peers(self._dead_mask)
            elif isinstance(msg, host.StartAllreduce):
                self._start_round = max(self._start_round, msg.round)
                eng = self.engine
                if eng is not None and self._init is not None:
                    behind = self._start_round - self._init.max_lag - 1
                    if behind >= self.round:  # fell more than maxLag behind: close the old rounds now
                        eng.force_complete(behind)
            self._cv.notify_all()

    # ---- lifecycle ------------------------------------------------------------------------------------
    def start(self) -> None:
        self._node.listen("127.0.0.1", 0)
        self._node.begin()
        self._thread = threading.Thread(target=self._run_guarded, name="aar-device-worker", daemon=True)
        self._thread.start()

    def wait(self, timeout: float = 3600.0) -> bool:
        ok = self.finished.wait(timeout)
        if self._error is not None:
            raise self._error
        return ok

    def stop(self) -> None:
        with self._cv:
            self._stop = True
            self._cv.notify_all()
        if self._thread is not None:
            self._thread.join(timeout=30)
        self._node.stop()
        if self.engine is not None:
            self.engine.close()

    def _run_guarded(self) -> None:
        try:
            self._run()
        except Exception as e:  # surfaced by wait(); never swallowed silently
            self._error = e
            self.log.error("device worker failed: %s", e)
        finally:
            self.finished.set()

    # ---- the round loop (worker thread) -----------------------------------------------------------
    def _run(self) -> None:
        with self._cv:
            while self._init is None and not self._stop:
                self._cv.wait(0.1)
            if self._stop:
                return
            init = self._init
        cfg, dev_cfg = self.cfg, self.cfg.device
        self.id = init.dest_id
        world, n, L = init.worker_num, init.data_size, init.max_lag + 1
        if self.device_index is None:
            local = os.environ.get("LOCAL_RANK")
            self.device_index = int(local) if local is not None else self.id % torch.cuda.device_count()
        torch.cuda.set_device(self.device_index)
        placement = bind_to_gpu(self.device_index)  # this thread and the ones it starts: cores next to the GPU
        if placement and placement.get("bound"):
            self.log.info("----bound to the %d cpus next to cuda:%d (NUMA node %s, from %s)", placement["cpus"],
                          self.device_index, placement.get("numa_node"), placement.get("source"))
        dtype = _DTYPES[dev_cfg.dtype]
        gate_on_device = dev_cfg.gate == "device"
        self.engine = DeviceAllreduce(
            self.id, world, n, init.max_chunk_size, init.max_lag, dtype=dtype,
            th_reduce=init.th_reduce, th_complete=init.th_complete,
            th_allreduce=cfg.thresholds.thAllreduce if gate_on_device else None,
            device=self.device_index, uds_dir=self.uds_dir or default_uds_dir(self.master_port),
            align_blocks=self.align_blocks, ctas=dev_cfg.ctas or None, timeout_ms=dev_cfg.timeoutMs,
            scale=dev_cfg.scale, normalize=dev_cfg.normalize, algo=dev_cfg.algo)
        eng = self.engine
        dev = eng.device
        with self._cv:
            if self._dead_mask:  # a peer was removed while the heap was coming up
                eng.set_dead_peers(self._dead_mask)
        self.log.info("----Actor id = %d on cuda:%d; peers = %d; thReduce = %s, thComplete = %s, maxLag = %d; "
                      "symmetric region %.1f MiB, multicast %s", self.id, self.device_index, world, init.th_reduce,
                      init.th_complete, init.max_lag, eng.launcher.sym_bytes / 2 ** 20, eng.has_multicast)
        stream = torch.cuda.Stream(device=dev)
        copy_stream = torch.cuda.Stream(device=dev)  # H2D of round r+1 overlaps the kernel of round r
        staged = [torch.cuda.Event() for _ in range(L)]
        # Sink: the round's symmetric sink slot itself when the layout allows (valid until the slot is
        # recycled maxLag+1 rounds later), else a private ring of L output buffers.
        outs = [eng.output_view(k) if eng.output_view(k) is not None else torch.empty(n, dtype=dtype, device=dev)
                for k in range(L)]
        counts = [torch.empty(n, dtype=torch.int32, device=dev) for _ in range(L)]
        # Staging for host inputs: the symmetric staging slot when the kernel can read it in place (NVLS).
        stage = [eng.input_view(k) for k in range(L)]
        events = [torch.cuda.Event() for _ in range(L)]
        inflight: list[int] = []

        def flush_oldest() -> None:
            r = inflight.pop(0)
            slot = r % L
            events[slot].synchronize()
            # The sink runs with the round stream current: whatever GPU work it enqueues on the views (a checksum,
            # an optimizer step) is ordered before the launch that recycles this slot maxLag + 1 rounds later, so a
            # sink need not synchronise to be safe.
            with torch.cuda.stream(stream):
                self.sink(AllReduceOutput(outs[slot], counts[slot], r))
            self._node.report_complete(self.id, r)
            self.round = r + 1
            self.rounds_completed += 1

        # the coordinator says where the job ends (InitWorkers.max_round); without that, the local configuration
        last = init.max_round if init.max_round >= 0 else cfg.data.maxRound
        r = init.start_round  # 0 unless the master initialised us into a running job
        while r <= last:
            with self._cv:
                if self._stop:
                    break
                started = self._start_round
            if started < 0 or (not gate_on_device and started < r):
                if inflight:
                    flush_oldest()  # the master needs our CompleteAllreduce before it starts more rounds
                else:
                    with self._cv:
                        if self._start_round == started and not self._stop:
                            self._cv.wait(0.05)
                continue
            while len(inflight) >= L or (inflight and events[inflight[0] % L].query()):
                flush_oldest()
            slot = r % L
            if self._nvtx:
                torch.cuda.nvtx.range_push(f"aar.round{r}.fetch")
            inp = self.source(AllReduceInputRequest(r))
            if self._nvtx:
                torch.cuda.nvtx.range_pop()
            data = inp.data if isinstance(inp, AllReduceInput) else inp
            if not isinstance(data, torch.Tensor):
                data = torch.as_tensor(data)
            if data.numel() != n:
                raise ValueError(f"Input data size {data.numel()} is different from initialization time {n}!")
            if data.device != dev or data.dtype != dtype:
                if stage[slot] is None:
                    stage[slot] = torch.empty(n, dtype=dtype, device=dev)
                with torch.cuda.stream(copy_stream):
                    # the slot was last read by round r-L, which has been flushed (its event synchronised)
                    stage[slot].copy_(data, non_blocking=True)  # H2D from (ideally pinned) host memory
                    staged[slot].record(copy_stream)
                stream.wait_event(staged[slot])
                if data.device.type == "cpu":
                    self.h2d_bytes += data.numel() * data.element_size()
                data = stage[slot]
            with torch.cuda.stream(stream):
                if self._nvtx:
                    torch.cuda.nvtx.range_push(f"aar.round{r}.launch")
                eng.allreduce(data, outs[slot], counts[slot], round_=r, stream=stream)
                events[slot].record(stream)
                if self._nvtx:
                    torch.cuda.nvtx.range_pop()
            inflight.append(r)
            r += 1
        while inflight:
            flush_oldest()


# ---- built-in benchmark source/sink (AllreduceWorker.scala:324-343) ------------------------------
def start_device_worker(config: AllreduceConfig, data_size: int, checkpoint: int = 50, assert_multiple: int = 0,
                        master_host: str | None = None, master_port: int | None = None) -> DeviceWorker:
    state = {"ramp": None, "tic": time.time()}
    stats = SinkStats()
    log = get_logger("worker")

    def source(req: AllReduceInputRequest) -> AllReduceInput:
        if state["ramp"] is None:
            state["ramp"] = torch.arange(data_size, dtype=_DTYPES[config.device.dtype], device="cuda")
        return AllReduceInput(state["ramp"])

    def sink(out: AllReduceOutput) -> None:
        if out.iteration % checkpoint != 0 or out.iteration == 0:
            return
        elapsed = time.time() - state["tic"]
        mbytes = out.data.numel() * 4.0 * checkpoint / 1e6
        print(f"----Data output at #{out.iteration} - {elapsed} s")
        print("%2.1f Mbytes in %2.1f seconds at %4.3f MBytes/sec" % (mbytes, elapsed, mbytes / elapsed), flush=True)
        stats.checkpoints += 1
        stats.last_mbytes_per_s = mbytes / elapsed
        world = config.worker.totalSize
        if world > 1:  # bus bandwidth, nccl-tests convention: algbw x 2(P-1)/P
            esz = out.data.element_size()
            alg = out.data.numel() * esz * checkpoint / 1e9 / elapsed
            print("     = %.2f GB/s algorithm bandwidth, %.2f GB/s bus bandwidth over %d workers"
                  % (alg, alg * 2 * (world - 1) / world, world), flush=True)
        if assert_multiple > 0:
            ok = bool((out.data == state["ramp"] * assert_multiple).all()) and bool((out.count == assert_multiple).all())
            if not ok:
                stats.failures += 1
                log.error("Expected output should be multiple of input, counts equal to the multiple. "
                          "Check if all thresholds are 1 (iteration %d)", out.iteration)
        state["tic"] = time.time()

    w = DeviceWorker(source, sink, config, master_host, master_port)
    w.stats = stats
    w.start()
    return w
