"""Worker of the ``shm`` backend: one process per worker on one host, no GPU.

Control plane as for the other backends — the C++ coordinator over TCP hands out ids (``InitWorkers``), starts rounds
(``StartAllreduce``) and collects completion reports.  The data plane is the DEVICE protocol (round-tagged flags,
slot ring, back-pressure, thresholds firing at ">= required", forced catch-up) in a POSIX shared-memory segment all
workers of the job attach to (``csrc/host/shm_cluster.h``): what ``DeviceWorker`` does over NVLink, done over host
memory.  It is the CPU stand-in for the CUDA backend (same semantics, same callbacks, numpy arrays instead of device
tensors) and the fastest way to run the reference's configurations on one machine.

Same contract as the reference's worker (AllreduceWorker.scala:197-210): ``source(AllReduceInputRequest(iteration))``
gives the round's ``dataSize`` floats, ``sink(AllReduceOutput(data, count, iteration))`` receives the reduced message
and the per-element contributor counts.
"""
from __future__ import annotations

import threading
import time

import numpy as np

from .. import host
from ..api import AllReduceInput, AllReduceInputRequest, AllReduceOutput, DataSink, DataSource
from ..config import AllreduceConfig
from ..utils.logging import get_logger


class SinkStats:
    """What the built-in throughput/assert sink has seen (mirror of the CPU backend's ThroughputStats)."""

    def __init__(self) -> None:
        self.checkpoints = 0
        self.failures = 0
        self.last_mbytes_per_s = 0.0


class ShmWorker:
    def __init__(self, data_source: DataSource, data_sink: DataSink, config: AllreduceConfig,
                 master_host: str | None = None, master_port: int | None = None, shm_name: str | None = None):
        self.source, self.sink, self.cfg = data_source, data_sink, config
        self.master_host = master_host or config.cluster.host
        self.master_port = master_port or config.cluster.port
        self.shm_name = shm_name or f"/aar_b200_{self.master_port}"
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
        self._thread: threading.Thread | None = None
        self._error: BaseException | None = None
        self.cluster = None
        self.id = -1
        self.round = 0
        self.rounds_completed = 0
        self.timeouts = 0
        self.finished = threading.Event()

    # ---- coordinator messages (IO thread) ---------------------------------------------------------
    def _on_message(self, msg) -> None:
        with self._cv:
            if isinstance(msg, host.InitWorkers):
                if self._init is None:
                    self._init = msg
                # later InitWorkers only refresh the peer map (AllreduceWorker.scala:87-89): ids that are missing from
                # it belong to workers the master has removed (AllreduceMaster.scala:46-52) — the protocol's dead-peer
                # mask, as on the GPU: nothing is pushed to them, nothing waits for them.  A replacement that takes
                # over a vacated id shows up in the next refresh and clears its bit again.
                alive = set(msg.workers.keys())
                self._dead_mask = sum(1 << q for q in range(msg.worker_num) if q not in alive and q != msg.dest_id)
                if self.cluster is not None and self.id >= 0:
                    self.cluster.set_dead_peers(self.id, self._dead_mask)
            elif isinstance(msg, host.StartAllreduce):
                self._start_round = max(self._start_round, msg.round)
                cl, init = self.cluster, self._init
                if cl is not None and init is not None:
                    behind = self._start_round - init.max_lag - 1
                    if behind >= self.round:  # more than maxLag behind: close the old rounds now (catch-up)
                        cl.force_complete(self.id, behind)
            self._cv.notify_all()

    # ---- lifecycle ------------------------------------------------------------------------------------
    def start(self) -> None:
        self._node.listen("127.0.0.1", 0)
        self._node.begin()
        self._thread = threading.Thread(target=self._run_guarded, name="aar-shm-worker", daemon=True)
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
        if self.cluster is not None and self.id >= 0:
            self.cluster.force_complete(self.id, 2 ** 30)  # whatever the round loop is waiting for, let it go
        if self._thread is not None:
            self._thread.join(timeout=30)
        self._node.stop()
        self.cluster = None  # unmaps (and, for the creator, unlinks) the segment

    def _run_guarded(self) -> None:
        try:
            self._run()
        except Exception as e:  # surfaced by wait(); never swallowed silently
            self._error = e
            self.log.error("shm worker failed: %s", e)
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
        cfg = self.cfg
        self.id = init.dest_id
        gate_in_memory = cfg.device.gate == "device"
        sc = host.ShmConfig()
        sc.world, sc.n, sc.max_chunk, sc.max_lag = init.worker_num, init.data_size, init.max_chunk_size, init.max_lag
        sc.th_reduce, sc.th_complete = init.th_reduce, init.th_complete
        sc.th_allreduce = cfg.thresholds.thAllreduce if gate_in_memory else 0.0
        sc.timeout_ms, sc.scale, sc.normalize = cfg.device.timeoutMs, cfg.device.scale, cfg.device.normalize
        # the worker holding id 0 at start-up creates the segment, everybody else (and any later replacement) attaches
        joining_late = init.start_round > 0
        self.cluster = host.ShmCluster.attach(sc, self.shm_name, self.id == 0 and not joining_late,
                                              accept_orphan=joining_late)
        cl = self.cluster
        with self._cv:
            cl.set_dead_peers(self.id, self._dead_mask)   # (a replacement clears what its predecessor was told)
        self.log.info("----Actor id = %d (shared memory %s); peers = %d; thReduce = %s, thComplete = %s, maxLag = %d; "
                      "region %.1f KiB per rank", self.id, self.shm_name, sc.world, sc.th_reduce, sc.th_complete,
                      sc.max_lag, cl.region_bytes / 1024)
        # the coordinator says where the job ends (InitWorkers.max_round); an older coordinator leaves it to the config
        n, last = sc.n, (init.max_round if init.max_round >= 0 else cfg.data.maxRound)
        r = max(init.start_round, cl.issued(self.id))  # a replacement continues after its predecessor's last round
        if joining_late:
            # ... and not behind the job: the coordinator's round is already history by the time this process has
            # attached, and a worker that starts behind stays behind (thresholds below 1 do not wait for it).  The
            # segment knows where the front-runner is; starting there, the gate lets the others catch up with us.
            r = max(r, cl.next_round)
        self.round = r
        while r <= last:
            with self._cv:
                # the coordinator's StartAllreduce opens the round (always for the first one, so that nobody runs
                # before the whole job is up; for every round when the gate is the coordinator's)
                while not self._stop and (self._start_round < 0 or (not gate_in_memory and self._start_round < r)):
                    self._cv.wait(0.05)
                if self._stop:
                    break
            inp = self.source(AllReduceInputRequest(r))
            data = np.ascontiguousarray(inp.data if isinstance(inp, AllReduceInput) else inp, dtype=np.float32)
            if data.size != n:
                raise ValueError(f"Input data size {data.size} is different from initialization time {n}!")
            out, counts, info = cl.rank_round(self.id, r, data)
            self.timeouts += 1 if info["timed_out"] else 0
            self.sink(AllReduceOutput(out, counts, r))
            self._node.report_complete(self.id, r)
            r += 1
            self.round = r
            self.rounds_completed += 1


def start_shm_worker(config: AllreduceConfig, data_size: int, checkpoint: int = 50, assert_multiple: int = 0,
                     master_host: str | None = None, master_port: int | None = None) -> ShmWorker:
    """Built-in ramp source and throughput/assert sink (AllreduceWorker.scala:324-343) on the shm backend."""
    ramp = np.arange(data_size, dtype=np.float32)
    stats = SinkStats()
    log = get_logger("worker")
    state = {"tic": time.time()}

    def source(req: AllReduceInputRequest) -> AllReduceInput:
        return AllReduceInput(ramp)

    def sink(out: AllReduceOutput) -> None:
        if checkpoint <= 0 or out.iteration % checkpoint != 0 or out.iteration == 0:
            return
        elapsed = time.time() - state["tic"]
        print(f"----Data output at #{out.iteration} - {elapsed} s", flush=True)
        mb = data_size * 4.0 * checkpoint / 1e6
        stats.checkpoints += 1
        stats.last_mbytes_per_s = mb / max(elapsed, 1e-9)
        print(f"{mb:.1f} Mbytes in {elapsed:.1f} seconds at {stats.last_mbytes_per_s:.3f} MBytes/sec", flush=True)
        if assert_multiple > 0:
            ok = bool((out.data == ramp * assert_multiple).all()) and bool((out.count == assert_multiple).all())
            if not ok:
                stats.failures += 1
                log.error("assertion failed at #%d: output != input * %d or count != %d", out.iteration,
                          assert_multiple, assert_multiple)
        state["tic"] = time.time()

    w = ShmWorker(source, sink, config, master_host, master_port)
    w.stats = stats
    w.start()
    return w
