"""AllreduceWorker: a worker process and its launcher.

Counterpart of the reference's worker actor + ``object AllreduceWorker`` launcher
(AllreduceWorker.scala:7-8,303-362).  A worker is built from a data-source callback and a
data-sink callback, registers with the master, and then runs rounds as the master starts them.

* ``backend="cpu"``  — the whole protocol (scatter / reduce / broadcast messages) runs in the C++
  host engine over TCP, like the reference's actors over Akka remoting.
* ``backend="cuda"`` — the data plane is the fused sm_100a kernel over symmetric memory
  (``akka_allreduce_b200.parallel.device_worker``); only membership and round bookkeeping go
  through the coordinator.

CLI (positional order of the reference, README.md:7)::

    python -m akka_allreduce_b200.worker [PORT=2553] [dataSize=10]
        [--master-host H] [--master-port 2551] [--checkpoint 50] [--assert-multiple 0] [--config FILE]

The built-in source is the constant ramp ``[0, 1, .., n-1]``; the built-in sink prints, every
``checkpoint`` iterations, the elapsed seconds and ``MBytes/sec = len*4*checkpoint/1e6/elapsed``
and optionally asserts ``output == input*k`` and ``count == k`` (AllreduceWorker.scala:324-343).
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

from . import host
from .api import AllReduceInput, AllReduceInputRequest, AllReduceOutput, DataSink, DataSource
from .config import DEFAULT_CONFIG_PATH, AllreduceConfig, load_config
from .utils.logging import get_logger


class AllreduceWorker:
    """CPU-backend worker: C++ engine + TCP node, user callbacks in Python."""

    def __init__(self, data_source: DataSource | None, data_sink: DataSink | None,
                 master_host: str = "127.0.0.1", master_port: int = 2551, port: int = 0,
                 heartbeat: float = 1.0, loglevel: str = "INFO", _node=None, _stats=None):
        self.log = get_logger("worker", loglevel)
        self._stats = _stats
        self._port = port
        if _node is not None:  # built by `builtin()`
            self._node = _node
            return
        if data_source is None or data_sink is None:
            raise ValueError("AllreduceWorker needs a data source and a data sink")
        opts = host.WorkerNodeOptions()
        opts.heartbeat_s = heartbeat

        def source(iteration: int):
            out = data_source(AllReduceInputRequest(iteration))
            return np.asarray(out.data, dtype=np.float32)

        def sink(data, count, iteration: int):
            data_sink(AllReduceOutput(data, count, iteration))

        self._node = host.WorkerNode(master_host, master_port, source, sink, opts)

    @classmethod
    def builtin(cls, data_size: int, checkpoint: int = 50, assert_multiple: int = 0,
                master_host: str = "127.0.0.1", master_port: int = 2551, port: int = 0,
                heartbeat: float = 1.0) -> "AllreduceWorker":
        """Worker with the native ramp source and throughput/assert sink (no Python per round)."""
        opts = host.WorkerNodeOptions()
        opts.heartbeat_s = heartbeat
        stats = host.ThroughputStats()
        node = host.WorkerNode.builtin(master_host, master_port, data_size, checkpoint, assert_multiple, stats, opts)
        return cls(None, None, master_host, master_port, port, heartbeat, _node=node, _stats=stats)

    def start(self) -> int:
        port = self._node.listen("127.0.0.1", self._port)
        self._node.begin()
        return port

    def wait(self, timeout: float = 3600.0) -> bool:
        """Blocks until the master says the run is over (or disappears)."""
        return self._node.wait_shutdown(timeout)

    def stop(self) -> None:
        self._node.stop()

    id = property(lambda self: self._node.worker_id)
    rounds_completed = property(lambda self: self._node.rounds_completed)
    stats = property(lambda self: self._stats)
    engine = property(lambda self: self._node.engine)


def create_worker(data_source: DataSource, data_sink: DataSink, config: AllreduceConfig,
                  master_host: str | None = None, master_port: int | None = None, **kwargs):
    """One constructor for every backend: ``config.device.backend`` picks the CPU worker (C++ protocol engine over
    TCP), the shared-memory worker (the device protocol over POSIX shared memory, one host) or the CUDA worker (fused
    kernel over NVLink); the callbacks and the contract are the same."""
    host_, port_ = master_host or config.cluster.host, master_port or config.cluster.port
    if config.device.backend == "cuda":
        from .parallel.device_worker import DeviceWorker

        return DeviceWorker(data_source, data_sink, config, host_, port_, **kwargs)
    if config.device.backend == "shm":
        from .parallel.shm_worker import ShmWorker

        return ShmWorker(data_source, data_sink, config, host_, port_, **kwargs)
    if config.device.backend != "cpu":
        raise ValueError(f"unknown backend {config.device.backend!r} (cpu | shm | cuda)")
    return AllreduceWorker(data_source, data_sink, host_, port_, heartbeat=config.cluster.heartbeat,
                           loglevel=config.cluster.loglevel, **kwargs)


def serve_worker_metrics(worker, port: int = 0, host_: str = "127.0.0.1"):
    """Prometheus endpoint for any of the three worker kinds (they share `id`, `rounds_completed` and, with the
    built-in sink, `stats`).  Returns the exporter (`.port`, `.close()`)."""
    from .utils.metrics import MetricsExporter

    def collect():
        m = {"aar_worker_id": worker.id, "aar_worker_rounds_completed": worker.rounds_completed}
        stats = getattr(worker, "stats", None)
        if stats is not None:
            m["aar_worker_checkpoints"] = stats.checkpoints
            m["aar_worker_sink_assertion_failures"] = stats.failures
            m["aar_worker_mbytes_per_s"] = stats.last_mbytes_per_s
        if hasattr(worker, "timeouts"):
            m["aar_worker_round_timeouts"] = worker.timeouts
        return m

    return MetricsExporter(collect, port, host_, {"role": "worker"})


def ramp_source(data_size: int) -> DataSource:
    floats = np.arange(data_size, dtype=np.float32)
    return lambda req: AllReduceInput(floats)


def start_up(port: int | str = 0, data_size: int = 10, checkpoint: int = 50, assert_multiple: int = 0,
             master_host: str = "127.0.0.1", master_port: int = 2551, backend: str = "cpu",
             config: AllreduceConfig | None = None):
    """Reference-style entry point (AllreduceWorker.scala:348-362)."""
    if backend == "cuda":
        from .parallel.device_worker import start_device_worker

        return start_device_worker(config or AllreduceConfig(), data_size, checkpoint, assert_multiple,
                                   master_host, master_port)
    if backend == "shm":
        from .parallel.shm_worker import start_shm_worker

        return start_shm_worker(config or AllreduceConfig(), data_size, checkpoint, assert_multiple, master_host,
                                master_port)
    w = AllreduceWorker.builtin(data_size, checkpoint, assert_multiple, master_host, master_port, int(port))
    w.start()
    return w


def main(argv: list[str] | None = None) -> int:
    ap = argparse.ArgumentParser(prog="akka_allreduce_b200.worker", description=__doc__,
                                 formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("port", nargs="?", type=int, default=2553)
    ap.add_argument("dataSize", nargs="?", type=int)
    ap.add_argument("--config")
    ap.add_argument("--master-host")
    ap.add_argument("--master-port", type=int)
    ap.add_argument("--checkpoint", type=int)
    ap.add_argument("--assert-multiple", type=int)
    ap.add_argument("--backend", choices=["cpu", "shm", "cuda"])
    ap.add_argument("--metrics-port", type=int, default=0, help="serve Prometheus metrics of this worker on this port")
    a = ap.parse_args(sys.argv[1:] if argv is None else argv)
    # like the reference, whose application.conf is picked up without being named: the checkout's (or the
    # packaged) application.conf is read unless --config / AAR_CONFIG names another file
    cfg = load_config(a.config or os.environ.get("AAR_CONFIG") or DEFAULT_CONFIG_PATH)
    data_size = a.dataSize if a.dataSize is not None else cfg.data.dataSize
    w = start_up(a.port, data_size,
                 a.checkpoint if a.checkpoint is not None else cfg.worker.checkpoint,
                 a.assert_multiple if a.assert_multiple is not None else cfg.worker.assertMultiple,
                 a.master_host or cfg.cluster.host, a.master_port or cfg.cluster.port,
                 a.backend or cfg.device.backend, cfg)
    exporter = serve_worker_metrics(w, a.metrics_port) if a.metrics_port > 0 else None
    try:
        w.wait()
    except KeyboardInterrupt:
        pass
    finally:
        if exporter is not None:
            exporter.close()
        w.stop()
    fails = w.stats.failures if getattr(w, "stats", None) is not None else 0
    return 1 if fails else 0


if __name__ == "__main__":
    sys.exit(main())
