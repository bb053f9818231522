"""AllreduceMaster: the rank-0 coordinator and its launcher.

Counterpart of the reference's master actor + ``object AllreduceMaster`` launcher
(AllreduceMaster.scala:12-146): listens on the configured port, waits for ``totalSize`` workers,
assigns ids in arrival order, broadcasts InitWorkers once, and gates every round on
``thAllreduce``.  The gate itself runs in the C++ ``MasterNode``.

CLI (positional order of the reference, README.md:3)::

    python -m akka_allreduce_b200.master [PORT=2551] [numOfWorkers=2] [dataSize=5*numOfWorkers] [maxChunkSize=2]
        [--config FILE] [--thAllreduce F] [--thReduce F] [--thComplete F] [--maxLag N] [--maxRound N]
"""
from __future__ import annotations

import argparse
import os
import sys

from . import host
from .config import DEFAULT_CONFIG_PATH, AllreduceConfig, DataConfig, ThresholdConfig, WorkerConfig, load_config
from .utils.logging import get_logger


class AllreduceMaster:
    def __init__(self, config: AllreduceConfig, verbose: bool = False):
        self.config = config
        self.log = get_logger("master", config.cluster.loglevel)
        opts = host.MasterNodeOptions()
        opts.down_after_s = config.cluster.downAfter
        opts.verbose = verbose
        mc = config.master_config()
        # Workers of the device protocol (GPU, or its shared-memory form) must be told when a peer is gone: they would
        # push into and wait on a dead process's memory.  The TCP backend keeps the reference's behaviour (stale peers
        # become dead letters) unless asked.
        mc.announce_removals = config.cluster.announceRemovals or config.device.backend in ("cuda", "shm")
        self._node = host.MasterNode(mc, opts)
        self.port = None
        self.metrics = None

    def start(self) -> int:
        c = self.config
        self.port = self._node.listen(c.cluster.host, c.cluster.port)
        self.log.info("-------\n Port = %s \n Number of Workers = %s \n Message Size = %s \n Max Chunk Size = %s",
                      self.port, c.worker.totalSize, c.data.dataSize, c.data.maxChunkSize)
        self._node.start()
        if c.cluster.metricsPort > 0:
            self.serve_metrics(c.cluster.metricsPort)
        return self.port

    def serve_metrics(self, port: int = 0) -> int:
        """Prometheus endpoint with the coordinator's view of the job; returns the port (0 = any free one)."""
        from .utils.metrics import MetricsExporter

        def collect():
            return {"aar_master_round": self.round, "aar_master_alive_workers": self.alive,
                    "aar_master_registered_workers": self.registered,
                    "aar_master_total_workers": self.config.worker.totalSize,
                    "aar_master_max_round": self.config.data.maxRound,
                    "aar_master_finished": 1.0 if self._node.wait_finished(0.0) else 0.0}

        self.metrics = MetricsExporter(collect, port, self.config.cluster.host, {"role": "master"})
        return self.metrics.port

    def wait(self, timeout: float = 3600.0) -> bool:
        """Blocks until round ``maxRound`` has passed the gate (or the timeout expires)."""
        return self._node.wait_finished(timeout)

    def stop(self) -> None:
        if self.metrics is not None:
            self.metrics.close()
            self.metrics = None
        self._node.stop()

    round = property(lambda self: self._node.round)
    registered = property(lambda self: self._node.registered)
    alive = property(lambda self: self._node.alive)

    # -- reference-style entry points (AllreduceMaster.scala:138-144) ----------------------------
    @staticmethod
    def start_up(port: int | str | None = None, thresholds: ThresholdConfig | None = None,
                 data_config: DataConfig | None = None, worker_config: WorkerConfig | None = None,
                 config: AllreduceConfig | None = None, verbose: bool = True) -> "AllreduceMaster":
        cfg = config or AllreduceConfig()
        if thresholds:
            cfg.thresholds = thresholds
        if data_config:
            cfg.data = data_config
        if worker_config:
            cfg.worker = worker_config
        if port is not None:
            cfg.cluster.port = int(port)
        m = AllreduceMaster(cfg, verbose=verbose)
        m.start()
        return m


def build_config(argv: list[str]) -> AllreduceConfig:
    ap = argparse.ArgumentParser(prog="akka_allreduce_b200.master", description=__doc__,
                                 formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("port", nargs="?", type=int)
    ap.add_argument("numOfWorkers", nargs="?", type=int)
    ap.add_argument("dataSize", nargs="?", type=int)
    ap.add_argument("maxChunkSize", nargs="?", type=int)
    ap.add_argument("--config")
    for k in ("thAllreduce", "thReduce", "thComplete"):
        ap.add_argument("--" + k, type=float)
    for k in ("maxLag", "maxRound"):
        ap.add_argument("--" + k, type=int)
    ap.add_argument("--host")
    a = ap.parse_args(argv)
    # like the reference, whose application.conf is picked up without being named: the checkout's (or the
    # packaged) application.conf is read unless --config / AAR_CONFIG names another file
    cfg = load_config(a.config or os.environ.get("AAR_CONFIG") or DEFAULT_CONFIG_PATH)
    if a.numOfWorkers is not None:
        cfg.worker.totalSize = a.numOfWorkers
        if a.dataSize is None and a.config is None:
            cfg.data.dataSize = 5 * a.numOfWorkers  # AllreduceMaster.scala:103
    overrides = {"port": a.port, "dataSize": a.dataSize, "maxChunkSize": a.maxChunkSize, "host": a.host,
                 "thAllreduce": a.thAllreduce, "thReduce": a.thReduce, "thComplete": a.thComplete,
                 "maxLag": a.maxLag, "maxRound": a.maxRound}
    return _with(cfg, overrides)


def _with(cfg: AllreduceConfig, overrides: dict) -> AllreduceConfig:
    from .config import _apply

    _apply(cfg, {k: v for k, v in overrides.items() if v is not None})
    return cfg


def main(argv: list[str] | None = None) -> int:
    cfg = build_config(sys.argv[1:] if argv is None else argv)
    master = AllreduceMaster(cfg, verbose=True)
    master.start()
    try:
        master.wait()
    except KeyboardInterrupt:
        pass
    finally:
        master.stop()
    return 0


if __name__ == "__main__":
    sys.exit(main())
