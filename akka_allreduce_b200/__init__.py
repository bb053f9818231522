"""akka_allreduce_b200 — a Blackwell-native threshold/lag allreduce framework.

Capabilities of GuixingLin/akka-allreduce (master-gated rounds, source/sink callbacks, chunked
scatter -> reduce -> broadcast with scatter/reduce/online-nodes thresholds, per-element contributor
counts, ``maxLag`` rounds in flight), rebuilt for 8xB200: a C++ coordinator and host protocol
engine, and a hand-written sm_100a kernel that fuses the pipeline over NVLink peer memory.
"""
from .api import AllReduceInput, AllReduceInputRequest, AllReduceOutput, DataSink, DataSource
from .config import (AllreduceConfig, ClusterConfig, DataConfig, DeviceConfig, ThresholdConfig,
                     WorkerConfig, load_config)

__version__ = "0.1.0"

__all__ = [
    "AllReduceInput", "AllReduceInputRequest", "AllReduceOutput", "DataSink", "DataSource",
    "AllreduceConfig", "ClusterConfig", "DataConfig", "DeviceConfig", "ThresholdConfig", "WorkerConfig",
    "load_config", "AllreduceMaster", "AllreduceWorker", "create_worker",
]


def __getattr__(name):  # lazy: importing the package must not trigger a native build
    if name == "AllreduceMaster":
        from .master import AllreduceMaster

        return AllreduceMaster
    if name == "AllreduceWorker":
        from .worker import AllreduceWorker

        return AllreduceWorker
    if name == "create_worker":
        from .worker import create_worker

        return create_worker
    raise AttributeError(name)
