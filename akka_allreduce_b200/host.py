"""Loader for the C++ host library (``_aar_host``): protocol engines, buffers, bus, TCP nodes.

Builds it in-tree on first use when the shared object is missing or stale.
"""
from __future__ import annotations

import importlib

from . import build as _build


def _load():
    _build.build_host()
    return importlib.import_module(f"akka_allreduce_b200.{_build.host_native_dir().name}._aar_host")


_lib = _load()

# protocol math
threshold_count = _lib.threshold_count
master_gate = _lib.master_gate
num_chunks = _lib.num_chunks
total_reduced_chunks_ref = _lib.total_reduced_chunks_ref
Partition = _lib.Partition
# messages
InitWorkers = _lib.InitWorkers
StartAllreduce = _lib.StartAllreduce
ScatterBlock = _lib.ScatterBlock
ReduceBlock = _lib.ReduceBlock
CompleteAllreduce = _lib.CompleteAllreduce
# buffers
ScatterBuffer = _lib.ScatterBuffer
ReduceBuffer = _lib.ReduceBuffer
# engines + bus
WorkerOptions = _lib.WorkerOptions
WorkerEngine = _lib.WorkerEngine
MasterConfig = _lib.MasterConfig
MasterEngine = _lib.MasterEngine
LocalBus = _lib.LocalBus
# TCP nodes
MasterNodeOptions = _lib.MasterNodeOptions
MasterNode = _lib.MasterNode
WorkerNodeOptions = _lib.WorkerNodeOptions
WorkerNode = _lib.WorkerNode
ThroughputStats = _lib.ThroughputStats
# device protocol restated for CPU threads in shared memory (csrc/host/shm_cluster.h)
ShmConfig = _lib.ShmConfig
ShmCluster = _lib.ShmCluster
