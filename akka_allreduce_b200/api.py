"""User-facing source/sink contract.

Same three payload types as the reference (DataWrapper.scala:3-7) and the same two callback
shapes (AllreduceWorker.scala:305-306):

    DataSource = AllReduceInputRequest -> AllReduceInput
    DataSink   = AllReduceOutput       -> None

``data`` is the raw **sum** over the contributors, ``count`` the per-element number of
contributors; nothing is divided unless ``device.normalize`` is set.  On the CPU backend the
payloads are numpy arrays; on the CUDA backend they are torch tensors on the worker's GPU and
``AllReduceOutput.data`` / ``count`` are views of the round's sink slot, valid until that slot
is recycled ``maxLag + 1`` rounds later.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Callable


@dataclass
class AllReduceInputRequest:
    iteration: int


@dataclass
class AllReduceInput:
    data: Any  # float array of length dataSize


@dataclass
class AllReduceOutput:
    data: Any       # summed values, length dataSize
    count: Any      # int32 contributors per element, length dataSize
    iteration: int


DataSource = Callable[[AllReduceInputRequest], AllReduceInput]
DataSink = Callable[[AllReduceOutput], None]
