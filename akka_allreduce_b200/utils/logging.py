"""Process-wide logging (the reference uses ActorLogging + println; AllreduceWorker.scala:8,79-86)."""
from __future__ import annotations

import logging
import sys

_FMT = "[%(levelname)s] [%(asctime)s] [%(name)s] %(message)s"


def get_logger(name: str, level: str = "INFO") -> logging.Logger:
    log = logging.getLogger("aar." + name)
    if not log.handlers:
        h = logging.StreamHandler(sys.stderr)  # stdout is for results (bench JSON line, throughput prints)
        h.setFormatter(logging.Formatter(_FMT))
        log.addHandler(h)
        log.propagate = False
    log.setLevel(getattr(logging, str(level).upper(), logging.INFO))
    return log
