#!/usr/bin/env bash
# Manual 4-worker cluster run, master side (counterpart of the reference's scripts/testAllreduceMaster.sc:
# thresholds 1/1/1, dataSize 778, maxChunkSize 3, maxRound 1000, totalSize 4, maxLag 3).
# Start this first, then scripts/test_allreduce_worker.sh four times.  BACKEND=cuda uses one GPU per worker.
set -euo pipefail
cd "$(dirname "$0")/.."
PORT="${PORT:-2551}"
exec python -m akka_allreduce_b200.master "$PORT" 4 778 3 \
  --thAllreduce 1.0 --thReduce 1.0 --thComplete 1.0 --maxLag 3 --maxRound "${MAX_ROUND:-1000}"
