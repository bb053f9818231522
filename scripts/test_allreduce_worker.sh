#!/usr/bin/env bash
# Manual cluster run, worker side (counterpart of scripts/testAllreduceWorker.sc: random port, dataSize 778,
# checkpoint 10, assertMultiple 4 — the sink asserts output == 4*input and count == 4 every 10 rounds).
set -euo pipefail
cd "$(dirname "$0")/.."
exec python -m akka_allreduce_b200.worker 0 778 --master-port "${PORT:-2551}" --checkpoint 10 --assert-multiple 4 \
  --backend "${BACKEND:-cpu}"
