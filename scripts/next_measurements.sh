#!/usr/bin/env bash
# The measurements DESIGN.md §7 asks for first, in the order of what they are expected to change (8 GPUs; about
# 4 minutes of box time).  Every line prints JSON; nothing here changes defaults.
set -u
cd "$(dirname "$0")/.."
N="${N:-8}"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
mkdir -p gpurun_out
echo "--- e2e with and without NUMA placement (utils/affinity.py)"
timeout 300 $TR --master-port 29701 bench.py --gpus $N --steps 50 --warmup 5 --no-extras 2>&1 | grep "^{" | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print("placement", d["config"].get("cpu_placement"), "e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], "kernel", d["ms_per_step"])'
AAR_NO_AFFINITY=1 timeout 300 $TR --master-port 29702 bench.py --gpus $N --steps 50 --warmup 5 --no-extras 2>&1 | grep "^{" | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print("no placement: e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], "kernel", d["ms_per_step"])'
echo "--- 0.5-2 MB: push one-shot vs pull one-shot (in-switch reduction) vs two-shot"
timeout 300 $TR --master-port 29711 benchmarks/sweep.py --sizes 512K,1M,2M --algo auto 2>&1 | grep "^{" | cut -c1-330
AAR_ONESHOT_BYTES=2097152 timeout 300 $TR --master-port 29712 benchmarks/sweep.py --sizes 512K,1M,2M --algo auto --no-nccl 2>&1 | grep "^{" | cut -c1-260
AAR_ONESHOT_PULL=1 AAR_ONESHOT_BYTES=2097152 timeout 300 $TR --master-port 29713 benchmarks/sweep.py --sizes 512K,1M,2M --algo auto --no-nccl 2>&1 | grep "^{" | cut -c1-260
echo "--- launch prologue (gate + destination mask) at 1 MB: gate_us of the trace"
timeout 300 $TR --master-port 29721 benchmarks/sweep.py --sizes 1M --algo twoshot --no-nccl --trace 2>&1 | grep -E "rank 0. trace" | cut -c1-400
AAR_KEEP_GATE=1 timeout 300 $TR --master-port 29722 benchmarks/sweep.py --sizes 1M --algo twoshot --no-nccl --trace 2>&1 | grep -E "rank 0. trace" | cut -c1-400
