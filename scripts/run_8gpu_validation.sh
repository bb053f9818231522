#!/usr/bin/env bash
# Everything that needs the full 8-GPU box, in one go (GPU minutes are charged x8): correctness, headline bench,
# sweeps vs NCCL, stragglers, DDP.  Outputs land in gpurun_out/; scripts/make_profiles.py turns them into profiles/.
set -u
cd "$(dirname "$0")/.."
N="${N:-8}"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -3
grep -E "FAIL" gpurun_out/multi_gpu_test.log | head -5
timeout 300 $TR --master-port 29601 bench.py --gpus $N --steps 100 --warmup 5 2>&1 | grep "^{" > gpurun_out/bench_${N}gpu.json; cut -c1-900 gpurun_out/bench_${N}gpu.json
timeout 400 $TR --master-port 29611 benchmarks/sweep.py --sizes 1K,8K,64K,256K,1M,4M,16M,64M,256M,1G --algo auto --zero-copy --out gpurun_out/sweep_${N}gpu_zero_copy.jsonl 2>&1 | grep "^{" | cut -c1-330
timeout 400 $TR --master-port 29621 benchmarks/sweep.py --sizes 1M,64M,1G --algo auto --out gpurun_out/sweep_${N}gpu_user_tensors.jsonl 2>&1 | grep "^{" | cut -c1-330
echo "--- experiments"
AAR_NVLS_UNROLL=16 timeout 300 $TR --master-port 29631 benchmarks/sweep.py --sizes 256M,1G --algo auto --zero-copy --no-nccl 2>&1 | grep "^{" | cut -c1-260
timeout 300 $TR --master-port 29641 benchmarks/sweep.py --sizes 1G --chunks 4M --algo auto --zero-copy --no-nccl --trace 2>&1 | grep -E "^\{|rank 0. trace" | cut -c1-500
timeout 300 $TR --master-port 29651 benchmarks/sweep.py --sizes 1G --algo auto --zero-copy --no-nccl --no-counts 2>&1 | grep "^{" | cut -c1-260
echo "--- stragglers"
timeout 300 $TR --master-port 29661 benchmarks/straggler.py --rounds 200 --out gpurun_out/straggler_${N}gpu.json 2>&1 | grep "^{" | cut -c1-700
echo "--- ddp"
timeout 400 $TR --master-port 29671 benchmarks/ddp_gpt2.py --batch 8 --steps 8 --out gpurun_out/ddp_gpt2_${N}gpu.json 2>&1 | grep -E "^\{|rror" | cut -c1-900
