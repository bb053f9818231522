#!/bin/bash
# One 8-GPU measurement pass (gpurun --gpus 8): every output lands in gpurun_out/.
NG=${NG:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1"
O=gpurun_out
mkdir -p $O
echo "== bench.py --gpus $NG"
timeout 400 $TR --master-port 29701 bench.py --gpus $NG --steps 50 --warmup 5 > $O/bench_n${NG}.json 2> $O/bench_n${NG}.err; tail -2 $O/bench_n${NG}.err; head -c 6000 $O/bench_n${NG}.json; echo
echo "== straggler (config 3)"
timeout 300 $TR --master-port 29711 benchmarks/straggler.py --rounds 400 --out $O/straggler_n${NG}.json 2>&1 | grep -E "^\{" | head -c 3000; echo
echo "== soak"
timeout 300 $TR --master-port 29721 benchmarks/soak.py --rounds 2000 --zero-copy --out $O/soak_n${NG}.jsonl 2>&1 | grep -E "^\{"
timeout 300 $TR --master-port 29722 benchmarks/soak.py --rounds 1000 --out $O/soak_n${NG}.jsonl 2>&1 | grep -E "^\{"
echo "== small messages: one-shot variants"
for pull in 0 1; do AAR_ONESHOT_PULL=$pull timeout 200 $TR --master-port 29731 benchmarks/sweep.py --sizes 256K,512K,1M,2M,4M --chunks 1M --iters 30 --algo oneshot --out $O/sweep${NG}_oneshot_pull$pull.jsonl 2>&1 | grep -E "^\{" | python -c "import sys,json; [print(' pull$pull', r['bytes']>>10, 'KB', r['ms'], 'nccl', r.get('nccl_ms'), 'x', r.get('speedup_vs_nccl')) for r in map(json.loads, sys.stdin)]"; done
echo "== small messages: stream NVLS, ctas"
timeout 200 $TR --master-port 29732 benchmarks/sweep.py --sizes 1M,4M --chunks 1M --iters 30 --algo nvls --ctas 8,16,32,64 --no-nccl 2>&1 | grep -E "^\{" | python -c "import sys,json; [print(' nvls', r['bytes']>>10, 'KB ctas', r['ctas'], r['ms']) for r in map(json.loads, sys.stdin)]"
echo "== hybrid push chunks (ordinary tensors)"
for pc in 0 1 2 3; do AAR_PUSH_CHUNKS=$pc timeout 200 $TR --master-port 29741 benchmarks/sweep.py --sizes 64M,256M --chunks 1M --iters 20 --algo nvls --no-nccl 2>&1 | grep -E "^\{" | python -c "import sys,json; [print(' push_chunks=$pc', r['bytes']>>20, 'MB', r['ms']) for r in map(json.loads, sys.stdin)]"; done
echo "== chunk-size sweep"
timeout 400 $TR --master-port 29751 benchmarks/sweep.py --sizes 16M,64M,256M --chunks 64K,256K,1M,4M --iters 15 --out $O/sweep${NG}_chunks.jsonl 2>&1 | grep -E "^\{" | python -c "import sys,json; [print(' ', r['bytes']>>20, 'MB chunk', r['chunk_bytes']>>10, 'KB', r['ms'], r['algo'], 'x', r.get('speedup_vs_nccl')) for r in map(json.loads, sys.stdin)]"
echo "== size sweep, ordinary tensors and registered buffers"
timeout 300 $TR --master-port 29761 benchmarks/sweep.py --sizes 1K,8K,64K,256K,1M,4M,16M,64M,256M,1G --chunks 1M --iters 20 --out $O/sweep${NG}_user.jsonl 2>&1 | grep -E "^\{" | python -c "import sys,json; [print(' user', r['bytes'], r['ms'], r['algo'], 'x', r.get('speedup_vs_nccl')) for r in map(json.loads, sys.stdin)]"
timeout 300 $TR --master-port 29762 benchmarks/sweep.py --sizes 1M,4M,16M,64M,256M,1G --chunks 1M --iters 20 --zero-copy --out $O/sweep${NG}_zero_copy.jsonl 2>&1 | grep -E "^\{" | python -c "import sys,json; [print(' zc', r['bytes'], r['ms'], r['algo'], 'x', r.get('speedup_vs_nccl')) for r in map(json.loads, sys.stdin)]"
