#!/bin/bash
# Final measurement pass for one GPU count: NG=8 bash benchmarks/run_final.sh   (outputs under gpurun_out/)
NG=${NG:-8}
O=gpurun_out
mkdir -p $O
if [ "$NG" = "1" ]; then TR="python"; else TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29801"; fi
echo "== bench.py --gpus $NG"
timeout 400 $TR bench.py --gpus $NG --steps ${STEPS:-40} --warmup 5 2> $O/bench_final_n${NG}.err | grep -E "^\{" > $O/bench_final_n${NG}.json; tail -2 $O/bench_final_n${NG}.err | cut -c1-200; python - <<PY
import json
d=json.loads(open("$O/bench_final_n${NG}.json").read())
print({k: d.get(k) for k in ("ms_per_step","value","vs_nccl","correct","algo","kernel","timeouts")})
print("registered", {k: (d.get("registered_buffers") or {}).get(k) for k in ("ms","vs_nccl_ordinary","vs_best_registered_baseline","correct")})
print("nccl", d.get("nccl"), "nccl_registered", d.get("nccl_registered"), "symm_mem", d.get("symm_mem"))
print("e2e", {k: (d.get("e2e") or {}).get(k) for k in ("ms_per_step","value","vs_nccl_same_path","min_count")})
PY
[ "$NG" = "1" ] && exit 0
TR2="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1"
if [ "$NG" -ge 4 ]; then
echo "== straggler (config 3)"
timeout 300 $TR2 --master-port 29811 benchmarks/straggler.py --rounds ${ROUNDS:-400} --out $O/straggler_final_n${NG}.json 2>&1 | grep -E "^\{" | head -c 2500; echo
fi
echo "== size sweep, ordinary tensors"
timeout 300 $TR2 --master-port 29821 benchmarks/sweep.py --sizes ${SIZES:-1K,64K,256K,1M,2M,4M,16M,64M,256M,1G} --chunks 1M --iters 20 --out $O/sweep_final_n${NG}_user.jsonl 2>&1 | grep -E "^\{" | python -c "import sys,json; [print(' user', r['bytes'], r['ms'], r['algo'], r['kernel'], 'x', r.get('speedup_vs_nccl')) for r in map(json.loads, sys.stdin)]"
if [ "$NG" = "8" ]; then
echo "== registered buffers, large"
timeout 300 $TR2 --master-port 29831 benchmarks/sweep.py --sizes ${ZC_SIZES:-64M,256M,1G} --chunks 1M --iters 20 --zero-copy --out $O/sweep_final_n${NG}_zero_copy.jsonl 2>&1 | grep -E "^\{" | python -c "import sys,json; [print(' zc', r['bytes'], r['ms'], r['algo'], 'x', r.get('speedup_vs_nccl')) for r in map(json.loads, sys.stdin)]"
fi
