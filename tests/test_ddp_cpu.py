"""Data parallelism on CPU: 2 processes, gloo for DDP's bookkeeping, gradients averaged by the device
protocol in shared memory through `ShmAllreduceCommHook`; parameters after 3 SGD steps match stock DDP."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_ddp_with_the_shared_memory_hook_matches_stock_ddp():
    port = 29500 + os.getpid() % 400
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "ddp_cpu_worker.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=dict(os.environ, PYTHONPATH=ROOT))
    out = p.stdout + p.stderr
    assert p.returncode == 0 and out.count("ok rank") == 2, out[-3000:]
