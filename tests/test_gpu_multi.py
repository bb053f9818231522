"""Real multi-GPU runs (NVLink peers, one process per GPU) of the fused kernel and the worker API."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port() -> int:
    """A rendezvous port nobody is listening on (a fixed one can still be in TIME_WAIT from the previous test)."""
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_all_visible_gpus():
    n = min(torch.cuda.device_count(), 8)
    port = _free_port()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "multi_gpu_worker.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "multi_gpu_test.log"), "w") as f:
        f.write(p.stdout + "\n--- stderr ---\n" + p.stderr)
    lines = [l for l in (p.stdout + p.stderr).splitlines() if "FAIL" in l or "Error" in l or "error" in l]
    assert p.returncode == 0, "\n".join(lines[-40:])


def test_worker_death_is_absorbed_without_timeouts():
    """tests/multi_gpu_kill_worker.py: one GPU worker dies mid-job, the coordinator's removal becomes the kernels'
    dead-peer mask, the survivors finish at full speed with count P-1."""
    n = min(torch.cuda.device_count(), 8)
    port = _free_port()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "multi_gpu_kill_worker.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    with open(os.path.join(ROOT, "gpurun_out", "multi_gpu_kill.log"), "w") as f:
        f.write(p.stdout + "\n--- stderr ---\n" + p.stderr)
    lines = [l for l in (p.stdout + p.stderr).splitlines() if "FAIL" in l or "Error" in l or "] ok" in l]
    assert p.returncode == 0, "\n".join(lines[-40:])
