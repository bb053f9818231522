"""Model family used by the DDP benchmark: shapes, parameter count, a CPU forward/backward."""
import torch

from akka_allreduce_b200.models.gpt2 import GPT2, GPT2Config


def test_gpt2_small_parameter_count():
    cfg = GPT2Config.small()
    with torch.device("meta"):
        m = GPT2(cfg)
    assert m.num_parameters() == 124_439_808  # GPT-2 small with tied output embedding


def test_gpt2_tiny_forward_backward_cpu():
    torch.manual_seed(0)
    cfg = GPT2Config.tiny()
    m = GPT2(cfg)
    x = torch.randint(0, cfg.vocab_size, (2, 17))
    loss = m(x[:, :-1], x[:, 1:])
    loss.backward()
    assert torch.isfinite(loss) and abs(float(loss) - torch.log(torch.tensor(float(cfg.vocab_size)))) < 1.0
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in m.parameters())
    assert m(x[:, :-1]).shape == (2, 16, cfg.vocab_size)


def test_numa_placement_helpers(tmp_path, monkeypatch):
    """utils/affinity.py: cpulist parsing, the sysfs lookups against a fake tree, and the no-op cases."""
    import os

    from akka_allreduce_b200.utils import affinity as A

    assert A.parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert A.parse_cpulist("") == set()
    dev = tmp_path / "bus/pci/devices/0000:19:00.0"
    dev.mkdir(parents=True)
    (dev / "numa_node").write_text("1\n")
    node = tmp_path / "devices/system/node/node1"
    node.mkdir(parents=True)
    allowed = sorted(os.sched_getaffinity(0))
    (node / "cpulist").write_text(",".join(str(c) for c in allowed[: max(1, len(allowed) // 2)]))
    assert A.numa_node_of_pci("0000:19:00.0", str(tmp_path)) == 1
    assert A.numa_node_of_pci("0000:1A:00.0", str(tmp_path)) is None          # unknown device
    (dev / "numa_node").write_text("-1\n")
    assert A.numa_node_of_pci("0000:19:00.0", str(tmp_path)) is None          # platform does not say
    (dev / "numa_node").write_text("1\n")
    assert A.cpus_of_node(1, str(tmp_path)) == set(allowed[: max(1, len(allowed) // 2)])
    assert A.cpus_of_node(7, str(tmp_path)) == set()
    # bind_to_gpu end to end with a faked PCI address, restoring the mask afterwards
    monkeypatch.setattr(A, "pci_address", lambda i: "0000:19:00.0")
    before = os.sched_getaffinity(0)
    try:
        got = A.bind_to_gpu(0, str(tmp_path))
        if len(allowed) // 2 >= A.MIN_LOCAL_CPUS:
            assert got == {"bound": True, "numa_node": 1, "cpus": len(allowed) // 2, "source": "sysfs"}
            assert os.sched_getaffinity(0) == set(allowed[: len(allowed) // 2])
        else:
            assert got["bound"] is False and "reason" in got
    finally:
        os.sched_setaffinity(0, before)
    monkeypatch.setenv("AAR_NO_AFFINITY", "1")
    assert A.bind_to_gpu(0, str(tmp_path)) is None
    monkeypatch.delenv("AAR_NO_AFFINITY")
    monkeypatch.setattr(A, "pci_address", lambda i: None)
    monkeypatch.setattr(A, "nvml_cpus_of_gpu", lambda i: set())
    got = A.bind_to_gpu(0, str(tmp_path))   # nothing names the GPU's cores: nothing bound, and the record says why
    assert got["bound"] is False and "reason" in got
    # a container that hides sysfs: NVML's CPU affinity for the device is used instead
    half = set(allowed[: len(allowed) // 2])
    monkeypatch.setattr(A, "nvml_cpus_of_gpu", lambda i: half)
    try:
        got = A.bind_to_gpu(0, str(tmp_path))
        if len(half) >= A.MIN_LOCAL_CPUS:
            assert got == {"bound": True, "numa_node": None, "cpus": len(half), "source": "nvml"}
    finally:
        os.sched_setaffinity(0, before)


def test_bench_roofline_bookkeeping(tmp_path):
    """bench.py's roofline fractions: HBM-bound on one GPU (against the measured copy bandwidth), NVLink-bound beyond
    (B, B x 2(P-1)/(P+1) for the in-switch variant, B x 2/P for one-shot)."""
    import json
    import os
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import bench

    peaks = tmp_path / "peaks.json"
    peaks.write_text(json.dumps({"hbm_gbs": 6000.0}))
    n = 64 << 20
    r = bench.roofline(1, "twoshot", n, 0.0865, True, str(peaks))
    assert r["bound"] == "HBM" and r["bytes_per_step"] == 3 * n and abs(r["fraction"] - 3 * n / 86.5e-6 / 1e9 / 6000) < 1e-3
    assert bench.roofline(1, "twoshot", n, 0.123, False, str(peaks))["bytes_per_step"] == 7 * n
    assert bench.roofline(1, "twoshot", n, 0.1, True, str(tmp_path / "missing.json"))["fraction"] is None
    r = bench.roofline(8, "nvls", n, 0.2047, True, str(peaks))
    assert r["roofline_busbw_GBps"] == 1400.0 and abs(r["busbw_GBps"] - 573.7) < 1.0 and abs(r["fraction"] - 0.41) < 0.01
    assert bench.roofline(2, "twoshot", n, 0.135, True, str(peaks))["roofline_busbw_GBps"] == 900.0
    assert bench.roofline(8, "oneshot", 1 << 20, 0.02, True, str(peaks))["roofline_busbw_GBps"] == 225.0


def test_bench_reference_arm_prints_one_json_line_and_probes_the_box():
    """Driver contract: `bench.py --impl reference` exits 0 with one JSON line; when the reference cannot run it says
    why, after actually looking for a JVM toolchain and dependency caches on this machine."""
    import json
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    p = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--gpus", "8", "--steps", "5",
                        "--warmup", "3"], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and "unavailable" in d
    for probe in ("java=", "sbt=", ".ivy2=", "baseline/_ref installed="):
        assert probe in d["unavailable"]
