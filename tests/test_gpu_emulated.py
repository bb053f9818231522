"""The fused sm_100a kernel, all ranks emulated on one GPU (one cooperative launch per round).

Numerics are checked against a plain PyTorch fp32 reference of the same op (sum over the ranks
named by ``count``), and against the reference's semantics for partial rounds: output equals the
sum over exactly ``count`` contributors, zeros where ``count == 0``.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _cluster(*a, **kw):
    from akka_allreduce_b200.ops.fused_allreduce import EmulatedCluster

    return EmulatedCluster(*a, **kw)


def _srcs(world, n, dtype, r, dev="cuda:0"):
    g = torch.Generator(device="cpu").manual_seed(1000 * r + world)
    return [torch.randn(n, generator=g, dtype=torch.float32).to(dtype).to(dev) for _ in range(world)]


def _ref_sum(srcs, ranks=None):
    ranks = range(len(srcs)) if ranks is None else ranks
    acc = torch.zeros_like(srcs[0], dtype=torch.float32)
    for i in ranks:  # fixed peer order, fp32 accumulate: same order as the kernel
        acc += srcs[i].float()
    return acc


@pytest.mark.parametrize("world,n,mc,max_lag", [
    (2, 10, 2, 1),            # BASELINE config 1 shape
    (4, 778, 3, 3),           # scripts/testAllreduceMaster.sc shape (scalar path, uneven tail)
    (3, 9, 2, 5),
    (4, 9, 2, 1),             # empty trailing block
    (8, 1 << 20, 1 << 14, 1),  # vector path, 8 chunks per block
    (8, 1 << 18, 1 << 18, 2),  # one chunk per block
    (5, 100003, 4096, 2),     # odd everything
])
def test_thresholds_one_matches_fp32_sum(world, n, mc, max_lag):
    cl = _cluster(world, n, mc, max_lag=max_lag, timeout_ms=5000)
    for r in range(2 * (max_lag + 1) + 1):  # wraps the lag ring twice
        srcs = _srcs(world, n, torch.float32, r)
        outs, counts = cl.run_round(srcs)
        want = _ref_sum(srcs)
        for i in range(world):
            assert torch.equal(outs[i], want), f"round {r} rank {i}: max err {(outs[i] - want).abs().max()}"
            assert bool((counts[i] == world).all())
    assert cl.timeouts_seen() == [0] * world
    assert cl.done_tags() == [cl.round] * world


def test_assert_multiple_ramp():
    """AllreduceWorker.scala:324-343: ramp source, output == input*k and count == k."""
    world, n = 4, 778
    cl = _cluster(world, n, 3, max_lag=3)
    ramp = torch.arange(n, dtype=torch.float32, device="cuda:0")
    for _ in range(10):
        outs, counts = cl.run_round([ramp] * world)
    for i in range(world):
        assert torch.equal(outs[i], ramp * world) and bool((counts[i] == world).all())


@pytest.mark.parametrize("wire", [torch.bfloat16, torch.float32])
def test_bf16_source(wire):
    world, n = 8, 1 << 19
    cl = _cluster(world, n, 1 << 15, dtype=torch.bfloat16, wire_dtype=wire, align_blocks=True)
    srcs = _srcs(world, n, torch.bfloat16, 3)
    outs, counts = cl.run_round(srcs)
    want = _ref_sum(srcs)
    for i in range(world):
        # bf16 output of an fp32 accumulation: one rounding of the result
        torch.testing.assert_close(outs[i].float(), want, rtol=2 ** -7, atol=1e-2)
        assert bool((counts[i] == world).all())
    # every rank holds bit-identical output (one owner reduces each chunk)
    assert all(torch.equal(outs[0], outs[i]) for i in range(1, world))


def test_scale_and_normalize():
    world, n = 4, 4096
    cl = _cluster(world, n, 512, scale=0.5, normalize=True)
    srcs = _srcs(world, n, torch.float32, 7)
    outs, _ = cl.run_round(srcs)
    want = sum((s * 0.5 for s in srcs[1:]), srcs[0] * 0.5) / world
    torch.testing.assert_close(outs[2], want, rtol=1e-6, atol=1e-6)


def test_scatter_threshold_with_muted_rank():
    """thReduce 0.75 of 4: a rank that never scatters is absorbed; count reports 3 contributors."""
    from akka_allreduce_b200.ops.fused_allreduce import Injection

    world, n = 4, 1 << 16
    cl = _cluster(world, n, 1 << 12, th_reduce=0.75, th_complete=1.0, max_lag=2)
    for r in range(4):
        srcs = _srcs(world, n, torch.float32, r)
        outs, counts = cl.run_round(srcs, Injection(mute=1 << 3))
        want = _ref_sum(srcs, ranks=[0, 1, 2])
        for i in range(world):
            assert torch.equal(outs[i], want) and bool((counts[i] == 3).all())
    assert cl.timeouts_seen() == [0] * world


def test_dropped_scatter_to_one_destination():
    """Rank 1 drops its ScatterBlocks to rank 2 only: block 2 is closed by the bounded wait with 3
    contributors, the rest has 4.  (Thresholds stay at 1 so the contributor sets are deterministic: with
    thReduce < 1 a chunk fires on the first poll that sees enough arrivals.)"""
    from akka_allreduce_b200.ops.fused_allreduce import Injection

    world, n = 4, 4000
    cl = _cluster(world, n, 256, timeout_ms=20)
    srcs = _srcs(world, n, torch.float32, 11)
    outs, counts = cl.run_round(srcs, Injection(skip={1: 1 << 2}))
    full, part = _ref_sum(srcs), _ref_sum(srcs, ranks=[0, 2, 3])
    lo, hi = cl.geo.start(2), cl.geo.end(2)
    for i in range(world):
        assert torch.equal(outs[i][:lo], full[:lo]) and torch.equal(outs[i][hi:], full[hi:])
        assert torch.equal(outs[i][lo:hi], part[lo:hi])
        assert bool((counts[i][lo:hi] == 3).all()) and bool((counts[i][:lo] == 4).all())


def test_completion_threshold_with_straggling_owner():
    """thComplete 0.75: a rank that starts 30 ms late misses the round on the other ranks — its block
    comes out as zeros with count 0 there (ReducedDataBuffer.scala:26-53), the others' blocks with
    count 3 (its scatter was late too)."""
    from akka_allreduce_b200.ops.fused_allreduce import Injection

    world, n = 4, 1 << 14
    cl = _cluster(world, n, 1 << 10, th_reduce=0.75, th_complete=0.75, timeout_ms=5000)
    srcs = _srcs(world, n, torch.float32, 5)
    outs, counts = cl.run_round(srcs, Injection(delay_ns={3: 30_000_000}))
    part = _ref_sum(srcs, ranks=[0, 1, 2])
    lo = cl.geo.start(3)
    for i in range(3):
        assert torch.equal(outs[i][:lo], part[:lo]) and bool((counts[i][:lo] == 3).all())
        assert bool((outs[i][lo:] == 0).all()) and bool((counts[i][lo:] == 0).all())
    # the straggler itself: the others' blocks were reduced without it (3 contributors).  Its own block is
    # timing dependent — its control CTA may close the round at 75% before its delayed CTAs reduce it.
    assert bool((counts[3][:lo] == 3).all()) and torch.equal(outs[3][:lo], part[:lo])
    own = counts[3][lo:]
    full = _ref_sum(srcs)
    want_own = torch.where(own == 4, full[lo:], torch.where(own == 3, part[lo:], torch.zeros_like(part[lo:])))
    assert bool(((own == 0) | (own == 3) | (own == 4)).all()) and torch.equal(outs[3][lo:], want_own)
    assert cl.timeouts_seen() == [0] * world


def test_timeout_force_completes_like_catch_up():
    """Thresholds 1 and a rank that never scatters: the wait is bounded, the round is closed with what
    arrived (count 3) — the device form of the catch-up path (AllreduceWorker.scala:100-106)."""
    from akka_allreduce_b200.ops.fused_allreduce import Injection

    world, n = 4, 1 << 12
    cl = _cluster(world, n, 1 << 10, timeout_ms=20)
    srcs = _srcs(world, n, torch.float32, 9)
    outs, counts = cl.run_round(srcs, Injection(mute=1 << 0))
    want = _ref_sum(srcs, ranks=[1, 2, 3])
    for i in range(world):
        assert torch.equal(outs[i], want) and bool((counts[i] == 3).all())
    assert any(t > 0 for t in cl.timeouts_seen())
    # the next round is healthy again: scratch and flags were left consistent
    srcs = _srcs(world, n, torch.float32, 10)
    outs, counts = cl.run_round(srcs)
    assert torch.equal(outs[1], _ref_sum(srcs)) and bool((counts[1] == 4).all())


def test_host_force_flag_closes_round_immediately():
    from akka_allreduce_b200.ops.fused_allreduce import Injection

    world, n = 4, 1 << 12
    cl = _cluster(world, n, 1 << 10, timeout_ms=10_000)
    srcs = _srcs(world, n, torch.float32, 12)
    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    outs, counts = cl.run_round(srcs, Injection(mute=1 << 2), force=True)
    t1.record()
    torch.cuda.synchronize()
    assert t0.elapsed_time(t1) < 2000  # did not sit out the 10 s timeout
    for i in range(world):
        got = counts[i] > 0
        assert bool((counts[i][got] <= 4).all())


def test_matches_host_engine_oracle():
    """Same inputs through the C++ host protocol engine (CPU oracle) and the kernel."""
    from akka_allreduce_b200 import host as h

    world, n, mc = 4, 103, 5
    rng = np.random.default_rng(0)
    data = [rng.integers(-50, 50, n).astype(np.float32) for _ in range(world)]
    bus = h.LocalBus()
    cfg = h.MasterConfig()
    cfg.total_workers, cfg.data_size, cfg.max_chunk_size, cfg.max_round, cfg.max_lag = world, n, mc, 0, 1
    cfg.th_allreduce = cfg.th_reduce = cfg.th_complete = 1.0
    _, master = bus.add_master(cfg)
    got = {}
    for i in range(world):
        ep, _w = bus.add_worker(lambda it, i=i: data[i], lambda d, c, it, i=i: got.__setitem__(i, (d.copy(), c.copy())))
        master.member_up(ep)
    bus.run()
    cl = _cluster(world, n, mc)
    outs, counts = cl.run_round([torch.from_numpy(d).cuda() for d in data])
    for i in range(world):
        assert np.array_equal(outs[i].cpu().numpy(), got[i][0])
        assert np.array_equal(counts[i].cpu().numpy(), got[i][1])


@pytest.mark.parametrize("world,n", [(2, 10), (8, 256), (4, 778), (8, 65536), (5, 100003)])
def test_oneshot_small_messages(world, n):
    """One-shot latency path: every rank sums all copies itself; thresholds 1, ring wrap, fp32-exact."""
    cl = _cluster(world, n, max(1, n // 3), max_lag=1, algo="oneshot", timeout_ms=5000)
    for r in range(5):
        srcs = _srcs(world, n, torch.float32, r)
        outs, counts = cl.run_round(srcs)
        want = _ref_sum(srcs)
        for i in range(world):
            assert torch.equal(outs[i], want) and bool((counts[i] == world).all())
    assert cl.timeouts_seen() == [0] * world and cl.done_tags() == [5] * world


def test_oneshot_timeout_leaves_out_the_silent_rank():
    from akka_allreduce_b200.ops.fused_allreduce import Injection

    world, n = 4, 4096
    cl = _cluster(world, n, 1024, algo="oneshot", timeout_ms=20)
    srcs = _srcs(world, n, torch.float32, 2)
    outs, counts = cl.run_round(srcs, Injection(mute=1 << 1))
    want = _ref_sum(srcs, ranks=[0, 2, 3])
    for i in range(world):
        assert torch.equal(outs[i], want) and bool((counts[i] == 3).all())


def test_cli_worker_cuda_backend_single_gpu():
    """`python -m akka_allreduce_b200.worker --backend cuda` against a CLI master: the reference's launcher
    pair (README.md:3,7) on the CUDA backend, with the built-in ramp source and throughput/assert sink."""
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, PYTHONPATH=root)
    port = 28000 + os.getpid() % 1500
    m = subprocess.Popen([sys.executable, "-m", "akka_allreduce_b200.master", str(port), "1", "4096", "512",
                          "--maxRound", "60", "--thComplete", "1.0"], env=env, stdout=subprocess.PIPE,
                         stderr=subprocess.STDOUT, text=True)
    w = subprocess.Popen([sys.executable, "-m", "akka_allreduce_b200.worker", "0", "4096", "--master-port", str(port),
                          "--checkpoint", "20", "--assert-multiple", "1", "--backend", "cuda"], env=env,
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    try:
        wout = w.communicate(timeout=120)[0]
        mout = m.communicate(timeout=30)[0]
    finally:
        for p in (w, m):
            if p.poll() is None:
                p.kill()
    assert w.returncode == 0, wout[-2000:]
    assert m.returncode == 0, mout[-2000:]
    assert "MBytes/sec" in wout and "----Data output at #40" in wout


def test_ddp_comm_hook_single_rank_matches_plain_backward():
    """DDP + the fused-allreduce communication hook in a one-rank process group: the gradients DDP ends up
    with must equal those of the same model without DDP (mean over one contributor = identity), and the hook
    must actually have launched the kernel for every bucket."""
    import os

    import torch.distributed as dist
    from torch.nn.parallel import DistributedDataParallel as DDP

    from akka_allreduce_b200.models.gpt2 import GPT2, GPT2Config
    from akka_allreduce_b200.parallel.ddp import AllreduceCommHook

    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", str(29800 + os.getpid() % 100))
    dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))
    try:
        torch.manual_seed(0)
        cfg = GPT2Config.tiny()
        ref = GPT2(cfg).cuda()
        model = GPT2(cfg).cuda()
        model.load_state_dict(ref.state_dict())
        ddp = DDP(model, device_ids=[0], bucket_cap_mb=1)
        hook = AllreduceCommHook(0, 1, 0)
        ddp.register_comm_hook(None, hook)
        x = torch.randint(0, cfg.vocab_size, (4, 33), device="cuda")
        for _ in range(2):  # the second iteration runs on DDP's rebuilt buckets
            ref.zero_grad()
            ddp.zero_grad()
            ref(x[:, :-1], x[:, 1:]).backward()
            ddp(x[:, :-1], x[:, 1:]).backward()
            torch.cuda.synchronize()
            for a, b in zip(ref.parameters(), model.parameters()):
                torch.testing.assert_close(b.grad, a.grad, rtol=1e-5, atol=1e-6)
        assert hook.buckets_reduced >= 2 and hook.kernel_launches == hook.buckets_reduced
        hook.close()
    finally:
        dist.destroy_process_group()


def test_random_straggler_stress_counts_always_explain_the_sums():
    """Stress version of the partial-completion checks (the reference has no fault-injection tooling;
    SURVEY.md 5.2/5.3 asks for randomised delays): 8 workers, thReduce = thComplete = 0.75, online-nodes gate
    6 of 8, maxLag 4, and every round a random subset of workers is late or silent.  Each worker's message is
    the constant 2^rank, so a reduced value is the bitmask of its contributors: on every rank, for every
    element, popcount(value) must equal the reported count, and count 0 must come with value 0."""
    import random

    from akka_allreduce_b200.ops.fused_allreduce import Injection

    P, n = 8, 1 << 15
    cl = _cluster(P, n, 1 << 11, max_lag=4, th_reduce=0.75, th_complete=0.75, th_allreduce=0.75, timeout_ms=200)
    srcs = [torch.full((n,), float(1 << r), device="cuda") for r in range(P)]
    rng = random.Random(7)
    seen = set()
    for r in range(120):
        late = rng.sample(range(P), rng.choice([0, 0, 1, 2]))
        inj = Injection(delay_ns={k: rng.choice([20_000, 200_000, 2_000_000]) for k in late[:1]},
                        mute=sum(1 << k for k in late[1:]))
        outs, counts = cl.run_round(srcs, inj)
        torch.cuda.synchronize()
        for i in range(P):
            bits = outs[i].to(torch.int64)
            pop = sum(((bits >> b) & 1) for b in range(P))
            assert torch.equal(pop, counts[i].to(torch.int64)), f"round {r} rank {i}"
            assert bool((bits.to(torch.float32) == outs[i]).all())
        seen.update(counts[0].unique().tolist())
    assert seen <= {0, 6, 7, 8} and 8 in seen and (6 in seen or 7 in seen)
    assert cl.done_tags() == [120] * P


@pytest.mark.parametrize("world,n", [(8, 256), (4, 778), (8, 1 << 18), (5, 100003)])
def test_oneshot_pull_variant(world, n, monkeypatch):
    """Pull form of the one-shot round (stage locally, peers read it; opt-in via AAR_ONESHOT_PULL): same sums,
    same counts, ring wrap.  (On one GPU there is no multicast alias, so this runs its unicast-pull path.)"""
    from akka_allreduce_b200.ops.fused_allreduce import Injection

    monkeypatch.setenv("AAR_ONESHOT_PULL", "2")
    cl = _cluster(world, n, max(1, n // 3), max_lag=1, algo="oneshot", timeout_ms=5000)
    assert cl.launcher.oneshot_pull == 2
    for r in range(5):
        srcs = _srcs(world, n, torch.float32, 40 + r)
        outs, counts = cl.run_round(srcs)
        want = _ref_sum(srcs)
        for i in range(world):
            assert torch.equal(outs[i], want) and bool((counts[i] == world).all())
    assert cl.timeouts_seen() == [0] * world
    # a silent rank is left out after the bounded wait, and the count says so
    cl2 = _cluster(4, 4096, 1024, algo="oneshot", timeout_ms=20)
    srcs = _srcs(4, 4096, torch.float32, 2)
    outs, counts = cl2.run_round(srcs, Injection(mute=1 << 1))
    want = _ref_sum(srcs, ranks=[0, 2, 3])
    assert all(torch.equal(outs[i], want) and bool((counts[i] == 3).all()) for i in range(4))
