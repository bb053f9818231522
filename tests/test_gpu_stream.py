"""The stream kernel (warp tiles, two CTAs per SM: csrc/cuda/stream_kernel.cuh) against the generic kernel and a
plain PyTorch fp32 reference, all ranks emulated on one GPU; plus the dead-peer mask and the one-shot mask."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _cluster(*a, **kw):
    from akka_allreduce_b200.ops.fused_allreduce import EmulatedCluster

    return EmulatedCluster(*a, **kw)


def _srcs(world, n, dtype, r, dev="cuda:0"):
    g = torch.Generator(device="cpu").manual_seed(31 * r + world)
    return [torch.randn(n, generator=g, dtype=torch.float32).to(dtype).to(dev) for _ in range(world)]


def _ref_sum(srcs, ranks=None):
    acc = torch.zeros_like(srcs[0], dtype=torch.float32)
    for i in (range(len(srcs)) if ranks is None else ranks):
        acc += srcs[i].float()
    return acc


@pytest.mark.parametrize("world,n,mc,max_lag,align", [
    (2, 1 << 20, 1 << 16, 1, False),
    (4, 1 << 20, 1 << 15, 2, False),
    (8, 1 << 21, 1 << 16, 1, False),      # 4 chunks per block: phases pipelined chunk by chunk
    (8, 1 << 18, 1 << 18, 1, False),      # one chunk per block
    (3, 3 * 4096 + 1024, 1000, 1, True),  # short last block, chunk size not a multiple of the warp sweep
    (1, 1 << 20, 1 << 16, 1, False),      # a single rank: pure HBM traffic
    (5, 5 * 40960, 8192, 3, False),
    (6, 100000, 4096, 2, True),           # trailing block much shorter than the others
])
def test_stream_kernel_matches_generic_and_fp32_sum(world, n, mc, max_lag, align):
    st = _cluster(world, n, mc, max_lag=max_lag, timeout_ms=5000, kernel="stream", align_blocks=align)
    ge = _cluster(world, n, mc, max_lag=max_lag, timeout_ms=5000, kernel="generic", align_blocks=align)
    assert st.kernel_used() == "stream" and ge.kernel_used() == "generic"
    for r in range(2 * (max_lag + 1) + 1):
        srcs = _srcs(world, n, torch.float32, r)
        outs, counts = st.run_round(srcs)
        outs_g, counts_g = ge.run_round(srcs)
        want = _ref_sum(srcs)
        for i in range(world):
            assert torch.equal(outs[i], want), f"round {r} rank {i}: max err {(outs[i] - want).abs().max()}"
            assert torch.equal(outs[i], outs_g[i]) and torch.equal(counts[i], counts_g[i])
            assert bool((counts[i] == world).all())
    assert st.timeouts_seen() == [0] * world
    assert st.done_tags() == [st.round] * world


def test_stream_kernel_is_the_default_for_aligned_shapes():
    assert _cluster(8, 1 << 18, 1 << 14).kernel_used() == "stream"
    assert _cluster(4, 778, 3).kernel_used() == "generic"                                   # unaligned blocks
    assert _cluster(8, 1 << 18, 1 << 14, dtype=torch.bfloat16, wire_dtype=torch.float32,
                    align_blocks=True).kernel_used() == "generic"                            # converting wire


def test_stream_kernel_bf16_scale_normalize():
    world, n = 8, 1 << 19
    cl = _cluster(world, n, 1 << 15, dtype=torch.bfloat16, align_blocks=True, kernel="stream")
    srcs = _srcs(world, n, torch.bfloat16, 3)
    outs, counts = cl.run_round(srcs)
    want = _ref_sum(srcs)
    for i in range(world):
        torch.testing.assert_close(outs[i].float(), want, rtol=2 ** -7, atol=1e-2)
        assert bool((counts[i] == world).all())
    assert all(torch.equal(outs[0], outs[i]) for i in range(1, world))
    cl = _cluster(4, 1 << 16, 1 << 13, scale=0.5, normalize=True, kernel="stream")
    srcs = _srcs(4, 1 << 16, torch.float32, 7)
    outs, _ = cl.run_round(srcs)
    want = sum((s * 0.5 for s in srcs[1:]), srcs[0] * 0.5) / 4
    torch.testing.assert_close(outs[2], want, rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("kernel", ["stream", "generic"])
def test_partial_thresholds_and_straggler(kernel):
    """thReduce 0.75 with a muted rank, then thComplete 0.75 with an owner that starts 30 ms late."""
    from akka_allreduce_b200.ops.fused_allreduce import Injection

    world, n = 4, 1 << 16
    cl = _cluster(world, n, 1 << 12, th_reduce=0.75, th_complete=1.0, max_lag=2, kernel=kernel)
    for r in range(4):
        srcs = _srcs(world, n, torch.float32, r)
        outs, counts = cl.run_round(srcs, Injection(mute=1 << 3))
        want = _ref_sum(srcs, ranks=[0, 1, 2])
        for i in range(world):
            assert torch.equal(outs[i], want) and bool((counts[i] == 3).all())
    assert cl.timeouts_seen() == [0] * world
    cl = _cluster(world, 1 << 14, 1 << 10, th_reduce=0.75, th_complete=0.75, timeout_ms=5000, kernel=kernel)
    srcs = _srcs(world, 1 << 14, torch.float32, 5)
    outs, counts = cl.run_round(srcs, Injection(delay_ns={3: 30_000_000}))
    part = _ref_sum(srcs, ranks=[0, 1, 2])
    lo = cl.geo.start(3)
    for i in range(3):
        assert torch.equal(outs[i][:lo], part[:lo]) and bool((counts[i][:lo] == 3).all())
        assert bool((outs[i][lo:] == 0).all()) and bool((counts[i][lo:] == 0).all())


@pytest.mark.parametrize("kernel", ["stream", "generic"])
def test_dropped_scatter_closes_on_the_clock_with_true_counts(kernel):
    from akka_allreduce_b200.ops.fused_allreduce import Injection

    world, n = 4, 4096
    cl = _cluster(world, n, 256, timeout_ms=20, kernel=kernel)
    srcs = _srcs(world, n, torch.float32, 11)
    outs, counts = cl.run_round(srcs, Injection(skip={1: 1 << 2}))
    full, part = _ref_sum(srcs), _ref_sum(srcs, ranks=[0, 2, 3])
    lo, hi = cl.geo.start(2), cl.geo.end(2)
    for i in range(world):
        assert torch.equal(outs[i][:lo], full[:lo]) and torch.equal(outs[i][hi:], full[hi:])
        assert torch.equal(outs[i][lo:hi], part[lo:hi])
        assert bool((counts[i][lo:hi] == 3).all()) and bool((counts[i][:lo] == 4).all())
    assert any(cl.timeouts_seen())


@pytest.mark.parametrize("kernel,algo", [("stream", "twoshot"), ("generic", "twoshot"), (None, "oneshot")])
def test_dead_peer_is_not_waited_for(kernel, algo):
    """A rank the host marked dead (coordinator: Terminated -> refreshed InitWorkers) is neither pushed to nor
    waited for: with every threshold at 1 and a 20 s timeout the round still closes at once, without a timeout,
    with count 3 — and the dead owner's block comes out as zeros with count 0 (nobody reduces it)."""
    import time

    world, n = 4, 1 << 14
    cl = _cluster(world, n, 1 << 10, max_lag=1, timeout_ms=20_000, kernel=kernel, algo=algo)
    dead = 1 << 3
    # in emulation rank 3 still runs (it is part of the launch): it is told that everybody else is dead, so it
    # talks to nobody, and nobody pushes to it
    cl.set_dead_peers({0: dead, 1: dead, 2: dead, 3: 0b0111})
    srcs = _srcs(world, n, torch.float32, 2)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    outs, counts = cl.run_round(srcs)
    torch.cuda.synchronize()
    assert time.perf_counter() - t0 < 5.0
    assert cl.timeouts_seen() == [0] * world
    part = _ref_sum(srcs, ranks=[0, 1, 2])
    lo = cl.geo.start(3)
    for i in range(3):
        if algo == "oneshot":   # every rank sums what it holds: the whole message from the three live ranks
            assert torch.equal(outs[i], part) and bool((counts[i] == 3).all())
        else:
            assert torch.equal(outs[i][:lo], part[:lo]) and bool((counts[i][:lo] == 3).all())
            assert bool((outs[i][lo:] == 0).all()) and bool((counts[i][lo:] == 0).all())


def test_all_gather_half_round():
    """The broadcast half on its own (AllreduceWorker.scala:252-268): every rank's block lands in everybody's output,
    count 1; then ordinary allreduce rounds keep working on the same ring."""
    from akka_allreduce_b200.ops import cuda as K

    world, n = 4, 1 << 16
    cl = _cluster(world, n, 1 << 12, max_lag=1, kernel="stream")
    full = _srcs(world, n, torch.float32, 1)
    for r in range(3):
        blocks = [full[i][cl.geo.start(i):cl.geo.end(i)].contiguous() + r for i in range(world)]
        outs, counts = cl.run_round(blocks, op=K.OP_ALL_GATHER)
        want = torch.cat(blocks)
        for i in range(world):
            assert torch.equal(outs[i], want) and bool((counts[i] == 1).all())
    outs, counts = cl.run_round(full)
    assert torch.equal(outs[0], _ref_sum(full)) and bool((counts[0] == world).all())
    assert cl.timeouts_seen() == [0] * world


@pytest.mark.parametrize("n,mc,max_lag", [(1 << 20, 1 << 16, 1), (1 << 18, 1 << 18, 2), (5 * 4096 + 512, 4096, 1)])
def test_direct_two_ranks(n, mc, max_lag):
    """Direct (streamed one-shot for two ranks): push my message to the peer chunk by chunk, sum mine + the peer's
    straight into the output; bit-identical to the two-shot result (same rank order)."""
    world = 2
    cl = _cluster(world, n, mc, max_lag=max_lag, timeout_ms=5000, algo="direct")
    assert cl.kernel_used() == "stream"
    for r in range(2 * (max_lag + 1) + 1):
        srcs = _srcs(world, n, torch.float32, r)
        outs, counts = cl.run_round(srcs)
        want = _ref_sum(srcs)
        for i in range(world):
            assert torch.equal(outs[i], want), f"round {r} rank {i}: max err {(outs[i] - want).abs().max()}"
            assert bool((counts[i] == world).all())
    assert cl.timeouts_seen() == [0] * world
    assert cl.done_tags() == [cl.round] * world


def test_direct_peer_that_never_shows_up():
    """A muted peer: the round closes on the clock with the rank's own data and count 1."""
    from akka_allreduce_b200.ops.fused_allreduce import Injection

    cl = _cluster(2, 1 << 14, 1 << 12, timeout_ms=20, algo="direct")
    srcs = _srcs(2, 1 << 14, torch.float32, 4)
    outs, counts = cl.run_round(srcs, Injection(mute=0b10))
    assert torch.equal(outs[0], srcs[0]) and bool((counts[0] == 1).all())
    assert torch.equal(outs[1], srcs[0]) and bool((counts[1] == 1).all())   # the muted rank still hears its peer


@pytest.mark.parametrize("kernel,algo,world,n,mc,align", [
    ("stream", "twoshot", 4, 1 << 14, 1 << 11, False), ("generic", "twoshot", 4, 1 << 14, 1 << 11, False),
    ("stream", "twoshot", 3, 3 * 4096 + 1024, 1000, True), ("generic", "twoshot", 4, 778, 3, False),
    (None, "oneshot", 4, 5000, 4096, False), ("stream", "direct", 2, 5 * 4096 + 512, 4096, False),
])
def test_no_write_outside_the_callers_buffers(kernel, algo, world, n, mc, align):
    """Guard bands around every output and count buffer stay untouched and the sources are not modified (the pool's
    compute-sanitizer is closed: this is the bounds check the suite can make on its own)."""
    G = 1024
    cl = _cluster(world, n, mc, max_lag=1, timeout_ms=5000, kernel=kernel, algo=algo, align_blocks=align)
    big_o = [torch.full((n + 2 * G,), -777.0, device="cuda:0") for _ in range(world)]
    big_c = [torch.full((n + 2 * G,), -777, dtype=torch.int32, device="cuda:0") for _ in range(world)]
    outs = [b[G:G + n] for b in big_o]
    cnts = [b[G:G + n] for b in big_c]
    for r in range(3):
        srcs = _srcs(world, n, torch.float32, r)
        keep = [s.clone() for s in srcs]
        cl.run_round(srcs, outs=outs, counts=cnts)
        torch.cuda.synchronize()
        want = _ref_sum(srcs)
        for i in range(world):
            assert torch.equal(outs[i], want) and bool((cnts[i] == world).all())
            assert bool((big_o[i][:G] == -777.0).all()) and bool((big_o[i][G + n:] == -777.0).all())
            assert bool((big_c[i][:G] == -777).all()) and bool((big_c[i][G + n:] == -777).all())
            assert torch.equal(srcs[i], keep[i])


def test_single_rank_engine_with_registered_buffers():
    """One real rank (no emulation): ordinary tensors, then source and sink aliased to the engine's own slots — the
    zero-copy path raises its ScatterBlock flags from the publisher warp at launch (nothing is scattered)."""
    import tempfile

    from akka_allreduce_b200.parallel.engine import DeviceAllreduce

    n = 1 << 20
    eng = DeviceAllreduce(0, 1, n, 1 << 16, max_lag=1, th_allreduce=1.0, device=0, uds_dir=tempfile.mkdtemp(),
                          timeout_ms=2000)
    try:
        src = torch.randn(n, device="cuda:0")
        out = torch.empty_like(src)
        cnt = torch.empty(n, dtype=torch.int32, device="cuda:0")
        for r in range(3):
            eng.allreduce(src + r, out, cnt)
            torch.cuda.synchronize()
            assert torch.equal(out, src + r) and bool((cnt == 1).all())
        for r in range(5):
            iv, ov = eng.input_view(eng.round), eng.output_view(eng.round)
            assert iv is not None and ov is not None
            iv.copy_(src * (r + 2))
            eng.allreduce(iv, ov, cnt)
            torch.cuda.synchronize()
            assert torch.equal(ov, src * (r + 2)) and bool((cnt == 1).all())
        assert eng.timeouts_seen() == 0
    finally:
        eng.close()


def test_message_length_not_a_multiple_of_the_unit_goes_to_the_generic_kernel():
    """Aligned blocks but a message whose last block ends in a partial 16-byte unit: the stream kernel moves whole
    units only, so the library must pick the generic kernel (scalar tail) — and the tail must come out right."""
    world, n = 4, 100003
    cl = _cluster(world, n, 4096, max_lag=1, timeout_ms=5000, align_blocks=True)
    assert cl.kernel_used() == "generic"
    for r in range(3):
        srcs = _srcs(world, n, torch.float32, r)
        outs, counts = cl.run_round(srcs)
        want = _ref_sum(srcs)
        for i in range(world):
            assert torch.equal(outs[i], want) and bool((counts[i] == world).all())
    with pytest.raises(Exception, match="stream kernel needs"):
        _cluster(world, n, 4096, align_blocks=True, kernel="stream").run_round(_srcs(world, n, torch.float32, 0))


@pytest.mark.parametrize("kernel", ["stream", "generic"])
def test_rounds_of_different_lengths_share_one_region(kernel):
    """One region laid out for the largest message serves shorter rounds too (how the DDP hook shares one heap across
    bucket sizes): strides and flags from the capacity, partition from the actual length, one ring for all of them."""
    world, cap, mc = 4, 1 << 18, 1 << 13
    cl = _cluster(world, cap, mc, max_lag=1, timeout_ms=5000, kernel=kernel, align_blocks=True)
    for r, m in enumerate([cap, 1 << 16, 5 * 4096 + 512, cap, 1024, 3 * (1 << 14) + 128, 512, cap]):
        srcs = _srcs(world, m, torch.float32, r)
        outs, counts = cl.run_part(srcs) if m != cap else cl.run_round(srcs)
        want = _ref_sum(srcs)
        for i in range(world):
            assert torch.equal(outs[i][:m], want), f"round {r} length {m} rank {i}"
            assert bool((counts[i][:m] == world).all())
    assert cl.timeouts_seen() == [0] * world
