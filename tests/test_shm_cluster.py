"""The device protocol restated for CPU threads in shared memory (csrc/host/shm_cluster.h): same region layout, same
partition / chunk / threshold arithmetic and the same rules as csrc/cuda/fused_allreduce.cu, checked here without a
GPU — against plain sums, against the message-level host engine, and under stragglers (sums must name exactly the
contributors the counts report).  scripts/sanitize_host.sh runs the native form under ThreadSanitizer."""
import numpy as np
import pytest

from akka_allreduce_b200 import host
from akka_allreduce_b200.ops.geometry import Geometry


def make(world, n, mc, lag, th=(1.0, 1.0, 1.0), timeout_ms=2000.0, notice=True):
    c = host.ShmConfig()
    c.world, c.n, c.max_chunk, c.max_lag = world, n, mc, lag
    c.th_allreduce, c.th_reduce, c.th_complete = th
    c.timeout_ms, c.skip_notice = timeout_ms, notice
    return host.ShmCluster(c)


@pytest.mark.parametrize("world,n,mc,lag", [(2, 10, 2, 1), (4, 778, 3, 3), (4, 9, 2, 1), (3, 17, 100, 0),
                                             (8, 1000, 7, 2), (5, 100003, 4096, 2), (1, 33, 4, 1), (16, 64, 1, 1)])
def test_full_thresholds_give_exact_sums_over_many_rounds(world, n, mc, lag):
    """Reference configs (BASELINE 1 and the manual 4-worker run), uneven partitions (n=9, P=4: an empty trailing
    block), one chunk per block, one element per chunk; the ring wraps several times."""
    cl = make(world, n, mc, lag)
    geo = Geometry.make(n, mc, world, lag, False)
    assert (cl.req_reduce, cl.total_chunks, cl.req_complete) == (world, geo.total_chunks, geo.total_chunks)
    rng = np.random.default_rng(7)
    for rnd in range(3 * (lag + 1) + 1):
        srcs = [rng.standard_normal(n).astype(np.float32) for _ in range(world)]
        outs, counts, infos = cl.run_round(rnd, srcs)
        want = np.zeros(n, dtype=np.float32)
        for s in srcs:                      # peer order, fp32, like the kernel and the reference
            want = want + s
        for r in range(world):
            assert np.array_equal(outs[r], want), (rnd, r)
            assert (counts[r] == world).all()
            assert not infos[r]["timed_out"] and infos[r]["reduced_partial"] == 0
        assert all(np.array_equal(outs[0], o) for o in outs)       # bit-identical on every rank


def test_matches_the_message_level_engine():
    """Same inputs through the C++ protocol engine (the reference restated message for message) and through the
    shared-memory device protocol: identical outputs and counts with thresholds 1."""
    from test_master_and_cluster import run_cluster

    P, n, mc = 4, 778, 3
    master, workers, outs_ref = run_cluster(P, n, mc, 3, 4, source=lambda it: np.arange(n, dtype=np.float32) * (it + 1))
    cl = make(P, n, mc, 3)
    for it in range(5):
        outs, counts, _ = cl.run_round(it, [np.arange(n, dtype=np.float32) * (it + 1)] * P)
        d_ref, c_ref = outs_ref[0][it]
        assert np.array_equal(outs[0], d_ref) and np.array_equal(counts[0], c_ref)


def test_muted_rank_and_threshold_below_one():
    """thReduce 0.75 on 4 ranks, rank 3 silent: every chunk is reduced from ranks 0..2 without waiting for a timeout,
    the count says 3, and rank 3 itself still reduces its block and receives everybody's."""
    cl = make(4, 64, 8, 1, th=(1.0, 0.75, 1.0), timeout_ms=5000.0)
    srcs = [np.full(64, float(1 << r), dtype=np.float32) for r in range(4)]
    outs, counts, infos = cl.run_round(0, srcs, mute_mask=1 << 3)
    for r in range(4):
        assert (outs[r] == 7.0).all() and (counts[r] == 3).all()
        assert not infos[r]["timed_out"] and infos[r]["reduced_partial"] == 2


def test_dropped_destination_and_timeout_with_full_thresholds():
    """Rank 1 never sends to rank 0 (skip_dest): with thReduce 1 rank 0 waits for its timeout, then reduces its block
    from ranks 0 and 2 only; the other blocks are complete.  Counts are per element."""
    cl = make(3, 30, 5, 1, timeout_ms=150.0)
    srcs = [np.full(30, float(1 << r), dtype=np.float32) for r in range(3)]
    outs, counts, infos = cl.run_round(0, srcs, skip_dest=[0, 1 << 0, 0])
    assert infos[0]["timed_out"] and infos[0]["reduced_partial"] == 2
    for r in range(3):
        assert (outs[r][:10] == 5.0).all() and (counts[r][:10] == 2).all()       # block 0: ranks 0 and 2
        assert (outs[r][10:] == 7.0).all() and (counts[r][10:] == 3).all()


def test_late_rank_is_left_out_with_partial_thresholds_but_waited_for_with_full_ones():
    srcs = [np.full(40, float(1 << r), dtype=np.float32) for r in range(4)]
    cl = make(4, 40, 5, 1, th=(0.75, 0.75, 0.75))
    outs, counts, _ = cl.run_round(0, srcs, delay_us=[0, 0, 0, 30000])
    # (3 contributors, or nothing for chunks past thComplete; 4 only if the scheduler held an owner back for 30 ms)
    assert np.isin(counts[0], (0, 3, 4)).all() and (counts[0] == 3).any()
    assert (outs[0][counts[0] == 3] == 7.0).all() and (outs[0][counts[0] == 4] == 15.0).all()
    cl = make(4, 40, 5, 1)
    outs, counts, _ = cl.run_round(0, srcs, delay_us=[0, 0, 0, 30000])
    assert (counts[0] == 4).all() and (outs[0] == 15.0).all()


def test_round_tags_never_repeat():
    cl = make(2, 8, 4, 1)
    srcs = [np.ones(8, dtype=np.float32)] * 2
    cl.run_round(0, srcs)
    cl.run_round(1, srcs)
    with pytest.raises(ValueError, match="already issued"):
        cl.run_round(1, srcs)
    assert cl.next_round == 2
    assert cl.run_free(50)["mismatches"] == 0 and cl.next_round == 52       # continues after the last round


@pytest.mark.parametrize("world,lag,th", [(4, 3, 0.75), (8, 2, 0.75), (8, 1, 0.75), (3, 0, 0.7), (5, 1, 1.0)])
def test_free_running_ranks_with_stragglers(world, lag, th):
    """Every rank runs rounds as fast as the protocol lets it, with random stalls; rank q contributes 2^q, so every
    output element names its contributors and must agree with its count.  With skip notices nothing waits for a
    timeout."""
    cl = make(world, 1024, 64, lag, th=(th, th, th), timeout_ms=2000.0)
    r = cl.run_free(400, stall_prob=0.02, stall_us=200, seed=11)
    assert r["mismatches"] == 0 and r["timeouts"] == 0, r
    if th == 1.0:
        assert r["partial_elements"] == 0 and r["missing_elements"] == 0


def test_skip_notices_replace_timeouts():
    """Same straggler run without the notices (the kernel's current protocol): a rank that was left out by its peers
    learns it only from the clock.  Still correct, just slow."""
    kw = dict(stall_prob=0.02, stall_us=200, seed=3)
    without = make(6, 512, 64, 1, th=(0.7, 0.7, 0.7), timeout_ms=20.0, notice=False).run_free(300, **kw)
    with_ = make(6, 512, 64, 1, th=(0.7, 0.7, 0.7), timeout_ms=5000.0, notice=True).run_free(300, **kw)
    assert without["mismatches"] == 0 and with_["mismatches"] == 0
    assert with_["timeouts"] == 0 and with_["seconds"] < 60       # nothing sat out a 5 s timeout


def test_native_shm_cluster_binary():
    import os
    import subprocess

    from akka_allreduce_b200.build import build_shm_cluster

    exe = str(build_shm_cluster())
    # (an uninstrumented binary under a pre-loaded sanitizer runtime only shows it the memcpys, not the flags)
    env = {k: v for k, v in os.environ.items() if k != "LD_PRELOAD"}
    for argv in (["4", "778", "3", "3", "500"], ["8", "4096", "64", "2", "300", "0.75", "0.75", "0.75", "0.01", "100", "500"]):
        p = subprocess.run([exe] + argv, capture_output=True, text=True, timeout=120, env=env)
        assert p.returncode == 0 and "0 mismatches" in p.stdout and "0 timeouts" in p.stdout, p.stdout + p.stderr


def test_scale_and_normalize_like_the_kernel():
    """`scale` multiplies every contribution on its way out, `normalize` divides a reduced chunk by its contributor
    count: with a silent rank the mean is over the three that spoke (what the DDP hook relies on)."""
    c = host.ShmConfig()
    c.world, c.n, c.max_chunk, c.max_lag = 4, 48, 6, 1
    c.th_allreduce, c.th_reduce, c.th_complete = 1.0, 0.75, 1.0
    c.scale, c.normalize, c.timeout_ms = 0.5, True, 5000.0
    cl = host.ShmCluster(c)
    srcs = [np.full(48, float(r + 1), dtype=np.float32) for r in range(4)]
    outs, counts, _ = cl.run_round(0, srcs, mute_mask=1 << 3)
    assert (counts[0] == 3).all() and np.allclose(outs[0], 0.5 * (1 + 2 + 3) / 3)
    outs, counts, _ = cl.run_round(1, srcs)
    full = counts[0] == 4            # a chunk fires as soon as 3 are in: 3 or 4 contributors, mean either way
    assert np.isin(counts[0], (3, 4)).all() and np.allclose(outs[0][full], 0.5 * 10 / 4)
    three = {round(0.5 * (10 - v) / 3, 5) for v in (1, 2, 3, 4)}        # whichever rank was not waited for
    assert all(round(float(x), 5) in three for x in outs[0][~full])


def test_any_shape_matches_plain_sums():
    """Property test over the geometry: any (world, n, maxChunkSize, maxLag), including n < world (empty blocks),
    chunks larger than blocks and prime sizes — the region layout and the partition arithmetic are the kernel's own
    headers, so this also walks `SymLayout` / `Partition` through shapes the fixed cases do not reach."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=40, deadline=None)
    @given(st.integers(1, 9), st.integers(1, 700), st.integers(1, 97), st.integers(0, 3), st.integers(0, 2 ** 31))
    def run(world, n, mc, lag, seed):
        cl = make(world, n, mc, lag, timeout_ms=10000.0)
        rng = np.random.default_rng(seed)
        for rnd in range(lag + 2):
            srcs = [rng.integers(-50, 50, n).astype(np.float32) for _ in range(world)]      # exact in fp32
            outs, counts, infos = cl.run_round(rnd, srcs)
            want = np.sum(srcs, axis=0, dtype=np.float32)
            for r in range(world):
                assert np.array_equal(outs[r], want) and (counts[r] == world).all(), (world, n, mc, lag, rnd, r)
                assert not infos[r]["timed_out"]

    run()


def test_dead_peer_is_neither_pushed_to_nor_waited_for():
    """The kernel's dead-peer mask in the shared-memory form of the protocol: rank 3 never runs; once the others have
    been told, every threshold is still 1 and the timeout 5 s, yet rounds close at once with count 3 and zeros /
    count 0 on the dead owner's block (AllreduceMaster.scala:46-52: the master drops the worker, AllreduceWorker.scala:
    213-216: the others skip it)."""
    import threading
    import time

    from akka_allreduce_b200 import host

    c = host.ShmConfig()
    c.world, c.n, c.max_chunk, c.max_lag = 4, 64, 4, 1
    c.th_reduce = c.th_complete = 1.0
    c.th_allreduce, c.timeout_ms = 0.75, 5000.0
    cl = host.ShmCluster(c)
    for r in range(3):
        cl.set_dead_peers(r, 1 << 3)
    results = {}

    def run(rank):
        src = np.full(64, float(1 << rank), dtype=np.float32)
        for rnd in range(6):
            results[(rank, rnd)] = cl.rank_round(rank, rnd, src)

    t0 = time.perf_counter()
    threads = [threading.Thread(target=run, args=(r,)) for r in range(3)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(20)
    assert time.perf_counter() - t0 < 4.0 and len(results) == 18      # nobody sat out the 5 s timeout
    for (rank, rnd), (out, cnt, info) in results.items():
        assert not info["timed_out"]
        assert (out[:48] == 7.0).all() and (cnt[:48] == 3).all()       # 2^0 + 2^1 + 2^2 from exactly three contributors
        assert (out[48:] == 0.0).all() and (cnt[48:] == 0).all()       # nobody owns block 3 any more
