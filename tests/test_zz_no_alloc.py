"""The entry point that brings its own scratch (``aar_allreduce_round_static_scratch``).

Some users of the C ABI cannot allocate device memory when a round is due: a server that gave all of
HBM to its own pool at start-up, a loop that must not take cudaMalloc's device-wide stall, an embedder
without an allocator to offer.  For them the kernel's scratch sits in a fixed array inside the library.
This file checks, through ctypes, that a round launched that way produces what the ordinary entry point
produces — twice in a row, because the array is shared and keeps the previous call's contents — and
that a configuration too big for the array is refused with a message instead of overrunning it.

It is collected last on purpose (``zz``): a failure here must not hide any other test.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _cluster(world, n, mc, **kw):
    from akka_allreduce_b200.ops.fused_allreduce import EmulatedCluster

    return EmulatedCluster(world, n, mc, max_lag=1, timeout_ms=5000, **kw)


def _launch_static(cl, srcs, outs, counts, round_):
    """Same launch struct the cluster would use, sent to the no-allocation entry point instead."""
    from akka_allreduce_b200.ops import cuda as K

    a = cl._a
    cl.launcher.fill(a, round_, None)
    for i in range(cl.world):
        a.src[i], a.out[i], a.counts[i] = srcs[i].data_ptr(), outs[i].data_ptr(), counts[i].data_ptr()
    stream = torch.cuda.current_stream()
    rc = K.lib().aar_allreduce_round_static_scratch(C.byref(a), C.c_void_p(stream.cuda_stream))
    return rc


@pytest.mark.parametrize("world,n,mc", [(4, 778, 3), (8, 1 << 18, 1 << 14)])
def test_static_scratch_matches_ordinary_entry_point_twice(world, n, mc):
    from akka_allreduce_b200.ops import cuda as K

    cl = _cluster(world, n, mc)
    outs = [torch.empty(n, device="cuda") for _ in range(world)]
    counts = [torch.empty(n, dtype=torch.int32, device="cuda") for _ in range(world)]
    rnd = 0
    for call in range(2):  # two calls in a row share the built-in array
        g = torch.Generator(device="cpu").manual_seed(31 + call)
        srcs = [torch.randn(n, generator=g).cuda() for _ in range(world)]
        ref_out, ref_cnt = cl.run_round(srcs, round_=rnd)          # ordinary entry point, caller's scratch
        ref_out = [t.clone() for t in ref_out]
        ref_cnt = [t.clone() for t in ref_cnt]
        rnd += 1
        rc = _launch_static(cl, srcs, outs, counts, rnd)            # same inputs, built-in scratch
        assert rc > 0, K.lib().aar_last_error().decode()
        rnd += 1
        torch.cuda.synchronize()
        for i in range(world):
            assert torch.equal(outs[i], ref_out[i]), f"call {call} rank {i}"
            assert torch.equal(counts[i], ref_cnt[i])
    assert cl.timeouts_seen() == [0] * world


def test_static_scratch_refuses_what_does_not_fit():
    from akka_allreduce_b200.ops import cuda as K

    world, n, mc = 8, 1 << 20, 16          # 8192 chunks per block: far more counters than 16 KB holds
    lib = K.lib()
    assert lib.aar_scratch_bytes(n, mc, world, 0) > 16384
    a = K.AarLaunch()
    a.world, a.n, a.max_chunk, a.block_step, a.emu = world, n, mc, 0, 1
    rc = lib.aar_allreduce_round_static_scratch(C.byref(a), None)
    assert rc == -4  # AAR_ETOOBIG
    msg = lib.aar_last_error().decode()
    assert "16384" in msg and "scratch" in msg
