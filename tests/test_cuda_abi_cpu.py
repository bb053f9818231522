"""Checks of the CUDA library that need no GPU: it loads without a driver, its C struct matches the ctypes
mirror, and the Python geometry mirror agrees with the C++ layout/threshold code."""
import ctypes as C

import pytest

from akka_allreduce_b200 import host
from akka_allreduce_b200.ops import cuda as K
from akka_allreduce_b200.ops.geometry import Geometry, gate_count, threshold_count


def test_library_loads_without_gpu_and_struct_matches():
    lib = K.lib()
    assert lib.aar_abi_sizeof_launch() == C.sizeof(K.AarLaunch)


@pytest.mark.parametrize("th,n", [(0.7, 10), (0.7, 9), (0.8, 6), (0.8, 4), (0.75, 8), (0.9, 2), (1.0, 8), (0.6, 2), (0.5, 4)])
def test_threshold_mirror_matches_cpp(th, n):
    assert threshold_count(th, n) == host.threshold_count(th, n)


@pytest.mark.parametrize("total,th", [(8, 0.75), (8, 1.0), (4, 0.75), (2, 1.0), (3, 0.5), (8, 0.8)])
def test_gate_count_is_first_count_that_passes_master_gate(total, th):
    k = gate_count(th, total)
    assert host.master_gate(k, total, th) and (k == 0 or not host.master_gate(k - 1, total, th))


@pytest.mark.parametrize("n,mc,P,lag", [(10, 2, 2, 1), (778, 3, 4, 3), (9, 2, 4, 1), (1 << 24, 1 << 18, 8, 1), (100003, 4096, 5, 2)])
def test_geometry_mirror_matches_cpp_partition(n, mc, P, lag):
    g = Geometry.make(n, mc, P, lag)
    part = host.Partition(n, P)
    assert [g.size(i) for i in range(P)] == [part.size(i) for i in range(P)]
    assert g.total_chunks == part.total_reduced_chunks(mc)
    lib = K.lib()
    # layout sizes are pure functions of the configuration: callable on the CPU
    assert lib.aar_sym_bytes(n, mc, P, lag + 1, K.F32, 0, 0) > 2 * (lag + 1) * n * 4
    assert lib.aar_sym_bytes(n, mc, P, lag + 1, K.BF16, 0, 0) <= lib.aar_sym_bytes(n, mc, P, lag + 1, K.F32, 0, 0)
    assert 0 < lib.aar_scratch_force_offset(n, mc, P, 0) < lib.aar_scratch_bytes(n, mc, P, 0)
    ranges = list(g.chunk_ranges())
    assert len(ranges) == g.total_chunks and ranges[-1][3] == n and all(a < b for _, _, a, b in ranges)


def test_static_scratch_limit_is_checked_before_any_cuda_call():
    """The no-allocation entry point validates the configuration first, so the refusal needs no GPU."""
    a = K.AarLaunch()
    a.world, a.n, a.max_chunk, a.block_step, a.emu = 8, 1 << 20, 16, 0, 1
    assert K.lib().aar_allreduce_round_static_scratch(C.byref(a), None) == -4


def test_in_place_round_is_refused_below_full_thresholds():
    """The guard sits in front of any CUDA call, so it can be checked without a device: in place is fine with
    thresholds 1 (a block of `out` is written only after its owner has this rank's slice of it) and refused below."""
    import types

    import torch

    from akka_allreduce_b200.parallel.engine import DeviceAllreduce

    eng = object.__new__(DeviceAllreduce)
    eng.geo = types.SimpleNamespace(n=64, world=4, total_chunks=8)
    eng.launcher = types.SimpleNamespace(dtype=torch.float32, req_reduce=3, req_complete=8)
    assert eng.partial_thresholds
    x = torch.zeros(64)
    with pytest.raises(ValueError, match="in place"):
        eng.allreduce(x, x)
    big = torch.zeros(96)
    with pytest.raises(ValueError, match="in place"):
        eng.allreduce(big[:64], big[32:])          # partial overlap counts too
    eng.launcher.req_reduce = 4
    assert not eng.partial_thresholds


@pytest.mark.parametrize("n,mc,P,lag", [(10, 2, 2, 1), (778, 3, 4, 3), (9, 2, 4, 1), (1 << 20, 1 << 16, 8, 1), (100003, 4096, 5, 2)])
def test_shared_memory_cluster_uses_the_kernels_region_layout(n, mc, P, lag):
    """host.ShmCluster (CPU threads) and libaar_cuda (the kernel) size one rank's symmetric region from the same
    header: same bytes, same thresholds, same chunk totals."""
    lib = K.lib()
    c = host.ShmConfig()
    c.world, c.n, c.max_chunk, c.max_lag = P, n, mc, lag
    c.th_reduce, c.th_complete = 0.75, 0.8
    cl = host.ShmCluster(c)
    assert cl.region_bytes == lib.aar_sym_bytes(n, mc, P, lag + 1, K.F32, 0, 0)
    geo = Geometry.make(n, mc, P, lag, False)
    assert (cl.total_chunks, cl.req_reduce, cl.req_complete) == (geo.total_chunks, geo.req_reduce(0.75), geo.req_complete(0.8))
    assert cl.gate_count == gate_count(1.0, P)


def _launch(**kw):
    a = K.AarLaunch()
    a.world, a.rank, a.n, a.max_chunk, a.L, a.slot, a.tag, a.ctas = 4, 0, 1 << 16, 1 << 12, 2, 0, 1, 8
    a.req_reduce, a.req_complete = 4, 16
    for k, v in kw.items():
        setattr(a, k, v)
    return a


@pytest.mark.parametrize("kw,needle", [
    (dict(slot=2), "slot"), (dict(slot=-1), "slot"), (dict(rank=4), "rank"), (dict(rank=-1), "rank"),
    (dict(world=0), "world"), (dict(world=17), "world"), (dict(tag=0), "tag"), (dict(ctas=0), "cta"),
    (dict(os_cap=-1), "one-shot"), (dict(block_step=1), "block_step"),
])
def test_launch_descriptions_are_validated_before_any_cuda_call(kw, needle):
    """A public C ABI must not turn a bad slot or rank into out-of-bounds stores on a peer (round-1 review): the
    checks sit in front of everything, so they can be exercised without a GPU."""
    lib = K.lib()
    a = _launch(**kw)
    assert lib.aar_round_kernel(C.byref(a)) == -1                     # AAR_EINVAL
    assert needle in lib.aar_last_error().decode()
    assert lib.aar_allreduce_round(C.byref(a), None) == -1


def test_scratch_words_the_host_writes():
    """Force tag and dead-peer mask are device words inside the scratch, next to each other, never reset by a kernel."""
    lib = K.lib()
    for n, mc, P in ((1 << 20, 1 << 16, 8), (778, 3, 4)):
        force, dead, total = (lib.aar_scratch_force_offset(n, mc, P, 0), lib.aar_scratch_dead_offset(n, mc, P, 0),
                              lib.aar_scratch_bytes(n, mc, P, 0))
        assert dead == force + 8 and dead + 4 <= total and force % 4 == 0


@pytest.mark.parametrize("kw,needle", [
    (dict(n=1 << 16, n_cap=1 << 15, block_step=1 << 14, block_step_cap=1 << 13), "capacity"),           # longer than the capacity
    (dict(n_cap=1 << 18), "explicit block steps"),                                                   # no steps given
    (dict(n_cap=1 << 18, block_step=1 << 17, block_step_cap=1 << 16), "block_step"),                  # step beyond the capacity's
    (dict(n_cap=1 << 18, block_step=1 << 14, block_step_cap=1 << 15), "cover n_cap"),                 # capacity step too small
    (dict(block_step_cap=1 << 14), "without n_cap"),
])
def test_capacity_layout_arguments_are_checked(kw, needle):
    """AarLaunch.n_cap (one region for several message lengths): inconsistent descriptions are refused up front."""
    lib = K.lib()
    a = _launch(**kw)
    assert lib.aar_round_kernel(C.byref(a)) in (-1, -4)
    assert needle in lib.aar_last_error().decode()


def test_slot_views_are_only_handed_out_with_full_thresholds():
    """Both views bypass the kernel's back-pressure (the caller touches the slot outside the launch).  Round 1 handed
    `input_view` out with thComplete < 1: a rank could finish round r - L without a slow owner's chunks and overwrite the
    staging rows that owner was still pulling.  The rule needs no GPU to check."""
    import types

    import torch

    from akka_allreduce_b200.parallel.engine import DeviceAllreduce

    eng = object.__new__(DeviceAllreduce)
    eng.rank, eng.world, eng.algo = 0, 4, "nvls"
    eng.geo = types.SimpleNamespace(n=4096, world=4, total_chunks=8, step=1024, slots=2)
    eng.launcher = types.SimpleNamespace(dtype=torch.float32, wire_dtype=torch.float32, scale=1.0, req_reduce=4,
                                         req_complete=6, off_scat=0, off_sink=0)
    eng._view = lambda off, numel: ("view", off, numel)
    assert eng.zero_copy_ok and eng.partial_thresholds
    assert eng.input_view(0) is None and eng.output_view(0) is None          # thComplete 0.75: neither
    eng.launcher.req_complete = 8
    assert eng.input_view(1) == ("view", 1 * 4 * 1024 * 4, 4096) and eng.output_view(0) == ("view", 0, 4096)
    eng.launcher.req_reduce = 3                                               # thReduce below 1: neither, either
    assert eng.input_view(0) is None and eng.output_view(0) is None


def test_engine_round_variants_fill_the_launch_description():
    """What each public entry of DeviceAllreduce puts into the launch description (no GPU: the launcher is a stub):
    whole rounds, the reduce-scatter / all-gather halves (op restored afterwards), and a shorter round on the
    capacity layout (n / block_step of the round, n_cap / block_step_cap of the region, req_complete of the shorter
    message; restored afterwards)."""
    import torch

    from akka_allreduce_b200.parallel.engine import DeviceAllreduce

    eng = object.__new__(DeviceAllreduce)
    eng.rank, eng.world, eng.algo, eng.round = 0, 4, "twoshot", 0
    eng.geo = Geometry.make(4096, 512, 4, 1, True)
    seen = []

    class Stub:
        dtype = wire_dtype = torch.float32
        scale, req_reduce, req_complete, th_complete, stream_expected = 1.0, 4, 8, 1.0, True

        def fill(self, a, r, inj):
            a.req_complete = self.req_complete

        def launch(self, a, s):
            seen.append((a.n, a.block_step, a.n_cap, a.block_step_cap, a.req_complete, a.op))

    eng.launcher, eng._launch, eng._inplace_src = Stub(), K.AarLaunch(), None
    eng._launch.n, eng._launch.block_step = 4096, eng.geo.block_step_arg
    x, y = torch.zeros(4096), torch.zeros(4096)
    eng.allreduce(x, y)
    eng.allreduce(x, x)                                        # in place: fine on two-shot with thresholds 1
    eng.allreduce_part(x[:1000], y[:1000])
    eng.allreduce_part(x, y)                                   # full length = an ordinary round
    lo, hi = eng.block_range()
    eng.reduce_scatter(x, y[: hi - lo])
    eng.all_gather(x[: hi - lo], y)
    assert seen == [(4096, 1024, 0, 0, 8, K.OP_ALLREDUCE)] * 2 + [(1000, 256, 4096, 1024, 4, K.OP_ALLREDUCE),
                                                                   (4096, 1024, 0, 0, 8, K.OP_ALLREDUCE),
                                                                   (4096, 1024, 0, 0, 8, K.OP_REDUCE_SCATTER),
                                                                   (4096, 1024, 0, 0, 8, K.OP_ALL_GATHER)]
    a = eng._launch
    assert (a.n, a.block_step, a.n_cap, a.block_step_cap, a.op) == (4096, 1024, 0, 0, K.OP_ALLREDUCE) and eng.round == 6
    with pytest.raises(ValueError):
        eng.allreduce_part(x[:1000], y[:999])
    with pytest.raises(ValueError):
        eng.reduce_scatter(x, y)                                # out must be block-sized
