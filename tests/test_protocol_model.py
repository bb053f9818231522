"""Race detection for the device protocol: every interleaving of a small model of the flag
protocol is searched for a reader consuming a recycled slot, a stuck state, or an incomplete sum.  The model
and what one step stands for: akka_allreduce_b200/utils/protocol_model.py."""
import pytest

from akka_allreduce_b200.utils.protocol_model import ModelConfig, check


@pytest.mark.parametrize("cfg", [
    ModelConfig(P=2, L=2, rounds=4),                       # flagship: thresholds 1, maxLag 1, gate elided (implied)
    ModelConfig(P=3, L=2, rounds=3),
    ModelConfig(P=2, L=1, rounds=4, gate_count=2),         # maxLag 0: one slot, the gate is what keeps it safe
    ModelConfig(P=3, L=1, rounds=2, gate_count=3),
    ModelConfig(P=2, L=3, rounds=5),
], ids=str)
def test_full_thresholds_are_safe_live_and_exact(cfg):
    r = check(cfg)
    assert r.ok, f"{r.violation}\n" + "\n".join(r.trace or [])


def test_one_slot_without_gate_gets_stuck():
    """Why RoundLauncher.gate_implied asks for maxLag >= 1: with a single slot and no gate a fast rank finds the
    slot of a slower peer still occupied, skips it (back-pressure), and the peer then waits for data that is
    never sent.  Safe, but not live without timeouts."""
    r = check(ModelConfig(P=2, L=1, rounds=3))
    assert not r.ok and r.violation.startswith("stuck")


@pytest.mark.parametrize("cfg", [
    ModelConfig(P=2, L=2, rounds=5, timeouts=True),                                            # timeouts only
    ModelConfig(P=2, L=2, rounds=5, req_reduce=1, req_complete=1, gate_count=1, timeouts=True),  # thresholds 0.5
    ModelConfig(P=2, L=3, rounds=5, req_reduce=1, req_complete=1, timeouts=True),                # no gate at all
    ModelConfig(P=2, L=1, rounds=4, req_reduce=1, req_complete=1, gate_count=1, timeouts=True),
    ModelConfig(P=3, L=2, rounds=2, req_reduce=2, req_complete=2, gate_count=2, timeouts=True, giveups=2),
], ids=str)
def test_partial_thresholds_and_timeouts_never_read_a_recycled_slot(cfg):
    """Ranks may now be any number of rounds apart (a wait can give up at any point, which also covers a peer
    that died): back-pressure on `done` is what keeps a reader's slot from being overwritten."""
    r = check(cfg)
    assert r.ok, f"{r.violation}\n" + "\n".join(r.trace or [])


def test_search_finds_the_seeded_bug():
    """Same configuration with the back-pressure test removed: the search must come back with the overwrite."""
    r = check(ModelConfig(P=2, L=2, rounds=4, req_reduce=1, req_complete=1, gate_count=1, timeouts=True,
                          backpressure=False))
    assert not r.ok and "carries tag" in r.violation and r.trace


def test_oneshot_rules():
    assert check(ModelConfig(P=2, L=2, rounds=4, oneshot=True)).ok
    assert check(ModelConfig(P=3, L=2, rounds=3, oneshot=True)).ok
    assert check(ModelConfig(P=2, L=1, rounds=4, oneshot=True, gate_count=2)).ok
    # one slot and no gate: a peer overwrites a copy that is still being summed; aar_allreduce_round()
    # refuses this combination (AAR_EINVAL "one-shot needs maxLag >= 1 or the thAllreduce gate")
    r = check(ModelConfig(P=2, L=1, rounds=3, oneshot=True))
    assert not r.ok and "carries tag" in r.violation


def test_oneshot_known_gap_under_timeouts_and_its_fix():
    """Found by this search, documented in DESIGN.md 7: the one-shot kernel has no destination mask, so a rank
    that gives up two rounds in a row (>= 2 timeouts, 4 s by default) can overwrite a copy a stalled peer is
    still summing.  Modelled with the two-shot kernel's `done + L >= tag` rule the violation is gone."""
    r = check(ModelConfig(P=2, L=2, rounds=4, oneshot=True, timeouts=True))
    assert not r.ok and "carries tag" in r.violation
    assert sum("gave up" in s for s in r.trace) >= 2
    assert check(ModelConfig(P=2, L=2, rounds=5, oneshot=True, timeouts=True, oneshot_backpressure=True)).ok


def test_zero_copy_input_preconditions():
    """`DeviceAllreduce.input_view` lets the caller write a round's message straight into the staging slot, before
    the launch and therefore before the gate and the destination mask.  It is only handed out in NVLS mode, which
    needs thReduce = 1: then nobody can reduce round r+1 before every rank has finished reading round r, whatever
    thComplete is.  Below thReduce = 1, or once waits may time out, the early write can hit a slot a slow peer is
    still reading (same 2-timeouts corner as the one-shot gap)."""
    assert check(ModelConfig(P=2, L=2, rounds=4, early_write=True)).ok
    assert check(ModelConfig(P=3, L=2, rounds=3, early_write=True, req_complete=2)).ok
    assert check(ModelConfig(P=2, L=1, rounds=4, early_write=True, gate_count=2)).ok
    assert not check(ModelConfig(P=2, L=3, rounds=5, early_write=True, req_reduce=1, req_complete=1)).ok
    assert not check(ModelConfig(P=2, L=2, rounds=4, early_write=True, timeouts=True)).ok


def test_zero_copy_output_preconditions():
    """`DeviceAllreduce.output_view`: the caller reads round r from the sink slot until it launches round r+L.
    Safe with thReduce = 1 (owners cannot reduce round r+L before this rank has scattered it); with thReduce < 1
    they can, and overwrite the view — which is why the engine only hands the view out when both thresholds are 1
    (this search found the missing thReduce condition)."""
    assert check(ModelConfig(P=2, L=2, rounds=5, output_view=True)).ok
    assert check(ModelConfig(P=3, L=2, rounds=4, output_view=True)).ok
    assert check(ModelConfig(P=2, L=1, rounds=4, output_view=True, gate_count=2)).ok
    r = check(ModelConfig(P=3, L=2, rounds=4, output_view=True, req_reduce=2))
    assert not r.ok and "output view" in r.violation


def test_random_executions_at_eight_ranks():
    """Configurations too large to enumerate: random executions with a random speed per rank (the bugs live in
    skew).  The 8-rank production shapes pass; the seeded bug is still found."""
    assert check(ModelConfig(P=8, L=2, rounds=8), walks=20, seed=1).ok
    assert check(ModelConfig(P=8, L=3, rounds=8, req_reduce=6, req_complete=6, gate_count=6, timeouts=True),
                 walks=40, seed=2).ok
    r = check(ModelConfig(P=6, L=2, rounds=10, req_reduce=4, req_complete=4, gate_count=4, timeouts=True,
                          backpressure=False), walks=300, seed=3)
    assert not r.ok and "carries tag" in r.violation


def test_skip_notices_make_partial_thresholds_live_without_timeouts():
    """Without notices a rank that its peers left out (back-pressure) waits for data that was never sent: with
    thresholds below 1 and no timeouts the search finds stuck states.  With notices ("highest tag for which I left
    you out", one word per pair, outside the slot ring) every configuration below is live and still safe — even one
    slot without the gate.  Implemented in csrc/host/shm_cluster.cc; a candidate for the kernel."""
    stuck = check(ModelConfig(P=3, L=2, rounds=3, req_reduce=2, req_complete=2, gate_count=2))
    assert not stuck.ok and stuck.violation.startswith("stuck")
    assert check(ModelConfig(P=3, L=2, rounds=3, req_reduce=2, req_complete=2, gate_count=2, notices=True)).ok
    assert check(ModelConfig(P=2, L=1, rounds=4, notices=True)).ok
    assert check(ModelConfig(P=2, L=2, rounds=5, req_reduce=1, req_complete=1, notices=True, timeouts=True)).ok
    assert check(ModelConfig(P=8, L=2, rounds=8, req_reduce=6, req_complete=6, gate_count=6, notices=True),
                 walks=40, seed=4).ok


def test_dead_peer_mask_keeps_the_survivors_safe_and_live():
    """A rank stops for good; at an arbitrary later moment the coordinator's removal reaches the survivors as the
    dead-peer mask.  With every threshold at 1 and NO timeouts the survivors still finish every round (the waits close
    on "nothing more can come") and never read a recycled slot; without the announcement they are stuck — the check
    has teeth."""
    from akka_allreduce_b200.utils.protocol_model import ModelConfig, check

    for cfg in (ModelConfig(P=3, L=2, rounds=4, dead_rank=2, dies_after=1),
                ModelConfig(P=3, L=2, rounds=4, gate_count=3, dead_rank=2, dies_after=2),
                ModelConfig(P=3, L=1, rounds=3, gate_count=3, notices=True, dead_rank=0, dies_after=1),
                ModelConfig(P=3, L=2, rounds=3, req_reduce=2, req_complete=2, gate_count=2, notices=True, dead_rank=1, dies_after=1),
                ModelConfig(P=3, L=2, rounds=3, oneshot=True, oneshot_backpressure=True, dead_rank=0, dies_after=1)):
        r = check(cfg)
        assert r.ok, (cfg, r.violation)
    r = check(ModelConfig(P=3, L=2, rounds=4, dead_rank=2, dies_after=1, announce=False))
    assert not r.ok and r.violation.startswith("stuck")
