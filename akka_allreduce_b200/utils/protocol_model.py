"""Exhaustive interleaving check of the device flag protocol (race detection without a sanitizer).

The reference needs no race detector: an actor handles one message at a time (SURVEY.md 5.2,
AllreduceWorker.scala:33).  The device path does have real cross-GPU races to get right — flags versus data,
ring-slot recycling, ranks more than a round apart — and `compute-sanitizer` cannot see across GPUs (nor is it
available on the pool this was developed on).  So the PROTOCOL is checked here as a small explicit-state
model: every rank runs the step sequence of `csrc/cuda/fused_allreduce.cu` (two-shot) or of the one-shot
kernel, memory is the symmetric region's flag and data words, and a breadth-first search visits every
interleaving of the ranks' steps for a few ranks, slots and rounds.

What a step is.  Stores to different peers are separate steps where a peer can observe the difference (flags,
`done` words); the data stores of a phase are one step (no guard ever looks at data, and one location's hazard
is independent of the others).  A wait is split in two: the step that OBSERVES the flags and fixes the
contributor set, and the later step that READS the data — the window a recycled slot could be overwritten in.
Data words carry the tag of the round that wrote them; a read of a word whose tag is not the reader's is the
violation the search looks for.  With `timeouts=True` every wait may also give up with whatever it has (the
kernel's timeout / force flag / a dead peer), which is what lets ranks drift more than one round apart.

Checked properties:
  * safety   — no reader ever consumes a word written for another round (slot recycled too early, flag
               overtaking data at protocol level, stale flag taken for a fresh one);
  * liveness — without timeouts no reachable state is stuck before every rank finished every round;
  * exactness — with all thresholds 1 and no timeouts every reduce and every completion sees all P ranks.

`tests/test_protocol_model.py` runs the configurations the engine actually uses and two seeded bugs (no
back-pressure; one-shot without gate and with a one-slot ring) that the search must find.
"""
from __future__ import annotations

from collections import deque
from dataclasses import dataclass

# program counters
GATE, MASK, SDATA, SFLAG, RWAIT, RREAD, BDATA, BFLAG, CWAIT, CREAD, DONE, END = range(12)
_NAMES = ["gate", "dest-mask", "scatter-data", "scatter-flag", "reduce-wait", "reduce-read", "bcast-data",
          "bcast-flag", "complete-wait", "complete-read", "done-flag", "end"]


@dataclass(frozen=True)
class ModelConfig:
    P: int = 2                 # ranks
    L: int = 2                 # ring slots (maxLag + 1)
    rounds: int = 3
    req_reduce: int = 0        # 0 = P
    req_complete: int = 0      # 0 = P (one reduced chunk per owner in the model)
    gate_count: int = 0        # thAllreduce gate: peers that must have finished round r-1; 0 = no gate
    timeouts: bool = False     # any wait may give up with what it has
    giveups: int = -1          # with timeouts: at most this many give-ups in one execution (-1 = unbounded);
                               # bounds the search for 3+ ranks, where the unbounded one does not fit in memory
    oneshot: bool = False      # the one-shot kernel's step sequence instead of the two-shot's
    backpressure: bool = True  # False = seeded bug: push into a peer's slot without looking at its `done`
    early_write: bool = False  # zero-copy input (`input_view`): the caller fills the round's staging rows BEFORE the
                               # launch, i.e. before the gate and without the destination mask
    output_view: bool = False  # zero-copy output (`output_view`): the caller keeps reading round r's sink slot until it
                               # launches round r+L (the worker loop delivers sink(r) before that launch)
    notices: bool = False      # skip notices (csrc/cuda/round_common.cuh decide_dest_mask_warp / peers_gone, csrc/host/shm_cluster.cc): a sender
                               # that leaves a peer out records the tag in a per-(dest, src) word outside the slot ring,
                               # and a waiter treats "not arrived + notice >= my tag" as "will never arrive"
    dead_rank: int = -1        # >= 0: this rank stops for good after finishing round-tag `dies_after` (its process is gone);
    dies_after: int = 1        # at some later, arbitrary moment the coordinator's removal reaches the survivors as the
                               # dead-peer mask (one global step): from then on they neither push to it nor wait for it
                               # (round_common.cuh peers_gone / decide_dest_mask_warp; AllreduceMaster.scala:46-52)
    announce: bool = True      # False = seeded bug: the removal never reaches the survivors (no dead-peer mask)
    oneshot_backpressure: bool = False  # True = the shipped one-shot kernel (destination mask in its prologue); False = the round-1 kernel without it;
                                        # True models the planned fix (same `done + L >= tag` rule as two-shot)


@dataclass
class Result:
    ok: bool
    states: int
    violation: str | None = None
    trace: list | None = None


class _Layout:
    """Flat indices into the memory tuple: done[q][s], then arr/scat/rfl/sink[q][l][s]."""

    def __init__(self, P: int, L: int):
        self.P, self.L = P, L
        self.o_done = 0
        self.o_arr = P * P
        self.o_scat = self.o_arr + P * L * P
        self.o_rfl = self.o_scat + P * L * P
        self.o_sink = self.o_rfl + P * L * P
        self.o_note = self.o_sink + P * L * P
        self.size = self.o_note + P * P

    def done(self, q, s):
        return self.o_done + q * self.P + s

    def arr(self, q, l, s):
        return self.o_arr + (q * self.L + l) * self.P + s

    def scat(self, q, l, s):
        return self.o_scat + (q * self.L + l) * self.P + s

    def rfl(self, q, l, s):
        return self.o_rfl + (q * self.L + l) * self.P + s

    def sink(self, q, l, s):
        return self.o_sink + (q * self.L + l) * self.P + s

    def note(self, q, s):
        return self.o_note + q * self.P + s


def _put(mem: tuple, writes) -> tuple:
    m = list(mem)
    for k, v in writes:
        m[k] = v
    return tuple(m)


def check(cfg: ModelConfig, max_states: int = 3_000_000, walks: int = 0, seed: int = 0,
          giveup_weight: float = 0.02) -> Result:
    """Breadth-first search over every interleaving (`walks` = 0), or `walks` random executions for
    configurations too large to enumerate (8 ranks, tens of rounds); a give-up is then taken with relative
    weight `giveup_weight` so that runs are mostly normal with the occasional timeout."""
    P, L, T = cfg.P, cfg.L, cfg.rounds
    req_r = cfg.req_reduce or P
    req_c = cfg.req_complete or P
    lay = _Layout(P, L)
    full = (1 << P) - 1
    exact = req_r == P and req_c == P and not cfg.timeouts and not cfg.notices

    # rank state: (tag, pc, i, ok_mask, mask)
    init = (tuple((1, GATE, 0, 0, 0) for _ in range(P)), tuple([0] * lay.size), cfg.giveups if cfg.timeouts else 0, 0)
    parent = {init: None}
    queue = deque([init])

    def trace_of(state, last):
        out = [last]
        while parent[state] is not None:
            state, what = parent[state]
            out.append(what)
        return list(reversed(out))

    def successors(state):
        """Yields (label, next_state) or (label, None, violation)."""
        ranks, mem, left, ann = state
        may_give_up = cfg.timeouts and left != 0
        dead = cfg.dead_rank if ann else -1     # what the survivors KNOW (the mask), not what is true
        if cfg.dead_rank >= 0 and cfg.announce and not ann and ranks[cfg.dead_rank][1] == END:
            yield ("coordinator: removal announced (dead-peer mask set)", (ranks, mem, left, 1))
        for q in range(P):
            t, pc, i, ok, m = ranks[q]
            if pc == END:
                continue
            l = (t - 1) % L
            label = f"r{q} t{t} {_NAMES[pc]}" + (f"[{(q + i) % P}]" if pc in (SFLAG, BFLAG, DONE) else "")

            def go(new_rank, writes=(), label=label, gave_up=False):
                nr = ranks[:q] + (new_rank,) + ranks[q + 1:]
                if gave_up:
                    label += " (gave up)"
                return (label, (nr, _put(mem, writes) if writes else mem, left - 1 if gave_up and left > 0 else left, ann))

            if pc == GATE and cfg.output_view and i == 0 and t > L and not cfg.oneshot:
                bad = [o for o in range(P) if mem[lay.sink(q, l, o)] > t - L]
                if bad:
                    yield (label, None, f"rank {q}: output view of round-tag {t - L} (slot {l}) was overwritten by owner "
                                        f"{bad[0]} for tag {mem[lay.sink(q, l, bad[0])]} before the caller gave the slot back")
                    return
            if pc == GATE and cfg.early_write and i == 0:
                # the source callback has produced round t's message straight into the staging slot
                yield go((t, GATE, 1, 0, 0), [(lay.scat(d, l, q), t) for d in range(P)], label=label + " (caller fills staging)")
            elif pc == GATE:
                have = sum(1 for s in range(P) if s == q or mem[lay.done(q, s)] + 1 >= t)
                need = min(cfg.gate_count, P - 1) if dead >= 0 else cfg.gate_count   # no more than the live ranks can give
                opened = cfg.gate_count <= 0 or t <= 1 or have >= need
                if opened or may_give_up:
                    yield go((t, MASK, 0, 0, 0), gave_up=not opened)
            elif pc == MASK:
                if (cfg.oneshot and not cfg.oneshot_backpressure) or not cfg.backpressure:
                    okm = full   # the one-shot kernel has no destination mask
                else:
                    okm = 0
                    for d in range(P):
                        if d == q or (mem[lay.done(q, d)] + L >= t and d != dead):
                            okm |= 1 << d
                notes = [(lay.note(d, q), t) for d in range(P) if not (okm >> d) & 1 and d != dead] if cfg.notices else []
                yield go((t, SDATA, 0, okm, 0), notes)
            elif pc == SDATA:
                # two-shot: my slice for owner d into d's scat row [me]; one-shot: my whole message, same rows
                writes = [] if cfg.early_write else [(lay.scat(d, l, q), t) for d in range(P) if (ok >> d) & 1]
                yield go((t, SFLAG, 0, ok, 0), writes)
            elif pc == SFLAG:
                d = (q + i) % P
                writes = [(lay.arr(d, l, q), t)] if (ok >> d) & 1 else []
                nxt = (t, SFLAG, i + 1, ok, 0) if i + 1 < P else (t, RWAIT, 0, ok, 0)
                yield go(nxt, writes)
            elif pc == RWAIT:
                have = 0
                for s in range(P):
                    if mem[lay.arr(q, l, s)] == t:
                        have |= 1 << s
                met = bin(have).count("1") >= req_r
                if not met and (cfg.notices or dead >= 0):   # everyone who has not arrived is dead or has said it never will
                    gone = [s for s in range(P) if not (have >> s) & 1 and (s == dead or (cfg.notices and mem[lay.note(q, s)] >= t))]
                    met = bool(gone) and bin(have).count("1") + len(gone) == P
                if met or may_give_up:
                    yield go((t, RREAD, 0, ok, have), gave_up=not met)
            elif pc == RREAD:
                for s in range(P):
                    if (m >> s) & 1 and mem[lay.scat(q, l, s)] != t:
                        yield (label, None, f"rank {q} round-tag {t}: contribution of rank {s} in slot {l} carries "
                                            f"tag {mem[lay.scat(q, l, s)]}")
                        return
                if exact and m != full and cfg.dead_rank < 0:
                    yield (label, None, f"rank {q} round-tag {t}: reduced {m:b} instead of all ranks")
                    return
                if cfg.oneshot:   # one-shot: the local sum IS the result; nothing to broadcast or wait for
                    yield go((t, DONE, 0, ok, m))
                else:
                    yield go((t, BDATA, 0, ok, m))
            elif pc == BDATA:
                # a skipped peer's copy is redirected to my own sink row (it is never flagged there)
                writes = [(lay.sink(d if (ok >> d) & 1 else q, l, q), t) for d in range(P)]
                yield go((t, BFLAG, 0, ok, m), writes)
            elif pc == BFLAG:
                d = (q + i) % P
                writes = [(lay.rfl(d, l, q), t)] if (ok >> d) & 1 else []
                nxt = (t, BFLAG, i + 1, ok, m) if i + 1 < P else (t, CWAIT, 0, ok, 0)
                yield go(nxt, writes)
            elif pc == CWAIT:
                have = 0
                for o in range(P):
                    if mem[lay.rfl(q, l, o)] == t:
                        have |= 1 << o
                met = bin(have).count("1") >= req_c
                if not met and (cfg.notices or dead >= 0):
                    gone = [o for o in range(P) if not (have >> o) & 1 and (o == dead or (cfg.notices and mem[lay.note(q, o)] >= t))]
                    met = bool(gone) and bin(have).count("1") + len(gone) == P
                if met or may_give_up:
                    yield go((t, CREAD, 0, ok, have), gave_up=not met)
            elif pc == CREAD:
                for o in range(P):
                    if (m >> o) & 1 and mem[lay.sink(q, l, o)] != t:
                        yield (label, None, f"rank {q} round-tag {t}: reduced block of owner {o} in slot {l} "
                                            f"carries tag {mem[lay.sink(q, l, o)]}")
                        return
                if exact and m != full and cfg.dead_rank < 0:
                    yield (label, None, f"rank {q} round-tag {t}: completed with {m:b} instead of all owners")
                    return
                yield go((t, DONE, 0, ok, 0))
            elif pc == DONE:
                d = (q + i) % P
                writes = [(lay.done(d, q), t)] if d != dead else []
                if i + 1 < P:
                    yield go((t, DONE, i + 1, ok, 0), writes)
                elif q == cfg.dead_rank and t >= cfg.dies_after:
                    yield go((t, END, 0, 0, 0), writes, label=label + " (and dies)")
                elif t < T:
                    yield go((t + 1, GATE, 0, 0, 0), writes)
                else:
                    yield go((t, END, 0, 0, 0), writes)

    if walks > 0:
        import random

        rng = random.Random(seed)
        steps = 0
        for _ in range(walks):
            state, path = init, []
            # uniform scheduling keeps the ranks in lock step; the bugs live in skew, so every execution draws a
            # speed per rank (log-uniform over three decades) and redraws it now and then
            speed = [10 ** rng.uniform(-3, 0) for _ in range(P)]
            while True:
                nxt = list(successors(state))
                if not nxt:
                    if any(r[1] != END for r in state[0]) and not (cfg.timeouts and state[2] == 0):
                        where = ", ".join(f"r{q}: tag {r[0]} at {_NAMES[r[1]]}" for q, r in enumerate(state[0]))
                        return Result(False, steps, "stuck (no rank can move): " + where, path + ["stuck"])
                    break
                bad = [x for x in nxt if x[1] is None]
                if bad:
                    return Result(False, steps, bad[0][2], path + [bad[0][0]])
                if rng.random() < 0.01:
                    speed = [10 ** rng.uniform(-3, 0) for _ in range(P)]
                weights = [(giveup_weight if x[0].endswith("(gave up)") else 1.0) * speed[int(x[0][1:x[0].index(" ")])]
                           for x in nxt]
                label, state = rng.choices(nxt, weights)[0]
                path.append(label)
                steps += 1
        return Result(True, steps)

    states = 0
    while queue:
        state = queue.popleft()
        states += 1
        if states > max_states:
            return Result(False, states, f"state limit {max_states} reached (search incomplete)")
        moved = False
        for nxt in successors(state):
            moved = True
            if nxt[1] is None:
                return Result(False, states, nxt[2], trace_of(state, nxt[0]))
            label, ns = nxt
            if ns not in parent:
                parent[ns] = (state, label)
                queue.append(ns)
        if not moved and any(r[1] != END for r in state[0]) and not (cfg.timeouts and state[2] == 0):
            where = ", ".join(f"r{q}: tag {r[0]} at {_NAMES[r[1]]}" for q, r in enumerate(state[0]))
            return Result(False, states, "stuck (no rank can move): " + where, trace_of(state, "stuck"))
    return Result(True, states)


if __name__ == "__main__":
    import argparse
    import time

    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("-P", type=int, default=2)
    ap.add_argument("-L", type=int, default=2)
    ap.add_argument("--rounds", type=int, default=3)
    ap.add_argument("--req-reduce", type=int, default=0)
    ap.add_argument("--req-complete", type=int, default=0)
    ap.add_argument("--gate", type=int, default=0)
    ap.add_argument("--timeouts", action="store_true")
    ap.add_argument("--giveups", type=int, default=-1)
    ap.add_argument("--oneshot", action="store_true")
    ap.add_argument("--no-backpressure", action="store_true")
    ap.add_argument("--oneshot-backpressure", action="store_true")
    ap.add_argument("--early-write", action="store_true")
    ap.add_argument("--output-view", action="store_true")
    ap.add_argument("--notices", action="store_true")
    ap.add_argument("--dead-rank", type=int, default=-1, help="this rank stops for good after round-tag --dies-after")
    ap.add_argument("--dies-after", type=int, default=1)
    ap.add_argument("--walks", type=int, default=0, help="random executions instead of the exhaustive search")
    ap.add_argument("--seed", type=int, default=0)
    a = ap.parse_args()
    cfg = ModelConfig(a.P, a.L, a.rounds, a.req_reduce, a.req_complete, a.gate, a.timeouts, a.giveups, a.oneshot,
                      not a.no_backpressure, a.early_write, a.output_view, a.notices, a.dead_rank, a.dies_after, True,
                      a.oneshot_backpressure)
    t0 = time.time()
    r = check(cfg, walks=a.walks, seed=a.seed)
    print(f"{cfg}\n  {'OK' if r.ok else 'VIOLATION'}: {r.states} states in {time.time() - t0:.1f}s")
    try:
        if not r.ok:
            print("  " + r.violation)
            for step in (r.trace or [])[-40:]:
                print("    " + step)
    except BrokenPipeError:  # `... | head`
        pass
