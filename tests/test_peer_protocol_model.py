"""A model of the statistics collective's protocol (csrc/dm_peer.cuh: tagged lines, two line sets by parity of the
sequence number, lagged and complete exchanges) under adversarial scheduling: every remote store is delivered at an
arbitrary later time (any order), ranks run at arbitrary relative speeds, lagged and complete exchanges are mixed.
Checked: nobody ever reads a value of the wrong exchange, no store overwrites a line its receiver has not consumed yet,
no deadlock, and after a final complete exchange every rank's totals equal the sum of everything every rank accumulated.
(CPU-only; the device path is tests/test_gpu_multirank.py.)"""
import random

import pytest


class Rank:
    def __init__(self, r, world):
        self.r, self.world = r, world
        self.seq, self.pending = 0, False
        self.acc = 0            # running accumulator (the step kernels add into it)
        self.total = 0          # global totals
        self.line = [[(0, 0)] * world for _ in range(2)]  # [parity][source rank] = (tag, value)
        self.consumed = [[0] * world for _ in range(2)]   # tag of the last line consumed from that slot
        self.ops = []           # remaining program: "step", "lagged", "complete"
        self.phase = None       # progress inside the current exchange


def run(world, program_len, seed):
    rng = random.Random(seed)
    ranks = [Rank(r, world) for r in range(world)]
    # every rank runs the same sequence of exchanges (the API's contract), with its own steps in between
    kinds = [rng.choice(["lagged", "lagged", "lagged", "complete"]) for _ in range(program_len)] + ["complete"]
    contributed = 0
    for rk in ranks:
        for k in kinds:
            rk.ops += ["step"] * rng.randint(0, 2) + [k]
    in_flight = []  # (dst, parity, src, tag, value): remote stores not yet delivered

    def recv_ready(rk, seq):
        return world == 1 or all(rk.line[seq & 1][s][0] == seq for s in range(world))

    def recv(rk, seq):
        if world == 1:  # one rank: the snapshot went straight to the totals (ll_send)
            return
        for s in range(world):
            tag, v = rk.line[seq & 1][s]
            assert tag == seq, "rank %d read exchange %d from rank %d while waiting for %d" % (rk.r, tag, s, seq)
            rk.total += v
            rk.consumed[seq & 1][s] = seq

    def send(rk, seq):
        snap, rk.acc = rk.acc, 0
        if world == 1:
            rk.total += snap
            return
        for d in range(world):
            in_flight.append((d, seq & 1, rk.r, seq, snap))

    def deliver(i):
        d, par, s, tag, v = in_flight.pop(i)
        old_tag = ranks[d].line[par][s][0]
        assert ranks[d].consumed[par][s] == old_tag, \
            "store of exchange %d from rank %d overwrites exchange %d that rank %d has not consumed" % (tag, s, old_tag, d)
        assert tag > old_tag
        ranks[d].line[par][s] = (tag, v)

    def advance(rk):
        """One scheduling quantum of rank rk; returns False when it cannot move (spinning on its lines)."""
        nonlocal contributed
        if not rk.ops:
            return False
        op = rk.ops[0]
        if op == "step":
            x = rng.randint(1, 9)
            rk.acc += x
            contributed += x
            rk.ops.pop(0)
            return True
        if rk.phase is None:
            rk.phase = "recv_old" if rk.pending else "send"
        if rk.phase == "recv_old":
            if not recv_ready(rk, rk.seq):
                return False
            recv(rk, rk.seq)
            rk.phase = "send"
            return True
        if rk.phase == "send":
            send(rk, rk.seq + 1)
            rk.phase = "recv_new" if op == "complete" else "done"
            return True
        if rk.phase == "recv_new":
            if not recv_ready(rk, rk.seq + 1):
                return False
            recv(rk, rk.seq + 1)
            rk.phase = "done"
            return True
        rk.seq += 1
        rk.pending = op == "lagged" and world > 1
        rk.phase = None
        rk.ops.pop(0)
        return True

    idle = 0
    while any(rk.ops for rk in ranks) or in_flight:
        moved = False
        if in_flight and rng.random() < 0.4:
            deliver(rng.randrange(len(in_flight)))  # any order, any delay
            moved = True
        else:
            moved = advance(rng.choice(ranks))
        idle = 0 if moved else idle + 1
        if idle > 10000:  # nobody can move and nothing is in flight: a deadlock
            assert in_flight or any(advance(rk) for rk in ranks), "deadlock"
            if in_flight:
                deliver(rng.randrange(len(in_flight)))
            idle = 0
    for rk in ranks:
        assert not rk.pending and rk.acc == 0
        assert rk.total == contributed, (rk.r, rk.total, contributed)


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_protocol_model(world):
    for seed in range(60):
        run(world, program_len=12, seed=1000 * world + seed)
