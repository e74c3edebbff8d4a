"""A model of K1b's tile ring (csrc/dm_tile_kernel.cuh) under adversarial scheduling: G groups of a CTA walk the CTA's tiles
round-robin (tile j -> group j % G, slot j % S); each slot has one mbarrier whose phase completes when the slot's pending load
has landed; a consumer waits with mbarrier.try_wait.parity semantics -- the parity of the CURRENT phase means "not yet", the
parity of the immediately PRECEDING phase means "complete".  With S == G a slot is always consumed by the same group and the
one-bit wait is safe.  With S > G (the middle launch shape: 3 groups, 4 slots) consecutive tenants of a slot belong to
different groups, and a consumer that arrives before the PREVIOUS tenant's load has completed is let through at once: the race
that produced `unspecified launch failure` after thousands of tiles.  The model finds that interleaving without the fix and
never with it (the consumer first waits until its tile's load has been issued: s_issued).  CPU only."""
import random

import pytest


class Violation(Exception):
    pass


def run(G, S, n_tiles, seed, fix):
    rng = random.Random(seed)
    # per slot: completed phases (the barrier's phase counter), load in flight (tile or None), content (tile landed), issued
    phase = [0] * S
    in_flight = [None] * S
    content = [None] * S
    issued = [-1] * S
    prog = [g for g in range(G)]           # next tile of each group
    state = ["wait"] * G                    # wait -> compute -> release
    events = []

    def issue(j):
        s = j % S
        if in_flight[s] is not None:
            raise Violation("load of tile %d issued into slot %d while tile %d is still in flight" % (j, s, in_flight[s]))
        in_flight[s] = j
        issued[s] = j

    for j in range(min(S, n_tiles)):        # the ring is filled at the start
        issue(j)
    steps = 0
    while any(p < n_tiles for p in prog):
        steps += 1
        assert steps < 200000, "deadlock"
        choices = [("land", s) for s in range(S) if in_flight[s] is not None] + [("grp", g) for g in range(G) if prog[g] < n_tiles]
        kind, i = rng.choice(choices)
        if kind == "land":
            content[i] = in_flight[i]
            in_flight[i] = None
            phase[i] += 1
            continue
        g, j = i, prog[i]
        s = j % S
        if state[g] == "wait":
            if fix and S != G and issued[s] < j:
                continue                     # spinning on s_issued
            want_parity = (j // S) & 1       # this tile is the (j // S)-th tenant of its slot
            if (phase[s] & 1) == want_parity:
                continue                     # the current (incomplete) phase has the wanted parity: not yet
            if content[s] != j:              # let through on the parity of the PRECEDING phase
                raise Violation("group %d starts tile %d on slot %d holding tile %s" % (g, j, s, content[s]))
            state[g] = "compute"
        elif state[g] == "compute":
            state[g] = "release"
        else:
            if j + S < n_tiles:
                issue(j + S)                 # the slot's consumer refills it when it is done with it
            prog[g] = j + G
            state[g] = "wait"
    return True


@pytest.mark.parametrize("G,S", [(4, 4), (3, 3), (2, 2)])
def test_one_slot_per_group_is_safe_without_the_flag(G, S):
    for seed in range(300):
        run(G, S, 60, seed, fix=False)


def test_more_slots_than_groups_needs_the_flag():
    found = 0
    for seed in range(400):
        try:
            run(3, 4, 60, seed, fix=False)
        except Violation:
            found += 1
    assert found > 0, "the model should reach the stale-tenant interleaving"
    for G, S in [(3, 4), (2, 3), (3, 6), (4, 6)]:
        for seed in range(400):
            run(G, S, 60, seed, fix=True)
