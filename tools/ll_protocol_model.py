#!/usr/bin/env python
"""Randomised-schedule model of the LL (flag-in-data) exchange protocols of
``csrc/mgb_common.cuh`` / ``csrc/allreduce.cu``.

Every rank owns slots ``[parity][source][unit]``; a store carries the epoch
tag and the payload in one word.  The model interleaves the ranks' atomic
steps (one remote store, one local poll) in random order and checks, for a
sequence of collectives,

* **all-reduce** (every rank pushes to every peer, then polls its own slots):
  each rank reads exactly the payload its peer sent for *this* epoch, with
  only two parities -- a rank can be at most one exchange ahead of a peer;
* **broadcast, ready-first** (every receiver announces itself with a tagged
  word in a unit of its own, the root pushes only when all have, receivers
  poll): nobody can finish collective e+1 before everybody has finished e,
  whatever mix of collectives and roots -- the invariant the two parities
  rest on;
* **broadcast, acknowledge-after** (the root pushes at once and then waits for
  acknowledgements -- the previous version): closes back-to-back broadcasts of
  ONE root, but a fast rank that becomes the root two collectives after an
  all-reduce (or after acknowledging another root) overwrites words a slow
  peer has not read; the model finds both, and more often with the
  acknowledgement in the data unit;
* **broadcast without acknowledgements** (the round-1 kernel): the model finds
  the overrun the advisor described -- epoch e+2 overwrites the slot of epoch e
  before a slow receiver has read it, and that receiver waits for ever.

    python tools/ll_protocol_model.py
"""
import random


class Stuck(Exception):
    pass


def run(world, ops, seed, bcast_ack=True, ack_unit=1, ready_first=False,
        max_steps=200_000):
    """ops: list of 'ar' / ('bc', root).  Returns the number of steps."""
    rng = random.Random(seed)
    # slots[dst][parity][src][unit] = (tag, payload); unit 0 carries data,
    # acknowledgements go to unit `ack_unit` (0 = share the data unit, as the
    # first version of the acknowledged broadcast did)
    slots = [[[[(0, None), (0, None)] for _ in range(world)]
              for _ in range(2)] for _ in range(world)]

    def rank_program(me):
        epoch = 0
        for k, op in enumerate(ops):
            epoch += 1
            par = epoch & 1
            if op == "ar":
                for d in range(world):
                    if d != me:
                        slots[d][par][me][0] = (epoch, ("ar", k, me))
                        yield
                for s in range(world):
                    if s != me:
                        while slots[me][par][s][0][0] != epoch:
                            yield "wait"
                        assert slots[me][par][s][0][1] == ("ar", k, s), \
                            (me, s, k, slots[me][par][s][0])
                        yield
            else:
                root = op[1]
                if me == root:
                    if ready_first:
                        for s in range(world):
                            if s != me:
                                while slots[me][par][s][1][0] != epoch:
                                    if slots[me][par][s][1][0] > epoch:
                                        raise Stuck("ready token overwritten")
                                    yield "wait"
                                yield
                    for d in range(world):
                        if d != me:
                            slots[d][par][root][0] = (epoch, ("bc", k))
                            yield
                    if bcast_ack and not ready_first:
                        for s in range(world):
                            if s != me:
                                while slots[me][par][s][ack_unit][0] != epoch:
                                    if slots[me][par][s][ack_unit][0] > epoch:
                                        raise Stuck(
                                            f"root {me} waits for the ack of "
                                            f"rank {s} for broadcast {k} but "
                                            "the slot already carries a later "
                                            "epoch")
                                    yield "wait"
                                yield
                else:
                    if ready_first:
                        slots[root][par][me][1] = (epoch, ("ready", k, me))
                        yield
                    while slots[me][par][root][0][0] != epoch:
                        if slots[me][par][root][0][0] > epoch:
                            raise Stuck(
                                f"rank {me} waits for broadcast {k} (epoch "
                                f"{epoch}) but its slot already carries epoch "
                                f"{slots[me][par][root][0][0]}")
                        yield "wait"
                    assert slots[me][par][root][0][1] == ("bc", k)
                    yield
                    if bcast_ack and not ready_first:
                        slots[root][par][me][ack_unit] = (epoch, ("ack", k, me))
                        yield
        return

    progs = [rank_program(r) for r in range(world)]
    alive = list(range(world))
    # a biased scheduler: one random rank is slow for long stretches
    slow, steps, idle = rng.randrange(world), 0, 0
    while alive:
        steps += 1
        if steps > max_steps:
            raise Stuck("no progress (deadlock)")
        if rng.random() < 0.01:
            slow = rng.randrange(world)
        r = rng.choice(alive)
        if r == slow and rng.random() < 0.9:
            continue
        try:
            what = next(progs[r])
        except StopIteration:
            alive.remove(r)
            continue
        idle = idle + 1 if what == "wait" else 0
        if idle > 50_000:
            raise Stuck("every runnable rank only waits (deadlock)")
    return steps


def main():
    mixes = {
        "all-reduces": ["ar"] * 12,
        "broadcasts, alternating roots": [("bc", k % 3) for k in range(12)],
        "mixed": ["ar", ("bc", 0), ("bc", 0), "ar", ("bc", 1), ("bc", 2),
                  ("bc", 2), ("bc", 0), "ar", "ar", ("bc", 1)],
    }
    bad = shared = 0
    for name, ops in mixes.items():
        for world in (2, 3, 4):
            seq = [o if o == "ar" else ("bc", o[1] % world) for o in ops]
            fails = fails_shared = fails_ack = 0
            for seed in range(150):
                try:
                    run(world, seq, seed, ready_first=True)
                except (Stuck, AssertionError) as e:
                    fails += 1
                    print("  ", name, world, seed, e)
                try:
                    run(world, seq, seed)
                except (Stuck, AssertionError):
                    fails_ack += 1
                try:
                    run(world, seq, seed, ack_unit=0)
                except (Stuck, AssertionError):
                    fails_shared += 1
            print(f"{name:32s} W={world}: {fails}/150 schedules fail "
                  f"(acknowledge-after instead of ready-first: {fails_ack}, "
                  f"with the ack in the data unit: {fails_shared})")
            bad += fails
            shared += fails_shared + fails_ack
    # the round-1 broadcast (no acknowledgement) must be caught
    caught = 0
    for seed in range(150):
        try:
            run(3, [("bc", 0)] * 8, seed, bcast_ack=False)
        except Stuck:
            caught += 1
    print(f"broadcast WITHOUT acknowledgements, W=3: overrun found in "
          f"{caught}/150 schedules")
    raise SystemExit(1 if bad or caught == 0 or shared == 0 else 0)


if __name__ == "__main__":
    main()
