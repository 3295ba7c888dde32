#!/usr/bin/env python
"""Discrete-event model of the linear-link kernel's barrier protocol
(``csrc/linlink.cu``): producer (TMA), issuer (Jacobian UMMAs, in tile order)
and SETS compute sets over STAGES W stages and NGB G buffers, with mbarrier
*parity* waits exactly as the hardware evaluates them -- ``wait(p)`` passes
iff the barrier's current phase has the other parity -- and TMA completions
that arrive out of order.

It checks, on every wait that passes, that the phase the code *meant* is the
one that completed (a wait that passes on the phase before it is the aliasing
bug found at P = 128: wrong sums, barrier time-outs, launch failures), that no
buffer is overwritten while it is still being read, and that the pipeline
drains (no deadlock).

    python tools/linlink_pipeline_model.py            # current protocol: clean
    python tools/linlink_pipeline_model.py --old      # full_w waited first: fails
"""
import argparse
import heapq
import random


class Barrier:
    def __init__(self, count):
        self.count, self.pending, self.completed = count, count, 0

    def arrive(self):
        self.pending -= 1
        assert self.pending >= 0, "more arrivals than the phase expects"
        if self.pending == 0:
            self.pending = self.count
            self.completed += 1

    def passes(self, parity):
        return (self.completed & 1) != parity


class Violation(Exception):
    pass


def run(stages, sets, ngb, tiles, seed, old_order=False, tma_jitter=6.0):
    rng = random.Random(seed)
    full_w = [Barrier(1) for _ in range(stages)]
    empty = [Barrier(1) for _ in range(stages)]
    g_full = [Barrier(4) for _ in range(ngb)]
    g_free = [Barrier(1) for _ in range(ngb)]
    stage_tile = [None] * stages          # tile whose data the stage holds
    gbuf_tile = [None] * ngb
    events, now = [], [0.0]
    seq = [0]

    def at(dt, fn):
        seq[0] += 1
        heapq.heappush(events, (now[0] + dt, seq[0], fn))

    def wait(bar, parity, then, meant):
        """Poll until the parity test passes; check it passed on `meant`."""
        def poll():
            if bar.passes(parity):
                if bar.completed != meant + 1:
                    raise Violation(
                        f"parity wait passed with {bar.completed} completed "
                        f"phases, meant phase {meant}")
                then()
            else:
                at(0.05, poll)
        poll()

    done = {"tiles": 0}

    # ---- producer: tile it -> stage it % stages, after empty of it-stages
    def produce(it):
        if it >= tiles:
            return
        s, r = it % stages, it // stages

        def issue():
            def land():
                stage_tile[s] = it
                full_w[s].arrive()
            at(rng.uniform(0.5, 0.5 + tma_jitter), land)   # unordered landing
            produce(it + 1)
        if r == 0:
            issue()
        else:
            wait(empty[s], (r & 1) ^ 1, issue, r - 1)

    # ---- issuer: in tile order
    def issue_jac(n2):
        if n2 >= tiles:
            return
        s, gb = n2 % stages, n2 % ngb

        def go():
            if stage_tile[s] != n2 or gbuf_tile[gb] != n2:
                raise Violation(f"Jacobian product of tile {n2} reads stage "
                                f"of {stage_tile[s]} / G of {gbuf_tile[gb]}")

            def complete():
                empty[s].arrive()
                g_free[gb].arrive()
                done["tiles"] += 1
            at(rng.uniform(0.1, 0.3), complete)
            issue_jac(n2 + 1)
        wait(g_full[gb], (n2 // ngb) & 1, go, n2 // ngb)

    # ---- compute warps (4 per set: they wait separately, arrive separately)
    def compute(it, warp):
        if it >= tiles:
            return
        s, gb = it % stages, it % ngb

        def have_tile():
            if stage_tile[s] != it:
                raise Violation(f"compute of tile {it} reads stage {s} "
                                f"holding tile {stage_tile[s]}")

            def publish():
                if gbuf_tile[gb] is not None and gbuf_tile[gb] != it and \
                        g_free[gb].completed < it // ngb:
                    raise Violation(f"G buffer {gb} overwritten by tile {it} "
                                    "before its product completed")
                gbuf_tile[gb] = it
                g_full[gb].arrive()
                compute(it + sets, warp)
            at(rng.uniform(0.4, 1.6), publish)

        def wait_tile():
            wait(full_w[s], (it // stages) & 1, have_tile, it // stages)

        if old_order:                 # the tile first, the G buffer afterwards
            def after_tile():
                if stage_tile[s] != it:
                    raise Violation(f"compute of tile {it} reads stage {s} "
                                    f"holding tile {stage_tile[s]}")
                if it >= ngb:
                    wait(g_free[gb], ((it // ngb) + 1) & 1,
                         lambda: at(rng.uniform(0.4, 1.6), lambda: (
                             gbuf_tile.__setitem__(gb, it),
                             g_full[gb].arrive(),
                             compute(it + sets, warp))),
                         it // ngb - 1)
                else:
                    at(rng.uniform(0.4, 1.6), lambda: (
                        gbuf_tile.__setitem__(gb, it), g_full[gb].arrive(),
                        compute(it + sets, warp)))
            wait(full_w[s], (it // stages) & 1, after_tile, it // stages)
        elif it >= ngb:
            wait(g_free[gb], ((it // ngb) + 1) & 1, wait_tile, it // ngb - 1)
        else:
            wait_tile()

    produce(0)
    issue_jac(0)
    for st in range(sets):
        for w in range(4):
            compute(st, w)
    limit = 200.0 * tiles
    while events:
        t, _, fn = heapq.heappop(events)
        if t > limit:
            raise Violation(f"deadlock: {done['tiles']} of {tiles} tiles done")
        now[0] = t
        fn()
    if done["tiles"] != tiles:
        raise Violation(f"stalled: {done['tiles']} of {tiles} tiles done")
    return now[0]


CONFIGS = [(8, 3, 4), (5, 3, 3), (8, 4, 5), (5, 2, 2), (8, 3, 3), (5, 3, 5)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--old", action="store_true")
    ap.add_argument("--seeds", type=int, default=200)
    ap.add_argument("--tiles", type=int, default=120)
    args = ap.parse_args()
    bad = 0
    for stages, sets, ngb in CONFIGS:
        fails = 0
        for seed in range(args.seeds):
            try:
                run(stages, sets, ngb, args.tiles, seed, old_order=args.old)
            except Violation as e:
                fails += 1
                if fails == 1:
                    print(f"  stages={stages} sets={sets} ngb={ngb} seed={seed}: {e}")
        print(f"stages={stages} sets={sets} ngb={ngb}: {fails}/{args.seeds} runs violate")
        bad += fails
    raise SystemExit(1 if bad and not args.old else 0)


if __name__ == "__main__":
    main()
