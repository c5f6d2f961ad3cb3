"""Model check of the cross-GPU signal-pad protocol of csrc/comm_sm100.cu (rs_adam_ag_kernel, allreduce_small_kernel).

compute-sanitizer's racecheck is closed on the GPU pool, so the protocol itself is checked here by
exhaustive exploration of the interleavings of a small model:

* every rank owns a pad ``sig[slot][phase][src]``; a launch with epoch ``e = state + 1`` has each rank STORE ``e`` into
  entry ``[phase 0][rank]`` of EVERY rank's pad (one store per destination, in any order, at any time), then WAIT until
  every entry of its OWN phase-0 row is ``>= e`` (signed distance, as ``(int)(v - e) >= 0``), do its shard's work,
  and repeat the same with the phase-1 row; the launch ends with ``state = e``;
* launches of one rank are stream-ordered; different ranks run freely, so a fast rank may already be storing the
  signals of launch ``e + 1`` while a slow one still waits in launch ``e``.

Properties checked for every reachable interleaving (2 and 3 ranks, 2 consecutive launches):
  1. safety: when a rank passes the phase-0 wait of launch e, every rank has finished writing its gradients of
     launch e (modelled as "has started launch e"); when it passes the phase-1 wait, every rank has finished its
     shard update of launch e;
  2. no deadlock: every interleaving ends with all ranks done;
  3. epochs never run backwards in a pad (an old store cannot overwrite a newer one), which is what makes the
     ``>=`` comparison sound when a fast rank is one launch ahead.
"""
from __future__ import annotations

import itertools


def _explore(world: int, launches: int):
    """Depth-first search over all schedules.  A rank's program is a list of atomic steps."""
    # program of one rank: for each launch e: world stores (phase 0), wait0, work, world stores (phase 1), wait1
    def program(rank):
        steps = []
        for e in range(1, launches + 1):
            steps.append(("begin", e))
            for dst in range(world):
                steps.append(("store", 0, dst, e))
            steps.append(("wait", 0, e))
            steps.append(("work", e))
            for dst in range(world):
                steps.append(("store", 1, dst, e))
            steps.append(("wait", 1, e))
        return steps

    progs = [program(r) for r in range(world)]
    # state: (pc per rank, pad values [dst][phase][src], begun[e-1 per rank], worked[...])
    init_pad = tuple(tuple(tuple(0 for _ in range(world)) for _ in range(2)) for _ in range(world))
    init = (tuple(0 for _ in range(world)), init_pad, tuple(0 for _ in range(world)), tuple(0 for _ in range(world)))
    seen, stack, finals = set(), [init], 0
    while stack:
        st = stack.pop()
        if st in seen:
            continue
        seen.add(st)
        pcs, pad, begun, worked = st
        if all(pcs[r] == len(progs[r]) for r in range(world)):
            finals += 1
            continue
        progressed = False
        for r in range(world):
            if pcs[r] == len(progs[r]):
                continue
            op = progs[r][pcs[r]]
            npad, nbegun, nworked = pad, begun, worked
            if op[0] == "begin":
                nbegun = begun[:r] + (op[1],) + begun[r + 1:]
            elif op[0] == "store":
                _, phase, dst, e = op
                old = pad[dst][phase][r]
                assert e >= old, "an epoch ran backwards in a pad"          # property 3
                row = pad[dst][phase][:r] + (e,) + pad[dst][phase][r + 1:]
                d = pad[dst][:phase] + (row,) + pad[dst][phase + 1:]
                npad = pad[:dst] + (d,) + pad[dst + 1:]
            elif op[0] == "wait":
                _, phase, e = op
                if not all(v - e >= 0 for v in pad[r][phase]):
                    continue                                                 # blocked: cannot take this step now
                if phase == 0:                                               # property 1
                    assert all(b >= e for b in begun), "passed the gradient barrier before a peer's gradients"
                else:
                    assert all(w >= e for w in worked), "passed the parameter barrier before a peer's update"
            elif op[0] == "work":
                nworked = worked[:r] + (op[1],) + worked[r + 1:]
            progressed = True
            stack.append((pcs[:r] + (pcs[r] + 1,) + pcs[r + 1:], npad, nbegun, nworked))
        assert progressed, f"deadlock at {pcs}"                              # property 2
    return len(seen), finals


def test_two_ranks_two_launches():
    states, finals = _explore(2, 2)
    assert finals == 1 and states > 100


def test_three_ranks_two_launches():
    states, finals = _explore(3, 2)
    assert finals == 1 and states > 1000


def test_wrapping_epoch_comparison():
    """The wait compares with a signed distance, so the 32-bit epoch may wrap: (int)(v - e) >= 0."""
    def reached(v, e):
        d = (v - e) & 0xFFFFFFFF
        return d < 0x80000000
    assert reached(5, 5) and reached(6, 5) and not reached(4, 5)
    assert reached(0, 0xFFFFFFFF) and reached(1, 0xFFFFFFFF) and not reached(0xFFFFFFFE, 0xFFFFFFFF)
    assert all(reached((e + k) & 0xFFFFFFFF, e) for e, k in itertools.product((0, 7, 0xFFFFFFF0), (0, 1, 2)))
