"""Model check of the copy-stream protocol of hj_xchg_* (csrc/hj_api.cu, xchg_scatter_stage): a discrete-event
simulation of G copy streams that enqueue, per stage and in ring order, [wait for the successor's token of round k-1]
-> block copy to rank+k -> [token to the predecessor] -> flag, with random copy durations and start skews.

What the product relies on and this file checks for G = 2, 4, 8:
  * the schedule never deadlocks, whatever the timing;
  * inside a stage no GPU receives from two senders at once (the reason the tokens exist: two copy engines writing
    into one GPU halve each other, profiles/r02_ce_probe_ipc_n4.jsonl);
  * without tokens the same timings DO produce such overlaps (the model is not vacuous).
The product's C++ is exercised on the GPU by test_xchg_pipeline_logical_ranks; this is the protocol on paper."""
import random

import pytest


def simulate(G, stages, tokens, rng):
    """Returns (finished, overlaps): overlaps = pairs of copies of the SAME stage into one destination whose
    intervals intersect."""
    # per-rank op list, as xchg_scatter_stage enqueues it on the one copy stream
    ops = []
    for r in range(G):
        seq = []
        for t in range(stages):
            for k in range(1, G):
                if tokens and k >= 2:
                    seq.append(("wait", t, k - 1))                 # token written by the successor after ITS round k-1
                seq.append(("copy", t, k, (r + k) % G))
                if tokens and k + 2 <= G:
                    seq.append(("token", t, k, (r - 1) % G))      # my round k is done: the predecessor may ship k+1
        ops.append(seq)
    clock = [rng.uniform(0.0, 2.0) for _ in range(G)]              # start skew, in units of one copy
    pc = [0] * G
    tok_time = {}                                                  # (rank, t, k) -> time the token became visible
    copies = []                                                    # (dest, t, begin, end)
    progressed = True
    while progressed:
        progressed = False
        for r in sorted(range(G), key=lambda i: clock[i]):
            while pc[r] < len(ops[r]):
                op = ops[r][pc[r]]
                if op[0] == "wait":
                    key = (r, op[1], op[2])
                    if key not in tok_time:
                        break                                      # blocked: another rank has to run first
                    clock[r] = max(clock[r], tok_time[key])
                elif op[0] == "copy":
                    d = rng.uniform(0.7, 1.4)
                    copies.append((op[3], op[1], clock[r], clock[r] + d))
                    clock[r] += d
                else:
                    clock[r] += 0.01
                    tok_time[(op[3], op[1], op[2])] = clock[r]
                pc[r] += 1
                progressed = True
    finished = all(pc[r] == len(ops[r]) for r in range(G))
    overlaps = 0
    for i, (d1, t1, b1, e1) in enumerate(copies):
        for d2, t2, b2, e2 in copies[i + 1:]:
            if d1 == d2 and t1 == t2 and b1 < e2 - 1e-9 and b2 < e1 - 1e-9:
                overlaps += 1
    return finished, overlaps


@pytest.mark.parametrize("G", [2, 4, 8])
def test_token_ring_never_deadlocks_and_serialises_the_senders_of_a_stage(G):
    rng = random.Random(1234 + G)
    for _ in range(40):
        finished, overlaps = simulate(G, stages=5, tokens=True, rng=rng)
        assert finished
        assert overlaps == 0


def test_without_tokens_the_same_timings_collide():
    rng = random.Random(99)
    total = 0
    for _ in range(20):
        finished, overlaps = simulate(4, stages=5, tokens=False, rng=rng)
        assert finished
        total += overlaps
    assert total > 0
