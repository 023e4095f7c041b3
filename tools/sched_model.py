#!/usr/bin/env python
"""Lane model of k_render_pipe's traversal rounds (csrc/pipeline.cuh), fed with the REAL visit sequence of every ray
(tests/sim/path_sim ... <events file>): what a round schedule costs in warp-instructions per ray. No GPU.

  python tools/sched_model.py <events file> [thresh] [max_rounds]

A warp holds 32 rays; a round is a fixed sequence of visits, each executed by the whole warp (a node visit always, a leaf visit
only if some lane is at a leaf); a lane advances through its own sequence when the visit matches what it needs next. After a
round the warp votes: with fewer than `thresh` lanes busy, or after `max_rounds` rounds, it pushes the finished rays and refills
the idle lanes. Instruction counts are read off the SASS of the default kernel (profiles/r01_l_pipe_c2_by_line.txt): node
visit 56, leaf visit 32 (3 when no lane is at a leaf), vote + loop 15, refill of idle lanes 120, push of finished rays 25.
The model ignores the shade / generate batches (a constant per ray) and queue contention. It reproduces the lane counts ncu
measured for the default schedule (node visits 29.6 and 21.8 lanes, leaf visits 10.7), which is what it is calibrated by.
"""
import random
import sys

N_COST, L_COST, L_SKIP, VOTE, REFILL, PUSH = 56, 32, 3, 15, 120, 25


def load(path):
    data = open(path, "rb").read()
    rays = data.split(b"\0")
    if rays and rays[-1] == b"":
        rays.pop()
    return rays


def simulate(rays, schedule, thresh, max_rounds, seed=1):
    order = list(range(len(rays)))
    random.Random(seed).shuffle(order)  # the EXT queue mixes primary and scattered rays
    nxt = 0
    seq = [b""] * 32
    pos = [0] * 32
    cost = 0
    lanes = {k: [0, 0] for k in range(len(schedule))}  # visit index in the round -> [executions, busy lanes]
    while True:
        idle = [i for i in range(32) if pos[i] >= len(seq[i])]
        if idle and nxt < len(order):
            cost += REFILL
            for i in idle:
                if nxt >= len(order):
                    break
                seq[i], pos[i] = rays[order[nxt]], 0
                nxt += 1
        busy = sum(1 for i in range(32) if pos[i] < len(seq[i]))
        if busy == 0:
            break
        starving = nxt >= len(order)
        rounds = 0
        while True:
            for k, what in enumerate(schedule):
                want = ord(what)
                hit = [i for i in range(32) if pos[i] < len(seq[i]) and seq[i][pos[i]] == want]
                if what == "N":
                    cost += N_COST
                else:
                    cost += L_COST if hit else L_SKIP
                if hit or what == "N":
                    lanes[k][0] += 1
                    lanes[k][1] += len(hit)
                for i in hit:
                    pos[i] += 1
            cost += VOTE
            rounds += 1
            busy = sum(1 for i in range(32) if pos[i] < len(seq[i]))
            if busy < (1 if starving else thresh) or rounds >= max_rounds:
                break
        cost += PUSH
    return cost / len(rays), {k: (v[1] / max(v[0], 1)) for k, v in lanes.items()}


def main():
    rays = load(sys.argv[1])
    thresh = int(sys.argv[2]) if len(sys.argv) > 2 else 14
    max_rounds = int(sys.argv[3]) if len(sys.argv) > 3 else 16
    n = sum(r.count(b"N") for r in rays) / len(rays)
    l = sum(r.count(b"L") for r in rays) / len(rays)
    print(f"{len(rays)} rays: {n:.3f} node visits and {l:.3f} leaf visits per ray; thresh {thresh}, max_rounds {max_rounds}")
    base = None
    for name, schedule in (("N L N L (default)", "NLNL"), ("N N L (CRT_PIPE_SCHED=1)", "NNL"), ("N N N L", "NNNL"), ("N L", "NL"), ("N N L L", "NNLL")):
        c, lanes = simulate(rays, schedule, thresh, max_rounds)
        base = base or c
        occ = ", ".join(f"{schedule[k]}{k + 1}: {v:.1f}" for k, v in lanes.items())
        print(f"  {name:26s} {c:7.2f} warp-instructions per ray ({100 * (c / base - 1):+5.1f} %)   busy lanes per executed visit: {occ}")


if __name__ == "__main__":
    main()
