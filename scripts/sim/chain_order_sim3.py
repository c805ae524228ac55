#!/usr/bin/env python3
"""Model of a chained launch with measured per-tile costs (profiles/r02_chain_timeline_medium_b1.txt); explores tile orders
and static tile -> pair assignments.  A tile: dependency ready -> X/MMA mainloop (stage_us per K stage) -> epilogue -> release."""
import sys
PAIRS = 74
def order(kind, mt, prob, lags=None):
    t = []
    if kind == "problem":
        for j, pr in enumerate(prob):
            for rb in range(mt): t += [(j, rb)] * pr[1]
    else:
        nprob = len(prob)
        for r in range(mt + max(lags)):
            for j in range(nprob):
                rb = r - lags[j]
                if 0 <= rb < mt: t += [(j, rb)] * prob[j][1]
    return t
def simulate(tiles, prob, greedy, ttfm=0.8, sig=0.95):
    ml = [0.0] * PAIRS; epi = [0.0] * PAIRS; accs = [[0.0, 0.0] for _ in range(PAIRS)]; cnt = [0] * PAIRS
    fin = {}; assign = []
    for g, (j, rb) in enumerate(tiles):
        nm, npr, kst, st, e, dep = prob[j]
        ready = (max(fin[(dep, rb)]) + ttfm) if dep >= 0 else 0.0
        p = min(range(PAIRS), key=lambda q: (max(ml[q], accs[q][cnt[q] % 2]), q)) if greedy else g % PAIRS
        start = max(ml[p], ready, accs[p][cnt[p] % 2])
        mend = start + kst * st
        eend = max(mend + 0.5, epi[p]) + e + sig
        ml[p] = mend; epi[p] = eend; accs[p][cnt[p] % 2] = eend; cnt[p] += 1
        fin.setdefault((j, rb), []).append(eend); assign.append(p)
    return max(epi), assign
if __name__ == "__main__":
    D, F, rows, tm = 1024, 4096, 1500, 128
    mt = (rows + tm - 1) // tm
    # name, column pairs, K stages, us per stage, epilogue us, dep
    prob = [("out", D // 256, D // 128, 0.34, 4.6, -1), ("fc1", F // 256, D // 128, 0.32, 3.6, 0),
            ("fc2", D // 256, F // 128, 0.27, 3.7, 1), ("qkv", 3 * D // 256, D // 128, 0.32, 2.3, 2)]
    print("problem-major, round robin (what runs):", round(simulate(order("problem", mt, prob), prob, False)[0], 1))
    print("problem-major, greedy:", round(simulate(order("problem", mt, prob), prob, True)[0], 1))
    res = []
    for l1 in range(0, 7):
        for l2 in range(l1, l1 + 7):
            for l3 in range(l2, l2 + 10):
                for gr in (False, True):
                    res.append((simulate(order("lag", mt, prob, (0, l1, l2, l3)), prob, gr)[0], (0, l1, l2, l3), gr))
    res.sort()
    for r in res[:6]: print(round(r[0], 1), r[1], "greedy" if r[2] else "round robin")
