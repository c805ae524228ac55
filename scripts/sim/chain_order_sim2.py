#!/usr/bin/env python3
"""List-scheduling model of a chained launch: tiles in a global order, each assigned to the pair whose mainloop frees first."""
import sys, itertools
PAIRS = 74
def order_tiles(kind, mt, prob, lags=None):
    tiles = []
    if kind == "problem":
        for j, pr in enumerate(prob):
            for rb in range(mt):
                tiles += [(j, rb)] * pr[1]
    else:   # lagged wavefront: group r = out(r) fc1(r-l1) fc2(r-l2) qkv(r-l3)
        for r in range(mt + max(lags)):
            for j, lag in enumerate(lags):
                if j >= len(prob): continue
                rb = r - lag
                if 0 <= rb < mt: tiles += [(j, rb)] * prob[j][1]
    return tiles
def simulate(tiles, prob, stage_us, sig_us=1.3, ttfm=0.6, greedy=True):
    ml_end = [0.0] * PAIRS; epi_last = [0.0] * PAIRS; accs = [[0.0, 0.0] for _ in range(PAIRS)]; cnt = [0] * PAIRS
    finish = {}; assign = []
    for g, (j, rb) in enumerate(tiles):
        nm, npr, kst, e, dep = prob[j]
        ready = (max(finish[(dep, rb)]) + ttfm) if dep >= 0 else 0.0
        if greedy:
            p = min(range(PAIRS), key=lambda q: (max(ml_end[q], accs[q][cnt[q] % 2]), q))
        else:
            p = g % PAIRS
        start = max(ml_end[p], ready, accs[p][cnt[p] % 2])
        mend = start + kst * stage_us
        eend = max(mend, epi_last[p]) + e + sig_us
        ml_end[p] = mend; epi_last[p] = eend; accs[p][cnt[p] % 2] = eend; cnt[p] += 1
        finish.setdefault((j, rb), []).append(eend)
        assign.append(p)
    return max(epi_last), assign
if __name__ == "__main__":
    D, F, rows, tm = 1024, 4096, 1500, 128
    if len(sys.argv) > 1: D, F, rows, tm = (int(v) for v in sys.argv[1:5])
    mt = (rows + tm - 1) // tm
    sc = tm / 128
    prob = [("out", D // 256, D // 128, 5.0 * sc, -1), ("fc1", F // 256, D // 128, 4.3 * sc, 0), ("fc2", D // 256, F // 128, 5.0 * sc, 1), ("qkv", 3 * D // 256, D // 128, 2.1 * sc, 2)]
    st = 0.42 if tm == 128 else 0.54
    print("problem rr", round(simulate(order_tiles("problem", mt, prob), prob, st, greedy=False)[0], 1))
    print("problem greedy", round(simulate(order_tiles("problem", mt, prob), prob, st)[0], 1))
    best = []
    for l1 in range(0, 5):
        for l2 in range(l1, l1 + 6):
            for l3 in range(l2, l2 + 8):
                t, _ = simulate(order_tiles("lag", mt, prob, (0, l1, l2, l3)), prob, st)
                best.append((t, (0, l1, l2, l3)))
    best.sort()
    print(best[:8])
    work = sum(pr[1] * mt * pr[2] * st for pr in prob) / PAIRS
    print("mainloop-only bound", round(work, 1))
