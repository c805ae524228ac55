#!/usr/bin/env python3
"""Discrete-event model of one chained encoder-block launch (wq4_chain.cu) used to choose the tile ORDER.
74 CTA pairs walk a static tile list round-robin; a tile's mainloop needs its producer row block complete."""
import sys
PAIRS = 74
def build(order, mt, prob):
    # prob: list of (name, npairs, kst, epi_us, dep)
    tiles = []
    if order == "problem":
        for j, (nm, npr, kst, e, dep) in enumerate(prob):
            for rb in range(mt):
                for cp in range(npr): tiles.append((j, rb))
    elif order.startswith("lag"):
        # out first (no deps), then groups: fc1(r), fc2(r - l2), qkv(r - l3)
        l2, l3 = (int(v) for v in order[3:].split(","))
        for rb in range(mt):
            for cp in range(prob[0][1]): tiles.append((0, rb))
        for r in range(mt + max(l2, l3)):
            for j, lag in ((1, 0), (2, l2), (3, l3)):
                if j >= len(prob): continue
                rb = r - lag
                if 0 <= rb < mt:
                    for cp in range(prob[j][1]): tiles.append((j, rb))
    return tiles
def simulate(tiles, mt, prob, stage_us=0.42, sig_us=1.3, ttfm=0.6):
    import collections
    done = collections.defaultdict(float)      # (j, rb) -> completion time of the row block
    left = {}
    for j, (nm, npr, kst, e, dep) in enumerate(prob):
        for rb in range(mt): left[(j, rb)] = npr
    # per pair state
    ml_end = [0.0] * PAIRS; epi_end = [[0.0, 0.0] for _ in range(PAIRS)]; epi_last = [0.0] * PAIRS; cnt = [0] * PAIRS
    # tiles are processed in list order per pair; need global event ordering: iterate rounds, but deps may point to tiles of same round
    # -> process tiles in list order; a tile's dep always earlier in the list, so its completion is known when we get to it
    finish = {}
    for g, (j, rb) in enumerate(tiles):
        p = g % PAIRS
        nm, npr, kst, e, dep = prob[j]
        ready = done_time(finish, tiles, prob, mt, dep, rb) + ttfm if dep >= 0 else 0.0
        acc_free = epi_end[p][cnt[p] % 2]                      # accumulator buffer of two tiles ago
        start = max(ml_end[p], ready, acc_free)
        mend = start + kst * stage_us
        estart = max(mend, epi_last[p])
        eend = estart + e + sig_us
        ml_end[p] = mend; epi_last[p] = eend; epi_end[p][cnt[p] % 2] = eend; cnt[p] += 1
        finish.setdefault((j, rb), []).append(eend)
    return max(epi_last)
_cache = {}
def done_time(finish, tiles, prob, mt, dep, rb):
    v = finish.get((dep, rb))
    return max(v) if v else 0.0
if __name__ == "__main__":
    D, F, rows, tm = 1024, 4096, 1500, 128
    if len(sys.argv) > 1: D, F, rows, tm = (int(v) for v in sys.argv[1:5])
    mt = (rows + tm - 1) // tm
    sc = tm / 128
    prob = [("out", D // 256, D // 128, 5.0 * sc, -1), ("fc1", F // 256, D // 128, 4.3 * sc, 0), ("fc2", D // 256, F // 128, 5.0 * sc, 1), ("qkv", 3 * D // 256, D // 128, 2.1 * sc, 2)]
    st = 0.42 if tm == 128 else 0.54
    for order in ["problem", "lag1,2", "lag2,4", "lag2,6", "lag3,6", "lag4,8", "lag1,4", "lag0,2", "lag3,8"]:
        t = build(order, mt, prob)
        print(order, len(t), round(simulate(t, mt, prob, stage_us=st), 1))
