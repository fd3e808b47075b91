"""Phase breakdown of k_bt_select from an LGB_B200_TRACE timeline (ids 41 ready, 43 P0 done, 44 replay done, 45 bookkeeping done,
46 ranks done, 47 batch written, 42 end)."""
import sys, collections
ev = sorted((tuple(map(int, l.split())) for l in open(sys.argv[1])), key=lambda e: e[1])
seq = [41, 43, 44, 45, 46, 47, 42]
names = ["P0 loads", "P1 replay", "P2-P3 commit", "P4 + ranks", "batch items", "tail"]
agg = collections.defaultdict(list)
cur = {}
for id_, t in ev:
    if id_ == 41: cur = {41: t}
    elif id_ in seq and cur: cur[id_] = t
    if id_ == 42 and len(cur) == len(seq):
        for a, b, n in zip(seq, seq[1:], names): agg[n].append((cur[b] - cur[a]) / 1e3)
        cur = {}
for n in names:
    v = sorted(agg[n])
    if v: print("%-14s n=%4d mean %6.2f us median %6.2f p90 %6.2f" % (n, len(v), sum(v) / len(v), v[len(v) // 2], v[int(len(v) * .9)]))
