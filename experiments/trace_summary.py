"""Summarise an LGB_B200_TRACE timeline: per kernel (CTA 0): start->after-wait, after-wait->end, and gaps between kernels."""
import sys, collections
ev = [tuple(map(int, l.split())) for l in open(sys.argv[1])]
ev.sort(key=lambda e: e[1])
names = {1: "partition", 2: "hist", 3: "find", 4: "select"}
t0 = ev[0][1]
rows = []
cur = {}
last_end = None
agg = collections.defaultdict(list)
for id_, t in ev:
    k, ph = id_ // 10, id_ % 10
    if k not in names: continue
    if ph == 0: cur[k] = [t, None, None]
    elif ph == 1 and k in cur: cur[k][1] = t
    elif ph == 2 and k in cur and cur[k][1] is not None:
        s, w, _ = cur.pop(k)
        agg[names[k] + " wait(start->deps done)"].append((w - s) / 1e3)
        agg[names[k] + " body(CTA0)"].append((t - w) / 1e3)
        if last_end is not None: agg[names[k] + " gap(prev CTA0 end->deps done)"].append((w - last_end) / 1e3)
        last_end = t
        rows.append((names[k], (s - t0) / 1e3, (w - t0) / 1e3, (t - t0) / 1e3))
for r in rows[:60]:
    print("%-10s start %9.1f  ready %9.1f  end %9.1f" % r)
print()
for k, v in sorted(agg.items()):
    v2 = sorted(v)
    print("%-45s n=%5d  mean %7.2f us  median %7.2f  p90 %7.2f" % (k, len(v), sum(v) / len(v), v2[len(v2) // 2], v2[int(len(v2) * 0.9)]))
