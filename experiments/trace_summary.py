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
sub = {}
hist_ready, hist_end = [], []
for id_, t in ev:
    k, ph = id_ // 10, id_ % 10
    if k not in names: continue
    if k == 2 and ph in (3, 4):      # histogram kernel: 23 = set-up done, 24 = CTA 0's flush began
        sub.setdefault(ph, []).append(t)
        continue
    if ph == 0: cur[k] = [t, None, None]
    elif ph == 1 and k in cur: cur[k][1] = t
    elif ph == 2 and k in cur and cur[k][1] is not None:
        s, w, _ = cur.pop(k)
        agg[names[k] + " wait(start->deps done)"].append((w - s) / 1e3)
        agg[names[k] + " body(CTA0)"].append((t - w) / 1e3)
        if last_end is not None: agg[names[k] + " gap(prev CTA0 end->deps done)"].append((w - last_end) / 1e3)
        last_end = t
        rows.append((names[k], (s - t0) / 1e3, (w - t0) / 1e3, (t - t0) / 1e3))
        if k == 2: hist_ready.append(w); hist_end.append(t)
for r in rows[:60]:
    print("%-10s start %9.1f  ready %9.1f  end %9.1f" % r)
print()
for k, v in sorted(agg.items()):
    v2 = sorted(v)
    print("%-45s n=%5d  mean %7.2f us  median %7.2f  p90 %7.2f" % (k, len(v), sum(v) / len(v), v2[len(v2) // 2], v2[int(len(v2) * 0.9)]))

# histogram kernel phases of CTA 0: ready -> set-up done -> flush begins -> end (matched by time order)
import bisect
if sub.get(3):
    setup, rowsph, flush = [], [], []
    s3, s4 = sorted(sub.get(3, [])), sorted(sub.get(4, []))
    for w, e in zip(hist_ready, hist_end):
        i = bisect.bisect_left(s3, w)
        if i < len(s3) and s3[i] <= e:
            setup.append((s3[i] - w) / 1e3)
            j = bisect.bisect_left(s4, s3[i])
            if j < len(s4) and s4[j] <= e:
                rowsph.append((s4[j] - s3[i]) / 1e3); flush.append((e - s4[j]) / 1e3)
    for name, v in (("hist set-up", setup), ("hist rows", rowsph), ("hist flush + exit", flush)):
        if v:
            v2 = sorted(v)
            print("%-45s n=%5d  mean %7.2f us  median %7.2f  p90 %7.2f" % (name, len(v), sum(v) / len(v), v2[len(v2) // 2], v2[int(len(v2) * 0.9)]))
