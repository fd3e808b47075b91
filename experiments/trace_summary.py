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
    if k == 4 and ph >= 3:           # replay kernel: 43 loaded, 44 replayed, 45 bookkeeping done, 46 ranked, 47 batch built
        sub.setdefault(40 + ph, []).append(t)
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

# replay kernel phases (thread 0 of its one CTA): deps done -> 43 -> 44 -> 45 -> 46 -> 47 -> end
sel = [(w, e) for (n, s_, w, e) in [(r[0], r[1], r[2], r[3]) for r in rows] if n == "select"]
if sub.get(43):
    marks = {p: sorted(sub.get(p, [])) for p in (43, 44, 45, 46, 47)}
    names4 = {43: "select: reduce + load leaves", 44: "select: replay order", 45: "select: tree bookkeeping", 46: "select: write back + rank",
              47: "select: build batch"}
    spans = collections.defaultdict(list)
    # rows hold times relative to t0 in us; marks are absolute ns
    for w, e in sel:
        prev = w * 1e3 + t0
        end = e * 1e3 + t0
        for p in (43, 44, 45, 46, 47):
            i = bisect.bisect_left(marks[p], prev)
            if i < len(marks[p]) and marks[p][i] <= end:
                spans[names4[p]].append((marks[p][i] - prev) / 1e3)
                prev = marks[p][i]
        spans["select: exit"].append((end - prev) / 1e3)
    for name in list(names4.values()) + ["select: exit"]:
        v = spans.get(name)
        if v:
            v2 = sorted(v)
            print("%-45s n=%5d  mean %7.2f us  median %7.2f  p90 %7.2f" % (name, len(v), sum(v) / len(v), v2[len(v2) // 2], v2[int(len(v2) * 0.9)]))
