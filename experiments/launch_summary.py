"""Summarise one boosting iteration out of an ncu launch list (--metrics gpu__time_duration.sum --csv): kernels between the
second-to-last and the last k_bt_init of the capture, by kernel, in launch order."""
import csv, re, sys, collections
rows = []
for r in csv.reader(open(sys.argv[1])):
    if len(r) > 14 and r[0].isdigit():
        name = re.sub(r"^(void )?lgb::", "", r[4]).split("(")[0]
        rows.append((name, float(r[14])))
inits = [i for i, (n, _) in enumerate(rows) if n.startswith("k_bt_init")]
a, b = inits[-2], inits[-1]
# the iteration's quantize / scales launches sit in front of its k_bt_init
while a > 0 and rows[a - 1][0].startswith(("k_quantize", "k_scales")): a -= 1
while b > 0 and rows[b - 1][0].startswith(("k_quantize", "k_scales")): b -= 1
agg = collections.OrderedDict()
for n, t in rows[a:b]:
    c = agg.setdefault(n, [0, 0.0]); c[0] += 1; c[1] += t
tot = sum(v[1] for v in agg.values())
for n, (c, t) in agg.items():
    print("%-44s %4d %9.1f us %6.1f%%" % (n, c, t / 1e3, 100 * t / tot))
print("total %.1f us (one boosting iteration of the capture; ncu-serialised, cold caches)" % (tot / 1e3))
