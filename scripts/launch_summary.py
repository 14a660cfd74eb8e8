"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) per kernel. usage: launch_summary.py launches.csv"""
import collections, csv, re, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
h = rows[0]; ki, vi = h.index("Kernel Name"), h.index("Metric Value")
t = collections.defaultdict(float); c = collections.Counter()
for r in rows[1:]:
    name = re.sub(r"\(.*", "", r[ki]).replace("void ", "").replace("vsann::", "")
    t[name] += float(r[vi]) / 1e6; c[name] += 1
tot = sum(t.values())
print(f"{'ms':>10s} {'share':>6s} {'launches':>8s}  kernel")
for k, v in sorted(t.items(), key=lambda kv: -kv[1]):
    print(f"{v:10.2f} {100 * v / tot:5.1f}% {c[k]:8d}  {k}")
print(f"total {tot:.1f} ms over {sum(c.values())} launches")
