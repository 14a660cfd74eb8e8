"""Instructions and stall samples per CUDA source line from `ncu --page source --print-source=cuda,sass --csv`.
usage: src_hot.py file.csv [top]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 50
cur_file = None; hdr = None
agg = collections.OrderedDict()   # (file,line) -> [src, inst, samples]
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split("/")[-1]; hdr = None; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = {n: i for i, n in enumerate(r)}; src_i = r.index("Source"); continue
    if hdr is None: continue
    ln = r[0]
    if ln:   # a source line row: totals for that line
        try:
            inst = int(r[hdr["Instructions Executed"]] or 0); smp = int(r[hdr["# Samples"]] or 0)
        except ValueError:
            continue
        k = (cur_file, int(ln))
        a = agg.setdefault(k, [r[src_i].strip(), 0, 0]); a[1] += inst; a[2] += smp
ti = sum(a[1] for a in agg.values()); ts = sum(a[2] for a in agg.values())
print("total inst", ti, "samples", ts)
byfile = collections.Counter(); byfile_s = collections.Counter()
for (f, l), a in agg.items(): byfile[f] += a[1]; byfile_s[f] += a[2]
for f, c in byfile.most_common(): print(f"  {f:24s} inst {100*c/ti:5.1f}%  samples {100*byfile_s[f]/ts:5.1f}%")
for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{f:18s}:{l:4d} inst {100*a[1]/ti:5.2f}% smp {100*a[2]/ts:5.2f}%  {a[0][:110]}")
