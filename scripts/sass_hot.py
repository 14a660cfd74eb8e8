"""Top SASS instructions by stall samples from `ncu --page source --csv --print-source=sass`. usage: sass_hot.py file.csv [kernel-substr] [top]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
pat = sys.argv[2] if len(sys.argv) > 2 else ""
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
# split per kernel
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name": cur = {"name": r[1], "hdr": None, "rows": []}; blocks.append(cur); continue
    if cur is None: continue
    if cur["hdr"] is None: cur["hdr"] = r; continue
    cur["rows"].append(r)
for b in blocks:
    if pat not in b["name"]: continue
    h = {n: i for i, n in enumerate(b["hdr"])}
    si, ii = h["# Samples"], h["Instructions Executed"]
    tot_s = sum(int(r[si] or 0) for r in b["rows"]); tot_i = sum(int(r[ii] or 0) for r in b["rows"])
    print(b["name"][:100], "samples", tot_s, "inst", tot_i, "n_sass", len(b["rows"]))
    ops = collections.Counter()
    for r in b["rows"]:
        op = r[h["Source"]].split()[0] if r[h["Source"]] else "?"
        if op.startswith("@"): op = r[h["Source"]].split()[1]
        ops[op.split(".")[0]] += int(r[ii] or 0)
    print("  inst by opcode:", ", ".join(f"{o}:{100*c/tot_i:.1f}%" for o, c in ops.most_common(18)))
    idx = sorted(range(len(b["rows"])), key=lambda i: -int(b["rows"][i][si] or 0))[:top]
    for i in sorted(idx):
        r = b["rows"][i]
        stalls = {k[6:]: int(r[h[k]] or 0) for k in h if k.startswith("stall_") and "Not Issued" not in k}
        top2 = sorted(stalls.items(), key=lambda kv: -kv[1])[:2]
        print(f"  #{i:5d} s={100*int(r[si] or 0)/tot_s:5.2f}% n={int(r[ii] or 0):9d} thr={r[h['Avg. Threads Executed']]:>5s} {r[h['Source']][:70]:70s} {top2}")
    break
