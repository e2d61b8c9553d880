"""Summarises an `ncu --page source --csv --print-source sass` export: executed warp-instructions and stall samples per
opcode, and the hottest SASS lines (development helper for profiles/)."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
ops = collections.defaultdict(lambda: [0, 0, collections.Counter()])
tot_i = tot_s = 0
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
lines = []
for r in rows[2:]:
    if len(r) < len(hdr): continue
    src = r[ix["Source"]].strip()
    toks = src.split()
    op = toks[0]
    if op.startswith("@"): op = toks[1]
    op = op.rstrip(";")
    n = int(r[ix["Instructions Executed"]]); s = int(r[ix["# Samples"]])
    ops[op][0] += n; ops[op][1] += s
    for c in stall_cols:
        v = int(r[ix[c]] or 0)
        if v: ops[op][2][c] += v
    tot_i += n; tot_s += s
    lines.append((s, n, r[ix["Address"]][-5:], src, {c: int(r[ix[c]] or 0) for c in stall_cols if int(r[ix[c]] or 0)}))
print(f"total warp-instructions {tot_i}, samples {tot_s}")
print(f"{'opcode':28s} {'inst%':>7s} {'samples%':>9s}  top stalls")
for op, (n, s, st) in sorted(ops.items(), key=lambda kv: -kv[1][0])[:int(sys.argv[2]) if len(sys.argv) > 2 else 30]:
    print(f"{op:28s} {100*n/tot_i:7.2f} {100*s/tot_s:9.2f}  " + ", ".join(f"{k[6:]}={v}" for k, v in st.most_common(4)))
allst = collections.Counter()
for op, (n, s, st) in ops.items(): allst.update(st)
print("stall totals:", ", ".join(f"{k[6:]}={100*v/tot_s:.1f}%" for k, v in allst.most_common(12)))
if len(sys.argv) > 3:
    print("hottest lines:")
    for s, n, a, src, st in sorted(lines, key=lambda t: -t[0])[:int(sys.argv[3])]:
        print(f"  {a} {100*s/tot_s:5.2f}% n={n} {src[:60]:60s} {st}")
