"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump by CUDA source line.
usage: ncu -i rep --page source --csv --print-source cuda,sass > x.csv ; python tools/ncu_lines.py x.csv [N]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
H = rows[hdr]
isamp, iinst, inoti = H.index("# Samples"), H.index("Instructions Executed"), H.index("Warp Stall Sampling (Not-issued Samples)")
stall_cols = [i for i, h in enumerate(H) if h.startswith("stall_")]
agg = collections.OrderedDict(); cur = None; nsass = collections.Counter()
for r in rows[hdr + 1:]:
    if len(r) < len(H): continue
    if r[0] != "" and r[0].isdigit():
        cur = (int(r[0]), r[1].strip()[:100]); agg.setdefault(cur, [0, 0]); continue
    if cur is None or r[2] in ("-", "..."): continue
    try:
        agg[cur][0] += int(r[isamp]); agg[cur][1] += int(r[iinst]); nsass[cur] += 1
    except ValueError: pass
ts = sum(v[0] for v in agg.values()); ti = sum(v[1] for v in agg.values())
print("total samples", ts, "warp-instructions", ti, "sass rows", sum(nsass.values()))
print("--- by samples"); 
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:N]: print(f"{100*v[0]/ts:5.1f}%s {100*v[1]/ti:5.1f}%i sass={nsass[k]:5d} L{k[0]:4d} {k[1]}")
print("--- by static sass size")
for k, c in nsass.most_common(25): print(f"sass={c:5d} L{k[0]:4d} {k[1]}")
