"""Summarise an ncu source-page CSV: stall totals and the hottest SASS lines."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = {s: 0 for s in stalls}
lines = []
for r in rows[2:]:
    if len(r) < len(hdr) or not r[ix["# Samples"]].strip().isdigit(): continue
    n = int(r[ix["# Samples"]] or 0)
    for s in stalls: tot[s] += int(r[ix[s]] or 0)
    lines.append((n, r[ix["Source"]], {s: int(r[ix[s]] or 0) for s in stalls}))
T = sum(tot.values())
print("total samples", T)
for s, v in sorted(tot.items(), key=lambda kv: -kv[1])[:8]: print(f"  {s:28s} {v:8d} {100*v/max(T,1):5.1f}%")
print("hottest instructions:")
for n, src, d in sorted(lines, key=lambda x: -x[0])[:int(sys.argv[2]) if len(sys.argv) > 2 else 14]:
    top = max(d.items(), key=lambda kv: kv[1])
    print(f"  {n:7d} {top[0]:18s} {src[:90]}")
