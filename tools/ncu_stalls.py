"""Per-instruction warp-stall summary of one kernel from `ncu -i X.ncu-rep --page source --csv --print-source sass`."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = {s: 0 for s in stalls}
data = []
def num(x):
    try: return int(float(x))
    except ValueError: return 0
for r in rows[hi + 1:]:
    if len(r) < len(hdr) or r[0] == "Address": continue
    n = num(r[ix["# Samples"]])
    data.append((n, r))
    for s in stalls: tot[s] += num(r[ix[s]])
T = sum(n for n, _ in data) or 1
print("total samples", T, "instructions", len(data))
for s, v in sorted(tot.items(), key=lambda x: -x[1])[:12]:
    print(f"  {s:28s} {v:7d} {100*v/T:5.1f}%")
print("top instructions:")
for n, r in sorted(data, key=lambda x: -x[0])[:topn]:
    top = sorted(((num(r[ix[s]]), s) for s in stalls), reverse=True)[:2]
    print(f"{n:6d} {100*n/T:4.1f}% {r[ix['Address']][-5:]} {r[ix['Source']][:72]:72s} {top[0][1][6:]}:{top[0][0]} {top[1][1][6:]}:{top[1][0]}")
