"""Aggregate `ncu --page source --csv` samples by opcode and by 50-instruction region."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
h = next(i for i, r in enumerate(rows) if "Source" in r); hdr = rows[h]
ia, isamp, iex = hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
body = [r for r in rows[h + 1:] if len(r) > iex]
if len(sys.argv) > 2 and sys.argv[2] == "half": body = body[:len(body) // 2]
def num(x):
    try: return int(x)
    except Exception: return 0
tot = sum(num(r[isamp]) for r in body) or 1
byop, exop = collections.Counter(), collections.Counter()
for r in body:
    op = [o for o in r[ia].strip().split() if not o.startswith('@')]
    op = op[0] if op else ''
    byop[op] += num(r[isamp]); exop[op] += num(r[iex])
print("samples", tot, "instructions", len(body), "warp-inst", sum(exop.values()))
for op, s in byop.most_common(18): print(f"{op:36s} {s:8d} {100*s/tot:5.1f}%  ex={exop[op]}")
acc = collections.Counter()
for i, r in enumerate(body): acc[i // 50] += num(r[isamp])
print({k * 50: v for k, v in sorted(acc.items()) if v > tot * 0.01})
