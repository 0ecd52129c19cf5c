"""Stall-reason totals (and top instructions per reason) from `ncu --page source --csv`."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
h = next(i for i, r in enumerate(rows) if "Source" in r); hdr = rows[h]
body = [r for r in rows[h + 1:] if len(r) == len(hdr)]
reasons = [c for c in hdr if c.startswith("stall_") and "Not Issued" not in c]
ia = hdr.index("Source")
def num(x):
    try: return int(x)
    except Exception: return 0
tot = collections.Counter()
for r in body:
    for c in reasons: tot[c] += num(r[hdr.index(c)])
allc = sum(tot.values()) or 1
for c, v in tot.most_common(8):
    print(f"{c:24s} {v:8d} {100*v/allc:5.1f}%")
    ic = hdr.index(c)
    top = sorted(((num(r[ic]), i, r[ia].strip()[:70]) for i, r in enumerate(body)), reverse=True)[:4]
    for v2, i, src in top: print(f"      {v2:7d} #{i:5d} {src}")
