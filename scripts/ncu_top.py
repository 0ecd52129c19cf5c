"""Top stall sites from `ncu --page source --csv` output: python scripts/ncu_top.py file.csv [n]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
h = next(i for i, r in enumerate(rows) if "Source" in r)
hdr = rows[h]
ia, isamp, iex = hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
def num(x):
    try: return int(x)
    except Exception: return 0
data = [(num(r[isamp]), num(r[iex]), i, r[ia]) for i, r in enumerate(rows[h + 1:]) if len(r) > iex]
tot = sum(d[0] for d in data) or 1
print("total samples", tot, "total warp-inst", sum(d[1] for d in data))
for s, e, i, src in sorted(data, reverse=True)[:n]:
    print(f"{s:8d} {100*s/tot:5.1f}% ex={e:10d} #{i:5d} {src.strip()[:100]}")
