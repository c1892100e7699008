"""Top stall sites of an `ncu -i x.ncu-rep --page source --csv` export (SASS view).
Usage: python tools/ncu_source_top.py <source.csv> [top=25] [context=0]"""
import csv, sys
path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
ctx = int(sys.argv[3]) if len(sys.argv) > 3 else 0
rows = list(csv.reader(open(path, newline="")))
h = rows[1]
ix = {n: i for i, n in enumerate(h)}
body = rows[2:]
stall_cols = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
tot = sum(int(r[ix["# Samples"]]) for r in body)
texec = sum(int(r[ix["Instructions Executed"]]) for r in body)
print(f"{path}: {len(body)} SASS lines, {tot} samples, {texec} warp instructions executed")
agg = {n: sum(int(r[ix[n]]) for r in body) for n in stall_cols}
print("stall totals:", ", ".join(f"{k[6:]} {v * 100 / max(tot, 1):.1f}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v))
order = sorted(range(len(body)), key=lambda i: -int(body[i][ix["# Samples"]]))[:top]
for i in sorted(order):
    for j in range(max(0, i - ctx), min(len(body), i + ctx + 1)):
        r = body[j]
        n = int(r[ix["# Samples"]])
        st = sorted(((int(r[ix[c]]), c[6:]) for c in stall_cols), reverse=True)[:3]
        print(f"{'*' if j == i else ' '} {j:5d} {n:6d} {n * 100 / max(tot, 1):5.1f}%  exec {int(r[ix['Instructions Executed']]):8d}  {r[ix['Source']].strip()[:70]:70s} "
              + " ".join(f"{k}:{v}" for v, k in st if v))
    if ctx:
        print()
