"""Per-shape comparison of tools/shape_breakdown.py outputs of two library builds run alternately on one box.
Usage: python tools/ab_breakdown.py <A1.txt> <B1.txt> [<A2.txt> <B2.txt> ...]   (minimum per row over the runs of each side)"""
import re, sys


def load(f):
    d = {}
    for l in open(f):
        m = re.match(r"\s+([0-9.]+) ms\s+[0-9.]+% n=\s*(\d+)\s+[0-9.]+ TF/s\s+(\S+)\s+(.*?)\s{2,}(\S.*)?$", l)
        if m:
            d[(m.group(3), re.sub(r"\s+", " ", m.group(4).strip()))] = float(m.group(1))   # key: op + shape (kernel name may differ)
    return d


files = sys.argv[1:]
A = [load(f) for f in files[0::2]]
B = [load(f) for f in files[1::2]]
keys = set(A[0]) & set(B[0])
rows = []
for k in keys:
    a = min(d.get(k, 1e9) for d in A)
    b = min(d.get(k, 1e9) for d in B)
    rows.append((b - a, k, a, b))
rows.sort()
ta, tb = sum(r[2] for r in rows), sum(r[3] for r in rows)
print(f"common rows: A {ta:.2f} ms  B {tb:.2f} ms  ({100 * (tb / ta - 1):+.2f} %)")
for d, k, a, b in rows[:12] + rows[-8:]:
    print(f"{d:+.3f} ms  {a:7.3f} -> {b:7.3f}  {k[0]} {k[1]}")
