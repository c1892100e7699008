"""SASS evidence for profiles/: per kernel of libtopo_b200.so, the count of Blackwell-native mnemonics (B200_PROFILING.md:
tcgen05.mma -> UTCHMMA, tcgen05.ld/st -> LDTM/STTM, TMA -> UTMALDG/UTMASTG) plus the packed-FP32 and MUFU counts that the
softmax / GELU notes in DESIGN.md refer to.  Usage: python tools/sass_summary.py > profiles/r02_sass.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "interactive-stylization-of-topographic-maps-using-prompt-driven-diffusion-models_b200", "libtopo_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
KEYS = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTCBAR", "SYNCS", "FFMA2", "FADD2", "FMUL2", "MUFU.EX2", "MUFU.RCP", "HMMA"]
cur, counts, order = None, collections.defaultdict(collections.Counter), []
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(.*", "", cur).replace("void topo::", "").replace("(int)", "").replace("(bool)", "")
        order.append(cur)
        continue
    if cur is None:
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if not m:
        continue
    op = m.group(1)
    for k in KEYS:
        if op == k or op.startswith(k + ".") or (k in ("UTCHMMA.2CTA",) and op.startswith("UTCHMMA") and ".2CTA" in op):
            counts[cur][k] += 1
print(f"# cuobjdump -sass {os.path.relpath(lib, ROOT)} : static instruction counts per kernel (sm_100a); HMMA = legacy mma.sync path (must be 0)")
print(f"# {'kernel':58s} " + " ".join(f"{k:>12s}" for k in KEYS))
tot = collections.Counter()
for name in order:
    c = counts[name]
    if not any(c[k] for k in ("UTCHMMA", "LDTM", "UTMALDG", "UTMASTG")):
        continue
    print(f"  {name[:58]:58s} " + " ".join(f"{c[k]:12d}" for k in KEYS))
    tot.update(c)
print(f"  {'TOTAL (tensor-core / TMA kernels)':58s} " + " ".join(f"{tot[k]:12d}" for k in KEYS))
print(f"# kernels in the library: {len(order)}; with UTCHMMA: {sum(1 for n in order if counts[n]['UTCHMMA'])}")
