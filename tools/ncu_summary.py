"""Summaries of Nsight Compute output for profiles/.
  python tools/ncu_summary.py launches <launches.csv> ["header comment"]   # --metrics gpu__time_duration.sum --csv log
  python tools/ncu_summary.py full <raw.csv> [name-regex]                   # ncu -i x.ncu-rep --page raw --csv
"""
import csv, re, sys
from collections import OrderedDict

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem"]


def rows_of(path):
    with open(path, newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    return list(csv.reader(lines))


def short(name):
    name = re.sub(r"\(.*", "", name)
    return name.replace("void ", "").strip()


def launches(path, header=""):
    rows = rows_of(path)
    hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    h = rows[hdr]
    kn, mn, mv = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value")
    agg = OrderedDict()
    for r in rows[hdr + 1:]:
        if len(r) <= mv or r[mn] != "gpu__time_duration.sum":
            continue
        us = float(r[mv].replace(",", "")) / 1e3
        d = agg.setdefault(short(r[kn]), [0, 0.0])
        d[0] += 1; d[1] += us
    tot = sum(d[1] for d in agg.values())
    if header:
        print("# " + header)
    print(f"# total {tot / 1e3:.2f} ms over {sum(d[0] for d in agg.values())} launches (cold-cache, serialised: compare shares)")
    for k, d in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"  {d[1] / 1e3:8.3f} ms {100 * d[1] / tot:5.1f}% n={d[0]:4d} avg={d[1] / d[0]:8.1f} us  {k}")


def full(path, pat=None):
    rows = rows_of(path)
    hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    h = rows[hdr]
    kn = h.index("Kernel Name")
    cols = [(m, h.index(m)) for m in KEEP if m in h]
    extra = [h.index(c) for c in ("Block Size", "Grid Size") if c in h]
    for r in rows[hdr + 2:]:
        if len(r) <= kn or (pat and not re.search(pat, r[kn])):
            continue
        print(f"Kernel Name={short(r[kn])}; " + "; ".join(f"{h[i]}={r[i]}" for i in extra) + "; " +
              "; ".join(f"{m}={r[i]}" for m, i in cols))


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")
    else:
        full(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else None)
