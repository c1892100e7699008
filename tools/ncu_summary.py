"""Summaries of Nsight Compute output for profiles/.
  python tools/ncu_summary.py launches <launches.csv> ["header comment"]   # --metrics gpu__time_duration.sum --csv log
  python tools/ncu_summary.py full <raw.csv> [name-regex]                   # ncu -i x.ncu-rep --page raw --csv
  python tools/ncu_summary.py json <out.json> "<source note>" <raw.csv>...  # raw pages exported with --print-units base
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
    name = re.sub(r"\((int|bool|unsigned int)\)", "", name)   # negative template arguments keep their cast: "(int)-5"
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


def to_json(out, source, paths):
    """Per kernel template instance: mean over the captured launches of duration, DRAM bytes, pipe utilisations (the
    file bench.py reads `roofline.traffic` from).  Raw pages must be in base units (bytes, ns, %)."""
    import json
    want = {"duration_ns": "gpu__time_duration.sum", "dram_read": "dram__bytes_read.sum", "dram_write": "dram__bytes_write.sum",
            "tensor_pipe_pct": "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
            "dram_pct": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "xu_pct": "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
            "issue_pct": "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
            "l2_pct": "lts__throughput.avg.pct_of_peak_sustained_elapsed",
            "warps_active_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
            "registers": "launch__registers_per_thread"}
    launches_ = []
    for path in paths:
        rows = rows_of(path)
        hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
        h = rows[hdr]
        kn = h.index("Kernel Name")
        gi = h.index("Grid Size") if "Grid Size" in h else None
        for r in rows[hdr + 2:]:
            if len(r) <= kn:
                continue
            d = {"kernel": short(r[kn]).replace("topo::", ""), "grid": r[gi] if gi is not None else ""}
            for k, m in want.items():
                if m in h:
                    try:
                        d[k] = float(r[h.index(m)].replace(",", ""))
                    except ValueError:
                        pass
            launches_.append(d)
    kernels = {}
    for d in launches_:
        kernels.setdefault(d["kernel"], []).append(d)
    summ = {}
    for k, ls in kernels.items():
        def mean(key):
            v = [l[key] for l in ls if key in l]
            return sum(v) / len(v) if v else None
        summ[k] = {"launches": len(ls), "duration_us": round(mean("duration_ns") / 1e3, 2),
                   "dram_bytes_per_launch": round(mean("dram_read") + mean("dram_write")),
                   "tensor_pipe_pct": round(mean("tensor_pipe_pct"), 2), "dram_pct": round(mean("dram_pct"), 2),
                   "xu_pct": round(mean("xu_pct"), 2), "issue_pct": round(mean("issue_pct"), 2),
                   "l2_pct": round(mean("l2_pct"), 2), "warps_active_pct": round(mean("warps_active_pct"), 2),
                   "registers": mean("registers"), "grids": sorted({l["grid"] for l in ls})}
    json.dump({"source": source, "kernels": summ, "launches": launches_}, open(out, "w"), indent=1)
    for k, v in sorted(summ.items(), key=lambda kv: -kv[1]["duration_us"] * kv[1]["launches"]):
        print(f"{k:44s} n={v['launches']:2d} {v['duration_us']:9.1f} us  tensor {v['tensor_pipe_pct']:5.1f}%  dram {v['dram_pct']:5.1f}%  "
              f"xu {v['xu_pct']:5.1f}%  issue {v['issue_pct']:5.1f}%  {v['dram_bytes_per_launch'] / 1e6:8.1f} MB/launch  regs {v['registers']:.0f}")


if __name__ == "__main__":
    if sys.argv[1] == "json":
        to_json(sys.argv[2], sys.argv[3], sys.argv[4:])
    elif sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")
    else:
        full(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else None)
