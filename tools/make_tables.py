"""Markdown tables for DESIGN.md §5/§6 from the bench lines under profiles/ (python tools/make_tables.py r02z)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r02z"
P = os.path.join(ROOT, "profiles")


def load(name):
    p = os.path.join(P, name)
    if not os.path.exists(p):
        return None
    for line in open(p):
        line = line.strip()
        if line.startswith("{"):
            return json.loads(line)
    return None


d = load(f"{tag}_bench_1gpu.json")
if d:
    r = d["roofline"]
    print(f"`bench.py` on a fresh B200 (`profiles/{tag}_bench_1gpu.json`): **{d['value']:.2f} tiles/s** device-resident, **{d['e2e']['value']:.2f} tiles/s** "
          f"end to end (pinned host in/out), {d['ms_per_step']:.0f} ms per 19-tile step, SM {d['clocks']['sm_mhz']:.0f} of {d['clocks']['sm_max_mhz']:.0f} MHz "
          f"({', '.join(d['clocks']['reasons']) or 'no throttle reason'}), {d['gpu_launches']} kernel launches in the timed region; whole job "
          f"{d['end_to_end_roofline']['achieved_tflops']:.0f} TFLOP/s = {100 * d['end_to_end_roofline']['frac_of_sustained_bf16']:.1f} % of the measured sustained bf16 rate.")
    if d.get("cpu_baseline"):
        print(f"CPU arm (oracle port, {d['cpu_baseline']['cores']} host threads): {d['cpu_baseline']['value']:.4f} tiles/s.")
    print()
    mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    ncu_all = json.load(open(os.path.join(P, "r03_ncu.json")))["kernels"]
    print("Per kernel template instance, one instrumented eager `apply_model` (Be = 38), CUDA events per launch; `frac` = achieved ÷ measured peak "
          f"(sustained bf16 {mp['bf16_tflops_sustained']:.0f} TFLOP/s or copy {mp['hbm_gbs']:.0f} GB/s, MEASURED_PEAKS.json); ncu columns from "
          "`profiles/r03_ncu.json` (`ncu --set full` of the same build):\n")
    print("| kernel | ops | launches | ms | share | achieved | frac | ncu tensor-pipe % | ncu DRAM % | ncu DRAM MB/launch |")
    print("|---|---|---:|---:|---:|---:|---:|---:|---:|---:|")
    tot = sum(v["ms"] for v in r["by_kernel"].values())
    for k, v in r["by_kernel"].items():
        if v["ms"] < 0.12:
            continue
        n = ncu_all.get(k, {})
        print(f"| `{k}` | {', '.join(v['ops'])} | {v['launches']} | {v['ms']:.2f} | {100 * v['ms'] / tot:.1f} % | {v['achieved']:.0f} {v['unit']} | {v['frac']:.2f} | "
              f"{n.get('tensor_pipe_pct', '—')} | {n.get('dram_pct', '—')} | {('%.0f' % (n['dram_bytes_per_launch'] / 1e6)) if n else '—'} |")
    print(f"\nBy op class: " + "; ".join(f"{k} {v['ms']:.2f} ms ({v['achieved']:.0f} {v['unit']}, {v['frac']:.2f})" for k, v in r["by_op"].items() if v["ms"] > 0.3) + ".")
print()
rows = []
for w in ("addlrSD", "qkvo", "fullmap"):
    x = load(f"{tag}_bench_{w}.json") or load(f"{tag}_bench_{w}_1gpu.json")
    if x:
        rows.append((w, x))
if rows:
    print("| workload (1 GPU) | tiles/s | e2e tiles/s | ms/step | tiles/step | % of sustained bf16 | gemm_lora TFLOP/s | attn_self | attn_cross GB/s |")
    print("|---|---:|---:|---:|---:|---:|---:|---:|---:|")
    for w, x in rows:
        bo = x["roofline"]["by_op"]
        print(f"| {w} | {x['value']:.2f} | {x['e2e']['value']:.2f} | {x['ms_per_step']:.0f} | {x['config']['tiles_per_step_per_gpu']} | "
              f"{100 * x['end_to_end_roofline']['frac_of_sustained_bf16']:.1f} | {bo.get('gemm_lora', {}).get('achieved', 0):.0f} | {bo['attn_self']['achieved']:.0f} | {bo['attn_cross']['gbs']:.0f} |")
print()
print("| GPUs | weak: tiles/s (19 tiles per rank) | ms/step | strong: tiles/s (one 256-tile map) | s per map | speed-up |")
print("|---:|---:|---:|---:|---:|---:|")
base = None
for n in (1, 2, 4, 8):
    wk = load(f"{tag}_bench_{n}gpu.json")
    st = load(f"{tag}_bench_fullmap_{n}gpu.json")
    if not (wk or st):
        continue
    if st and base is None:
        base = st["ms_per_step"] * (st["n_gpus"])  # normalise to 1 GPU if the first point is not N=1
        base = st["ms_per_step"] if st["n_gpus"] == 1 else None
    sp = f"{base / st['ms_per_step']:.2f}×" if (st and base) else "—"
    print(f"| {n} | {wk['value']:.2f} | {wk['ms_per_step']:.0f} |" if wk else f"| {n} | — | — |", end="")
    print(f" {st['value']:.2f} | {st['ms_per_step'] / 1e3:.2f} | {sp} |" if st else " — | — | — |")
