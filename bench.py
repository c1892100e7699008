#!/usr/bin/env python
"""bench.py — 512x512 map tiles/s of the ControlNet+LoRA SD-1.5 stylization hot path (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W                 # this repo's sm_100a path, default workload
  python bench.py --workload addlrSD|qkvo|fullmap ...           # the other BASELINE configs (not the driver's line)
  python bench.py --impl reference --steps K --warmup W         # the reference's CPU path (oracle port) on host cores

Default workload = BASELINE.json configs[1] ("ctrlora": ControlNet qkvo LoRA rank 16, 19-prompt multiseg batch = 19 tiles
of 512x512, nominal 30 DDIM steps = 31 model calls, classifier-free guidance 7.5 with the pair batched, VAE decode, uint8
conversion).  One "step" = one pass of the hot path over one batch of synthetic tiles.  N > 1: every rank stylizes its own
tiles (weak scaling, no collective in the denoising loop) and the finished uint8 tiles are all-gathered once per step.
`--workload fullmap` is configs[4]: ONE 16x16 mosaic (256 tiles) sharded over the N ranks in contiguous row blocks, one
all-gather, stitched on every rank — strong scaling.  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [ROOT]

METRIC = "512x512 map tiles/sec (30-step CN+LoRA SD1.5)"
UNIT = "tiles/s"
STEPS_DDIM = 30
# algorithmic work, SURVEY.md §8(d): GFLOP per effective-batch element per apply_model with step-invariant work hoisted
GFLOP_HOIST, GFLOP_VAE = 14.72 + 4.31, 2514.5
WORKLOADS = {
    # name: variant, model overrides, tiles per step per GPU, CFG scale, GFLOP per element-call (None: counted live), text
    "ctrlora": dict(variant="ctrlora", kw=dict(control_lora_rank=16), tiles=19, scale=7.5, gflop=1069.4,
                    text="ctrlora (configs[1]): random-init SD1.5-arch UNet+ControlNet, ControlNet q/k/v/out LoRA rank 16, "
                         "19 tiles 512x512 per step, DDIM S=30 (31 model calls), CFG 7.5 batched [uncond;cond], VAE decode, uint8"),
    "addlrSD": dict(variant="addlrSD", kw={}, tiles=8, scale=7.5, gflop=1072.8,
                    text="addlrSD (configs[2]): ControlNet qkvo LoRA rank 32 + UNet lora_qkv(r_q=8, r_kv=8) (96 wraps, s=2), "
                         "8 tiles 512x512 per step, DDIM S=30 (31 model calls), CFG 7.5 batched, VAE decode, uint8"),
    "qkvo": dict(variant="qkvo", kw={}, tiles=8, scale=5.0, gflop=None,
                 text="qkvo (configs[3]): stock ControlNet + UNet lora_attn_qkvo(r=8, s=4) (128 wraps), 8 tiles 512x512 per "
                      "step, DDIM S=30 (31 model calls), CFG 5.0 batched, VAE decode, uint8"),
    "fullmap": dict(variant="ctrlora", kw=dict(control_lora_rank=16), tiles=256, scale=7.5, gflop=1069.4,
                    text="full map (configs[4]): one 8192x8192 sheet = 16x16 mosaic of 256 tiles 512x512, ctrlora rank 16, DDIM "
                         "S=30 (31 model calls), CFG 7.5 batched, per-tile x_T (seed 1234 + tile index), tiles sharded over the "
                         "ranks in contiguous row blocks (<= 32 tiles = Be 64 per batch), ONE all-gather, stitch"),
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(src="measured", hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sust=d["bf16_tflops_sustained"])
    return dict(src="fallback", hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0)


def ncu_table():
    """profiles/r03_ncu.json (tools/ncu_named.sh + tools/ncu_summary.py json): per kernel template instance, the mean
    of `ncu --set full` metrics over the captured launches of the FINAL build.  None when the file is absent."""
    p = os.path.join(ROOT, "profiles", "r03_ncu.json")
    return json.load(open(p)) if os.path.exists(p) else None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "200"], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        self.p.wait()
        self.f.flush()
        rows = [r.split(",") for r in open(self.f.name).read().strip().splitlines() if r.count(",") >= 6]
        os.unlink(self.f.name)
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(r[0]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].strip() == "Active" for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][1]), "power_w_max": max(float(r[2]) for r in rows),
                "samples": len(rows), "reasons": reasons}


# ----------------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the oracle (CPU port of the reference's PyTorch path) on the host cores
# ----------------------------------------------------------------------------------------------------------
def cpu_reference(wl: dict, n_calls: int, warm: int):
    """Times ``n_calls`` apply_model calls at Be=2 (the CFG pair of ONE tile, 64x64 latent) + one VAE decode with the
    fp32 oracle on all host threads and extrapolates: tile time = 31 * t_call + t_vae (+ hint/K-V not hoisted: the
    reference recomputes them every call, and so does the oracle)."""
    import torch
    from oracle import ldm_oracle as O
    from topo_b200 import synth
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    model = synth.build_model(wl["variant"], **wl["kw"])
    sd = {k: v.float() for k, v in model.state_dict().items()}
    scales = synth.lora_scales(model)
    del model
    hint, ctx, xT = synth.make_hint(1), synth.make_ctx(1), synth.make_xT(1)
    cond = {"c_crossattn": [torch.cat([torch.zeros_like(ctx), ctx])], "c_concat": [torch.cat([hint, hint])]}
    x2, t2 = torch.cat([xT, xT]), torch.full((2,), 501, dtype=torch.long)
    times = []
    with torch.no_grad():
        for i in range(warm + n_calls):
            t0 = time.perf_counter()
            O.apply_model(sd, x2, t2, cond, heads=8, scales=scales)
            if i >= warm:
                times.append(time.perf_counter() - t0)
        t0 = time.perf_counter()
        O.decode_first_stage(sd, xT)
        t_vae = time.perf_counter() - t0
    t_call = sum(times) / len(times)
    t_tile = 31 * t_call + t_vae
    return dict(value=1.0 / t_tile, t_call=t_call, t_vae=t_vae, cores=cores, calls=n_calls,
                sample=f"{n_calls} apply_model calls at Be=2 (one tile's CFG pair, 64x64 latent, fp32, {cores} threads; "
                       f"{t_call:.2f} s/call) + 1 VAE decode ({t_vae:.2f} s); tile time extrapolated as 31*t_call + t_vae")


def config_of(args, wl, world=1):
    """The `config` object both arms print (same keys and values, so the driver can compare them)."""
    full = args.workload == "fullmap"
    return {"workload": wl["text"], "tiles_per_step_per_gpu": (wl["tiles"] // world) if full else wl["tiles"],
            "model_calls_per_tile": 31, "effective_batch": 2 * (min(32, wl["tiles"] // world) if full else wl["tiles"]),
            "l2": "working set per step (2.4 GB weights + activations) exceeds the 126 MB L2; no explicit flush",
            "cuda_graphs": True,
            "gather": ("all_gather of uint8 tiles once per step" if (world > 1 or full) else "none (1 GPU)")}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = WORKLOADS[args.workload]
    r = cpu_reference(wl, max(1, args.steps), max(0, min(args.warmup, 1)))
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * r["t_call"], "higher_is_better": True,
            "scaling": "strong" if args.workload == "fullmap" else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": config_of(args, wl, max(1, args.gpus)),
            "reference_note": "reference = oracle/ldm_oracle.py, the CPU fp32 restatement of the reference's PyTorch path (its "
                              "own code is not importable: empty ControlNet/ctrlora submodules); one CPU process, all host threads",
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------------------
# this repo's arm
# ----------------------------------------------------------------------------------------------------------
def roofline_leg(E, eng, model, pk, ctx_d, hints_d, xT_d):
    """One extra eager apply_model at the bench batch with every launch bracketed by CUDA events on the launching stream
    (after the timed region, so the headline number carries no event overhead), keyed by kernel template instance."""
    import torch
    dev = ctx_d.device
    B = ctx_d.shape[0]
    c2 = torch.cat([torch.zeros_like(ctx_d), ctx_d])
    h2 = torch.cat([hints_d, hints_d])
    x2 = torch.cat([xT_d, xT_d])
    t2 = torch.full((2 * B,), 501.0, device=dev)
    P = eng.prepare(c2, hints_d)              # as the sampler does: one hint per tile, context-free prefix shared
    for _ in range(2):
        eng.apply_model(x2, t2, c2, h2, model.control_scales, False, prepared=P, x_rep=2)
    torch.cuda.synchronize()
    E.PROF.recs, E.PROF.on = [], True
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    eng.apply_model(x2, t2, c2, h2, model.control_scales, False, prepared=P, x_rep=2)
    t1.record()
    torch.cuda.synchronize()
    E.PROF.on = False
    ncu = ncu_table() or {}

    def row(v, name=None):
        tensor = v["flops"] > 0
        ach = (v["flops"] / 1e12 if tensor else v["bytes"] / 1e9) / (v["ms"] / 1e3)
        peak = pk["tf_sust"] if tensor else pk["hbm"]
        r = {"ms": round(v["ms"], 3), "launches": v["launches"], "bound": "tensor" if tensor else "hbm",
             "achieved": round(ach, 1), "unit": "TFLOP/s" if tensor else "GB/s", "frac": round(ach / peak, 4),
             "gbs": round(v["bytes"] / 1e6 / max(v["ms"], 1e-9), 1), "ops": sorted(v["kinds"])}
        n = ncu.get("kernels", {}).get(name or "")
        if n:
            r["ncu"] = {k: n[k] for k in ("launches", "dram_bytes_per_launch", "tensor_pipe_pct", "dram_pct", "duration_us") if k in n}
        return r

    by_kernel = E.prof_summary(by="kernel")
    by_kind = E.prof_summary(by="kind")
    total = sum(d["ms"] for d in by_kernel.values())
    top = max(by_kernel, key=lambda k: by_kernel[k]["ms"])
    d = by_kernel[top]
    r = row(d, top)
    flops_all = sum(v["flops"] for v in by_kernel.values())
    nk = ncu.get("kernels", {}).get(top)
    roof = {"bound": r["bound"], "kernel": top, "achieved": r["achieved"], "peak": pk["tf_sust"] if r["bound"] == "tensor" else pk["hbm"],
            "unit": r["unit"], "frac": r["frac"],
            "traffic": nk["dram_bytes_per_launch"] if nk else None,
            "traffic_source": (f"profiles/r03_ncu.json ({ncu.get('source', 'ncu --set full')}): mean dram__bytes_read.sum + "
                               f"dram__bytes_write.sum over {nk['launches']} captured launches of this kernel") if nk else
                              "no ncu capture of this kernel in profiles/r03_ncu.json",
            "peak_source": pk["src"] + (" (sustained bf16)" if r["bound"] == "tensor" else " (copy)"),
            "launches": d["launches"], "avg_launch_us": 1e3 * d["ms"] / d["launches"],
            "algorithmic_per_launch": (d["flops"] if r["bound"] == "tensor" else d["bytes"]) / d["launches"],
            "share_of_apply_model": d["ms"] / total,
            "by_kernel": {k: row(v, k) for k, v in sorted(by_kernel.items(), key=lambda kv: -kv[1]["ms"])},
            "by_op": {k: row(v) for k, v in sorted(by_kind.items(), key=lambda kv: -kv[1]["ms"])},
            "how": f"one eager apply_model (Be={2 * B}) after the timed region, every launch bracketed by CUDA events on the "
                   "launching stream and keyed by the kernel template instance the C ABI reports (topo_last_kernel); achieved "
                   "= algorithmic FLOPs (true head dim / true LoRA rank) or bytes / summed duration; peak = MEASURED_PEAKS",
            "apply_model_ms_eager_instrumented": t0.elapsed_time(t1)}
    return roof, flops_all / (2 * B) / 1e9


def run_ours(args):
    import torch
    import torch.distributed as dist
    from topo_b200 import engine as E
    from topo_b200 import synth, tiles

    wl = WORKLOADS[args.workload]
    full = args.workload == "fullmap"
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the sm_100a path has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    pk = peaks()

    model = synth.build_model(wl["variant"], **wl["kw"]).to(dev)
    eng = model.engine()
    T = wl["tiles"]
    if full:
        # one synthetic 8192^2 sheet, tiled row-major as BaseModel/datasetCreation00.py:32-59; this rank keeps its shard
        lo, hi = tiles.shard_range(T, rank, world)
        sheet = synth.make_hint(1, 8192, seed=7)[0].permute(1, 2, 0).contiguous()
        hints_h = tiles.split_sheet(sheet, 512)[lo:hi].permute(0, 3, 1, 2).contiguous().pin_memory()
        del sheet
        ctx_h = synth.make_ctx(1, seed=11).expand(hi - lo, -1, -1).contiguous().pin_memory()    # one prompt for the whole sheet
        xT_h = None                                                                              # per-tile seed 1234 + global index
        n_local, batch = hi - lo, min(32, hi - lo)
    else:
        hints_h = synth.make_hint(T, 512, seed=7 + 1000 * rank).pin_memory()
        ctx_h = synth.make_ctx(T, seed=11 + 1000 * rank).pin_memory()
        xT_h = synth.make_xT(T, 64, seed=1234 + rank).pin_memory()
        n_local, batch = T, T
    out_h = torch.empty(n_local, 512, 512, 3, dtype=torch.uint8).pin_memory()
    hints_d, ctx_d = hints_h.to(dev), ctx_h.to(dev)
    xT_d = None if xT_h is None else xT_h.to(dev)
    kw = dict(steps=STEPS_DDIM, scale=wl["scale"], eta=0.0, batch=batch)
    if full:
        kw.update(noise="per_tile", tile_offset=tiles.shard_range(T, rank, world)[0])
    total_tiles = T if full else T * world

    def step_device():
        u8 = tiles.stylize_tiles(model, ctx_d, hints_d, xT_d, **kw)
        if full:
            return tiles.stitch_tiles(tiles.gather_tiles(u8, T, world=world), 16)
        return tiles.gather_tiles(u8, T * world) if world > 1 else u8

    def step_e2e():
        tiles.stylize_tiles(model, ctx_h, hints_h, xT_h, out_host=out_h, **kw)
        return out_h

    def timed(fn, k):
        """barrier + sync, k steps between CUDA events on the launching stream, sync + barrier; max over ranks."""
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0 = E.LAUNCHES[0]
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item(), E.LAUNCHES[0] - n0

    for _ in range(max(args.warmup, 1)):
        step_device()
    clk = ClockSampler(local) if rank == 0 else None
    ms, launches = timed(step_device, args.steps)
    clocks = clk.stop() if clk else None
    value = total_tiles * args.steps / (ms / 1e3)

    step_e2e()  # warm the pinned-memory path
    ms_e2e, _ = timed(step_e2e, args.steps)
    e2e = total_tiles * args.steps / (ms_e2e / 1e3)
    h2d = hints_h.numel() * 4 + ctx_h.numel() * 4 + (0 if xT_h is None else xT_h.numel() * 4)
    d2h = out_h.numel()

    roof = end_to_end = None
    if rank == 0:
        nb = min(batch, ctx_d.shape[0])
        xr = xT_d[:nb] if xT_d is not None else tiles.start_latents(nb, (4, 64, 64), 1234, dev, "per_tile", 0)
        roof, gflop_live = roofline_leg(E, eng, model, pk, ctx_d[:nb], hints_d[:nb], xr)
        gflop = wl["gflop"] or gflop_live
        tflop_tile = (31 * 2 * gflop + 2 * GFLOP_HOIST + GFLOP_VAE) / 1e3
        end_to_end = {"achieved_tflops": value / world * tflop_tile, "frac_of_sustained_bf16":
                      value / world * tflop_tile / pk["tf_sust"], "tflop_per_tile": tflop_tile,
                      "gflop_per_element_call": gflop, "gflop_per_element_call_counted_live": gflop_live,
                      "gflop_source": "SURVEY.md §8(d)" if wl["gflop"] else "engine's per-launch algorithmic FLOP count"}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r = cpu_reference(wl, 3, 1)
        cpu = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 1), "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "strong" if full else "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": config_of(args, wl, world),
                "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "ms_per_step": ms_e2e / args.steps, "api": "topo_b200.tiles.stylize_tiles(pinned host in, host out)"},
                "gpu_launches": launches, "clocks": clocks, "roofline": roof, "end_to_end_roofline": end_to_end,
                "cpu_baseline": cpu}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="ctrlora", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
