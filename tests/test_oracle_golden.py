"""not-gpu: pin the oracle against fixtures generated from the reference's own code (tests/golden/make_golden.py)
and against the known answers of the upstream DDIM schedule."""
import json
import os

import numpy as np
import torch

from oracle import ldm_oracle as O

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_lora_linear_matches_reference_loralinear():
    """oracle.lora_linear == LoRA_modified/LoRA_utils.py:40-42 on the reference's own outputs (fp32, 1e-5)."""
    cases = torch.load(os.path.join(G, "lora_linear.pt"))
    assert len(cases) == 5
    for c in cases:
        sd = {"p." + k: v for k, v in c["sd"].items()}
        y = O.lora_linear(sd, "p", c["x"], {"p": c["scaling"]})
        assert torch.allclose(y, c["y"], atol=1e-5, rtol=1e-5)
    # default scaling of the reference is (alpha or 2r)/r = 2.0 (LoRA_utils.py:38)
    assert [c["scaling"] for c in cases] == [2.0, 2.0, 2.0, 2.0, 2.0]
    sd = {"p." + k: v for k, v in cases[0]["sd"].items()}
    assert torch.allclose(O.lora_linear(sd, "p", cases[0]["x"]), cases[0]["y"], atol=1e-5)


def test_reference_lora_qkv_on_our_module_tree():
    """The reference's lora_qkv, run on this repo's UNet tree, wrapped 32/32/32 projections and produced the
    *.base / *.A / *.B key layout the '.A.'/'.B.' checkpoint filter relies on (…_addlrSD.py:539-543)."""
    ref = json.load(open(os.path.join(G, "lora_qkv.json")))
    assert ref["counts"] == {"q": 32, "k": 32, "v": 32}
    from topo_b200 import ldm, lora
    torch.manual_seed(0)
    unet = ldm.ControlledUnetModel(ldm.ModelConfig.tiny())
    counts = lora.lora_qkv(unet, r_q=8, r_kv=8)
    assert counts == ref["counts"]
    keys = sorted(k for k in unet.state_dict() if ".attn" in k and ("to_q" in k or "to_k" in k or "to_v" in k))
    assert len(keys) == ref["n_keys"] and keys[:12] == ref["sample_keys"]
    assert sum(1 for k in unet.state_dict() if ".A." in k or ".B." in k) == ref["lora_filter_keys"]


def test_stitch_matches_reference_order():
    d = torch.load(os.path.join(G, "stitch.pt"))
    from topo_b200 import tiles
    assert torch.equal(tiles.stitch_tiles(d["tiles"], d["cols"]), d["mosaic"])
    assert np.array_equal(O.stitch_tiles(d["tiles"].numpy(), d["cols"]), d["mosaic"].numpy())
    # split_sheet is the inverse (datasetCreation00.py row-major tiling)
    assert torch.equal(tiles.split_sheet(d["mosaic"], 8), d["tiles"])


def test_ddim_schedule_known_answers():
    """ldm 'uniform' schedule: c = 1000 // S, ts = range(0, 1000, c) + 1 → S=30 gives 31 steps."""
    for S, n, first3, last in [(4, 4, [1, 251, 501], 751), (12, 13, [1, 84, 167], 997), (16, 17, [1, 63, 125], 993),
                               (20, 20, [1, 51, 101], 951), (30, 31, [1, 34, 67], 991)]:
        ts, a, a_prev, sig = O.make_schedule(S)
        assert len(ts) == n and ts[:3].tolist() == first3 and int(ts[-1]) == last
        assert a_prev[0] == a_prev[0] and torch.all(sig == 0)
    betas = np.linspace(0.00085 ** 0.5, 0.012 ** 0.5, 1000, dtype=np.float64) ** 2
    ac = np.cumprod(1 - betas)
    ts, a, a_prev, _ = O.make_schedule(30)
    assert np.allclose(a.numpy(), ac[ts.numpy()], rtol=1e-6)
    assert np.isclose(float(a_prev[0]), ac[0], rtol=1e-6) and np.allclose(a_prev[1:].numpy(), ac[ts.numpy()[:-1]], rtol=1e-6)
    from topo_b200.sampling import make_ddim_timesteps
    for S in (4, 12, 15, 16, 20, 30):
        assert make_ddim_timesteps(S).tolist() == O.make_schedule(S)[0].tolist()


def test_oracle_tiny_model_runs_and_is_deterministic():
    """Whole-path oracle on the tiny config: apply_model + 2-step CFG DDIM + VAE decode; non-trivial, finite, repeatable."""
    import synth
    from topo_b200 import ldm
    model = synth.build_model(ldm.ModelConfig.tiny())
    model.control_model  # noqa
    sd = {k: v.float() for k, v in model.state_dict().items()}
    ctx, hint, xT = synth.make_ctx(1, dim=64), synth.make_hint(1, 64), synth.make_xT(1, 8)
    img, x, inter = O.sample_preview(sd, ctx, hint, steps=2, scale=7.5, x_T=xT, heads=2)
    img2, _, _ = O.sample_preview(sd, ctx, hint, steps=2, scale=7.5, x_T=xT, heads=2)
    assert img.shape == (1, 3, 64, 64) and torch.isfinite(img).all() and torch.equal(img, img2)
    assert len(inter) == 2 and float(img.std()) > 1e-3
    u8 = O.tiles_to_uint8(img)
    assert u8.dtype == np.uint8 and u8.shape == (1, 64, 64, 3)


def _golden_module():
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(G, "make_golden.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_control_lora_matches_reference_attention_lora():
    """oracle.control_lora_encoder / lora_control_attention == the reference's own ControlLoRA and
    CrossAttention_LoRA_Control (LoRA_modified/attention_lora.py:35-101) on its own outputs (fp32, 1e-5)."""
    d = torch.load(os.path.join(G, "control_lora.pt"))
    y = O.control_lora_encoder({"e." + k: v for k, v in d["enc_sd"].items()}, "e", d["enc_x"])
    assert y.shape == d["enc_y"].shape == (2, 256, 16, 16) and torch.allclose(y, d["enc_y"], atol=1e-5, rtol=1e-5)
    sd = {"a." + k: v for k, v in d["attn_sd"].items()}
    assert d["scaling"] == 2.0 and d["y_bad_equals_none"]
    y0 = O.lora_control_attention(sd, "a", d["x"], d["ctx"], 2, None)
    y1 = O.lora_control_attention(sd, "a", d["x"], d["ctx"], 2, d["cs"], d["scale"])
    yb = O.lora_control_attention(sd, "a", d["x"], d["ctx"], 2, d["cs"][:, :, :2], d["scale"])
    assert torch.allclose(y0, d["y_none"], atol=1e-5, rtol=1e-5) and torch.allclose(y1, d["y_ctrl"], atol=1e-5, rtol=1e-5)
    assert torch.allclose(y1 - y0, d["ctrl"], atol=1e-5) and torch.equal(yb, y0)
    assert float((d["y_ctrl"] - d["y_none"]).abs().max()) > 0.05          # the control term is not negligible in the fixture


def test_drivers_match_reference_vis_metrics(monkeypatch):
    """This repo's ``validate_step`` / ``sample_preview`` against outputs of the REFERENCE's own functions
    (TextEncoder_Finetuning/vis_metrics.py:73-116, 407-452) run on the same deterministic mock model and recording
    sampler (tests/golden/make_golden.py): identical eps-MSE, identical call sequence, identical sampler arguments
    (zeros uc, shared hint object, S, shape, scale, eta, x_T=None after manual_seed(seed)), identical decoded preview."""
    mg = _golden_module()
    d = torch.load(os.path.join(G, "drivers.pt"), weights_only=False)
    from topo_b200 import sampling
    batch, model = mg.driver_batch(), mg.MockLDM()
    torch.manual_seed(5)
    val = sampling.validate_step(model, mg.MockEnc(), batch, "cpu", iters=3)
    assert val == d["validate"] and model.calls == d["validate_calls"]
    val_fixed = sampling.validate_step(model, mg.MockEnc(), batch, "cpu", iters=2, fixed_t=d["fixed_t"], fixed_noise=d["fixed_noise"])
    assert val_fixed == d["validate_fixed"]
    monkeypatch.setattr(sampling, "DDIMSampler", mg.RecordingSampler)
    pred = sampling.sample_preview(model, mg.MockEnc(), batch, "cpu", steps=12, scale=6.5, eta=0.0, seed=4321)
    assert torch.equal(pred, d["preview"]) and float(pred.max()) == 1.0 and float(pred.min()) == -1.0
    got, want = mg.RecordingSampler.last, d["sampler_kwargs"]
    assert sorted(got) == sorted(want)
    for k in want:
        if isinstance(want[k], dict):
            assert sorted(got[k]) == sorted(want[k])
            for n in want[k]:
                assert all(torch.equal(a, b) for a, b in zip(got[k][n], want[k][n])), (k, n)
        else:
            assert got[k] == want[k], k
    assert got["same_hint_object"] and not bool(got["unconditional_conditioning"]["c_crossattn"][0].any())


def _build_proc(M, rec, C, ctx):
    """This repo's processor object for one golden case, loaded with the reference's state dict."""
    cls = getattr(M, rec["cls"])
    proc = cls(C, cross_attention_dim=ctx if rec["cross"] else None, **rec["kw"])
    proc.load_state_dict(rec["sd"])
    if "control" in rec:
        proc.inject_control_states(rec["control"].clone())
    if "pre_sd" in rec:
        pre = M.LoRACrossAttnProcessor(C, ctx, rank=4)
        pre.load_state_dict(rec["pre_sd"])
        post = M.LoRACrossAttnProcessor(C, ctx, rank=8, post_add=True)
        post.load_state_dict(rec["post_sd"])
        proc.inject_pre_lora(pre)
        proc.inject_post_lora(post)
    return proc


def test_models_oracle_matches_reference_models_py():
    """oracle/models_oracle.py == the reference's own LoRA_modified/models.py (processor algebra :70-431 and the
    ControlLoRA pyramid :434-833) on outputs generated by running that file (tests/golden/make_golden_models.py; the
    diffusers primitives it imports are stand-ins restating diffusers 0.12, everything on top is the reference's code).
    The same state dicts load into this repo's classes: constructor signatures and key names are the reference's."""
    from oracle import models_oracle as MO
    from topo_b200 import models as M
    d = torch.load(os.path.join(G, "models_ref.pt"), weights_only=False)
    dims = d["dims"]
    C, ctx = dims["C"], dims["ctx"]
    attns = {}
    for cross in (False, True):
        a = M.CrossAttention(C, ctx if cross else None, dims["heads"], dims["d"])
        a.load_state_dict(d["attn_cross" if cross else "attn_self"])
        attns[cross] = a
    assert len(d["cases"]) == 9
    for name, rec in d["cases"].items():
        proc = _build_proc(M, rec, C, ctx)
        out = MO.processor_call(attns[rec["cross"]], proc, d["hidden"], d["enc"] if rec["cross"] else None, rec["scale"])
        err = (out - rec["out"]).abs().max().item()
        assert out.shape == rec["out"].shape and err <= 2e-5 * max(1.0, rec["out"].abs().max().item()), (name, err)
    p = d["pyramid"]
    pyr = M.ControlLoRA(**p["kw"])
    missing = pyr.load_state_dict(p["sd"], strict=True)
    maps = MO.control_lora_forward(pyr, p["x"], inject=True)
    assert p["return_dict_true_is_none"] and len(maps) == 4
    for got, want in zip(maps, p["maps"]):
        assert got.shape == want.shape and torch.allclose(got, want, atol=2e-5, rtol=1e-4)
    assert [tuple(l[0].control_states.shape) for l in pyr.lora_layers] == [tuple(s) for s in p["injected_shapes"]]
