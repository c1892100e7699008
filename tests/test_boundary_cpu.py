"""not-gpu: the drop-in boundary — C-ABI exports, Python-face names, state-dict layout, loud failure without a GPU."""
import ctypes
import os
import re

import pytest
import torch
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from topo_b200 import build, capi
    lib_path = build.build()
    lib = ctypes.CDLL(lib_path)
    hdr = open(os.path.join(ROOT, "include", "topo_b200.h")).read()
    declared = sorted(set(re.findall(r"\b(topo_[a-z0-9_]+)\s*\(", hdr)))
    assert declared, "no declarations parsed"
    for sym in declared:
        assert hasattr(lib, sym), f"{sym} declared in include/topo_b200.h but not exported"
    assert sorted(capi.SYMBOLS) == declared
    lib.topo_version.restype = ctypes.c_int
    assert lib.topo_version() == 100


def test_no_cpu_fallback_is_loud():
    from topo_b200 import capi, ldm, lora
    lin = lora.LoRALinear(nn.Linear(64, 64), r=4)
    with pytest.raises(capi.TopoError):
        lin(torch.randn(2, 64))
    model = ldm.create_model(ldm.ModelConfig.tiny())
    with pytest.raises(capi.TopoError):
        model.apply_model(torch.randn(1, 4, 8, 8), torch.zeros(1, dtype=torch.long),
                          {"c_crossattn": [torch.randn(1, 77, 64)], "c_concat": [torch.rand(1, 3, 64, 64)]})
    with pytest.raises(RuntimeError):
        model.model.diffusion_model(torch.randn(1, 4, 8, 8))
    if not torch.cuda.is_available():
        h = ctypes.c_void_p()
        rc = capi.load_library().topo_create(ctypes.byref(h), 0)
        assert rc != 0 and b"no CPU fallback" in capi.load_library().topo_last_error()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "interactive-stylization-of-topographic-maps-using-prompt-driven-diffusion-models_b200")
    for f in os.listdir(pkg):
        if f.endswith(".py") and f != "smoke.py":  # smoke() is the one sanctioned checker call (uses the oracle to CHECK)
            src = open(os.path.join(pkg, f)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), f"{f} imports the oracle"


def test_state_dict_keys_follow_upstream_names():
    from topo_b200 import ldm
    m = ldm.create_model("ctrlora", model_channels=32, num_heads=2, context_dim=64, vae_ch=32, vae_ch_mult=(1, 2, 2, 2))
    keys = set(m.state_dict().keys())
    for k in ["model.diffusion_model.time_embed.0.weight", "model.diffusion_model.input_blocks.0.0.weight",
              "model.diffusion_model.input_blocks.1.0.in_layers.0.weight", "model.diffusion_model.input_blocks.1.0.emb_layers.1.bias",
              "model.diffusion_model.input_blocks.1.1.transformer_blocks.0.attn1.to_q.weight",
              "model.diffusion_model.input_blocks.1.1.transformer_blocks.0.attn2.to_out.0.bias",
              "model.diffusion_model.input_blocks.1.1.transformer_blocks.0.ff.net.0.proj.weight",
              "model.diffusion_model.input_blocks.1.1.transformer_blocks.0.ff.net.2.weight",
              "model.diffusion_model.input_blocks.3.0.op.weight", "model.diffusion_model.middle_block.1.proj_in.weight",
              "model.diffusion_model.output_blocks.2.1.conv.weight", "model.diffusion_model.output_blocks.11.1.norm.weight",
              "model.diffusion_model.out.2.weight", "control_model.input_hint_block.14.weight",
              "control_model.zero_convs.11.0.weight", "control_model.middle_block_out.0.bias",
              "control_model.input_blocks.1.1.transformer_blocks.0.attn1.to_q.lora_layer.down.weight",
              "control_model.middle_block.1.transformer_blocks.0.attn2.to_out.0.lora_layer.up.weight",
              "first_stage_model.post_quant_conv.weight", "first_stage_model.decoder.mid.attn_1.q.weight",
              "first_stage_model.decoder.up.3.upsample.conv.weight", "first_stage_model.decoder.up.0.block.0.nin_shortcut.weight",
              "first_stage_model.decoder.conv_out.bias"]:
        assert k in keys, k
    names = [n for n, _ in m.named_parameters()]
    # the trainable-parameter filters of the config scripts (…_ctrlora.py:250-258, …_onlyCNLora.py:244-248)
    assert sum("lora_layer" in n and n.startswith("control_model") for n in names) == 7 * 2 * 4 * 2
    assert any("zero_convs" in n for n in names) and any("middle_block_out" in n for n in names)
    assert len(m.control_scales) == 13 and m.num_timesteps == 1000


def test_variants_and_surgery_counts():
    from topo_b200 import ldm, lora
    small = dict(model_channels=32, num_heads=2, context_dim=64, with_vae=False)
    m = ldm.create_model("addlrSD", **small)
    unet = m.model.diffusion_model
    assert sum(isinstance(x, lora.LoRALinear) for x in unet.modules()) == 96
    a = unet.input_blocks[1][1].transformer_blocks[0].attn1
    assert a.to_q.scaling == 2.0 and a.to_q.A.weight.shape == (8, 32) and not bool(a.to_q.B.weight.any())
    assert m.cfg.control_lora_rank == 32
    m4 = ldm.create_model("qkvo", **small)
    u4 = m4.model.diffusion_model
    assert sum(isinstance(x, lora.LoRALinear) for x in u4.modules()) == 128
    assert u4.middle_block[1].transformer_blocks[0].attn2.to_out[0].scaling == 4.0
    assert ldm.create_model("configs/ctrlora_finetune_sd15_rank12.yaml", **small).cfg.control_lora_rank == 12
    # attention_lora.py injector: every CrossAttention replaced, base weights carried over
    m5 = ldm.create_model("stock", **small)
    u5 = m5.model.diffusion_model
    w_before = u5.input_blocks[1][1].transformer_blocks[0].attn1.to_q.weight.clone()
    n = lora.inject_control_lora_into_ldm(u5, rank=4, control_rank=2)
    assert n == 32
    a5 = u5.input_blocks[1][1].transformer_blocks[0].attn1
    assert isinstance(a5, lora.CrossAttention_LoRA_Control) and torch.equal(a5.to_q.base.weight, w_before)
    assert a5.to_control[0].weight.shape == (2, 32) and not bool(a5.to_control[1].weight.any())
    n_train = lora.freeze_all_but_lora(u5)
    assert n_train == sum(p.numel() for p in lora.get_lora_parameters(u5)) > 0
    lora.ensure_lora_to_device(u5, "cpu")
    hf = nn.Sequential(nn.Dropout(0.1), nn.Linear(4, 4), nn.Dropout(0.3))
    lora.disable_bert_dropout(hf)
    assert all(d.p == 0.0 for d in hf if isinstance(d, nn.Dropout))


def test_engine_fingerprint_tracks_weight_loads_and_surgery():
    from topo_b200 import engine, ldm, lora
    m = ldm.create_model(ldm.ModelConfig.tiny())
    f0 = engine.fingerprint(m)
    m.load_state_dict(m.state_dict(), strict=False)
    f1 = engine.fingerprint(m)
    assert f1 != f0
    lora.lora_qkv(m.model.diffusion_model, 4, 4)
    assert engine.fingerprint(m) != f1
