"""CPU: structure / state-dict / error behaviour of topo_b200/models.py and self-consistency of its oracle
(oracle/models_oracle.py restates LoRA_modified/models.py; diffusers is absent, so that file's parity is unpinned)."""
import pytest
import torch

from topo_b200 import models as M
from topo_b200.capi import TopoError
from oracle import models_oracle as O


def test_control_lora_structure():
    m = M.ControlLoRA()
    sd = m.state_dict()
    assert [len(l) for l in m.lora_layers] == [10, 10, 10, 2]
    assert "down_blocks.0.0.downsamplers.0.conv.weight" in sd and "pre_lora_layers.3.convnets.0.conv1.weight" in sd
    assert sd["pre_lora_layers.1.convnets.0.conv1.weight"].shape == (640, 256, 1, 1)
    assert sd["lora_layers.0.1.to_k_lora.down.weight"].shape == (4, 768)      # cross-attention processors read the context
    assert sd["lora_layers.0.0.to_k_lora.down.weight"].shape == (4, 320)
    assert sd["lora_layers.2.0.to_control.down.weight"].shape == (4, 1280)
    assert m.config["lora_rank"] == 4 and m.config["block_out_channels"] == (32, 64, 128, 256)
    m2 = M.ControlLoRA(lora_control_version=2, lora_rank=8, lora_control_rank=16)
    p = m2.lora_layers[1][0]
    assert isinstance(p, M.ControlLoRACrossAttnProcessorV2) and p.to_control_out.down.weight.shape == (16, 1280)
    m3 = M.ControlLoRA(lora_pre_conv_skipped=True)
    assert isinstance(m3.pre_lora_layers[0], torch.nn.Identity)
    assert m3.lora_layers[0][0].to_control.down.weight.shape == (4, 256)        # control_channels = lora_block_in_channels


def test_oracle_pyramid_shapes_and_padding():
    torch.manual_seed(0)
    m = M.ControlLoRA()
    outs = O.control_lora_forward(m, torch.rand(1, 3, 64, 64), inject=True)
    assert [tuple(o.shape) for o in outs] == [(1, 320, 8, 8), (1, 640, 4, 4), (1, 1280, 2, 2), (1, 1280, 1, 1)]
    assert m.lora_layers[3][1].control_states is outs[3]


def test_oracle_processor_algebra():
    torch.manual_seed(0)
    attn = M.CrossAttention(64, 32, heads=2, dim_head=32)
    h, e = torch.randn(2, 16, 64), torch.randn(2, 5, 32)
    plain = O.processor_call(attn, M.LoRACrossAttnProcessor(64, 32, rank=4), h, e)      # zero-init up: identity LoRA
    q, k, v = O._lin(attn.to_q, h), O._lin(attn.to_k, e), O._lin(attn.to_v, e)
    ref = O._lin(attn.to_out[0], O.attention_core(attn, q, k, v))
    assert torch.allclose(plain, ref, atol=1e-6)
    # the value path of a pre-LoRA ignores `scale` (reference :260): doubling scale must not double its contribution
    proc, pre = M.ControlLoRACrossAttnProcessor(64, 32, rank=4), M.LoRACrossAttnProcessor(64, 32, rank=4)
    with torch.no_grad():
        pre.to_v_lora.up.weight.normal_(0, 0.1)
    proc.inject_pre_lora(pre)
    proc.inject_control_states(torch.randn(2, 64, 4, 4))
    a = O.processor_call(attn, proc, h, e, 1.0)
    proc.inject_control_states(torch.randn(2, 64, 4, 4))
    b = O.processor_call(attn, proc, h, e, 2.0)
    assert torch.allclose(a, b, atol=1e-5)        # every other LoRA `up` is zero, so only the unscaled v term acts
    assert proc.control_states.shape == (2, 16, 64)


def test_no_cpu_forward():
    attn = M.CrossAttention(320, 768, 8, 40)
    attn.set_processor(M.LoRACrossAttnProcessor(320, 768))
    with pytest.raises(TopoError):
        attn(torch.randn(1, 64, 320), encoder_hidden_states=torch.randn(1, 77, 768))
    with pytest.raises(TopoError):
        M.ControlLoRA()(torch.rand(1, 3, 64, 64))
    with pytest.raises(TopoError):
        M.get_down_block("CrossAttnDownBlock2D", 1, 32, 32, None, True, 1e-6, "silu", None)
    p = M.LoRACrossAttnProcessor(320, 768, key_states_skipped=True)
    with pytest.raises(AssertionError):
        p.skip_key_states(False)


def test_text_encoder_containers():
    from transformers import BertConfig, BertModel
    from topo_b200.textenc import BERTTextEncoder, LastStateAdapter, build_text_encoder
    assert list(LastStateAdapter(768, 768, 2).state_dict())[:4] == ["ff.0.weight", "ff.0.bias", "ff.2.weight", "ff.2.bias"]
    assert "ff.15.weight" in LastStateAdapter(768, 768, 5).state_dict()          # addlrSD: 5 hidden layers, ff.15 = last Linear
    bert = BertModel(BertConfig(num_hidden_layers=1, hidden_dropout_prob=0.1), add_pooling_layer=False)
    te = build_text_encoder(hf_bert=bert, tokenizer=None, device="cpu")
    assert te.hf_bert is bert and te.max_len == 77 and te.encode == te.encode_on_device
    assert all(m.p == 0.0 for m in bert.modules() if isinstance(m, torch.nn.Dropout))   # disable_bert_dropout
    assert not any(p.requires_grad for p in bert.parameters()) and all(p.requires_grad for p in te.adapter.parameters())
    with pytest.raises(TopoError):
        te.encode_ids(torch.zeros(1, 8, dtype=torch.long))      # CPU: no forward
    with pytest.raises(TopoError):
        te.encode(["a prompt"])                                   # no tokenizer
    with pytest.raises(TopoError):
        BERTTextEncoder(BertModel(BertConfig(num_hidden_layers=1, hidden_act="relu"), add_pooling_layer=False))


def test_bert_align_container_matches_reference_names():
    """token_utils.py:95-112: parameters ``proj.weight`` (zeros), ``proj.bias`` (zeros), scalar ``beta`` = gate_init."""
    import torch
    from topo_b200 import capi
    from topo_b200.textenc import BertAlign
    m = BertAlign(768, gate_init=-2.0)
    sd = m.state_dict()
    assert sorted(sd) == ["beta", "proj.bias", "proj.weight"]
    assert float(sd["beta"]) == -2.0 and float(sd["proj.weight"].abs().max()) == 0.0
    import pytest
    with pytest.raises(capi.TopoError):
        m(torch.zeros(1, 4, 768))
