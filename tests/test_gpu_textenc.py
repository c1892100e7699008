"""-m gpu: text conditioning producer (topo_b200/textenc.py) vs HuggingFace's own BertModel.forward in fp32 — the
reference's text encoder IS that library call (multiseg_19prompts_batch1_shuffle_ctrlora.py:181-186), so this parity is
pinned to the real implementation (random-init weights: bert-base-uncased's checkpoint is not available offline).
Tolerance: 12 post-LN transformer layers in bf16 -> 3e-2 of the max magnitude, cosine >= 0.999."""
import pytest
import torch

from gpu_util import assert_close

pytestmark = pytest.mark.gpu


def _bert(layers, seed=0):
    from transformers import BertConfig, BertModel
    torch.manual_seed(seed)
    cfg = BertConfig(num_hidden_layers=layers, hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    m = BertModel(cfg, add_pooling_layer=False).eval()
    with torch.no_grad():   # default init (std 0.02) gives near-identity layers; make the test bite
        for n, p in m.named_parameters():
            if p.ndim == 2 and "embeddings" not in n:
                p.mul_(2.5)
            elif n.endswith("bias"):
                p.normal_(0, 0.05)
    return m


def _ids(B, T, lens, seed=1):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(1000, 30000, (B, T), generator=g)
    mask = (torch.arange(T)[None, :] < torch.tensor(lens)[:, None]).long()
    return ids * mask, mask   # [PAD] = 0 beyond the prompt


def _cos(a, b):
    return torch.nn.functional.cosine_similarity(a.flatten().float().cpu(), b.flatten().float().cpu(), dim=0).item()


@pytest.mark.parametrize("layers", [2, 12])
def test_bert_last_hidden_state_vs_hf(layers):
    from topo_b200.textenc import BERTTextEncoder
    bert = _bert(layers)
    ids, mask = _ids(3, 77, [9, 77, 40])
    with torch.no_grad():
        ref = bert(input_ids=ids, attention_mask=mask, return_dict=True).last_hidden_state
    te = BERTTextEncoder(bert).cuda()
    got = te.bert_last_hidden_state(ids, mask).view(3, 77, -1).float().cpu()
    c = _cos(got, ref)
    print(f"[parity] BERT {layers} layers: cosine={c:.6f}")
    assert c >= 0.999
    assert_close(got, ref, 3e-2, f"BERT last_hidden_state ({layers} layers, lengths 9/77/40)")


@pytest.mark.parametrize("n_hidden", [2, 5])
def test_adapter_and_encode(n_hidden, tmp_path):
    from transformers import BertTokenizer
    from topo_b200.textenc import build_text_encoder
    vocab = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]", "forest", "lake", "road", "rail", "##way", "contour", "lines",
             "in", "green", "blue", "swiss", "map", "style", ",", "a", "with", "building", "##s"]
    vf = tmp_path / "vocab.txt"
    vf.write_text("\n".join(vocab + [f"tok{i}" for i in range(30522 - len(vocab))]))
    tok = BertTokenizer(vocab_file=str(vf), do_lower_case=True)
    bert = _bert(2, seed=3)
    te = build_text_encoder(hf_bert=bert, tokenizer=tok, adapter_hidden_layers=n_hidden)
    assert len([m for m in te.adapter.ff if isinstance(m, torch.nn.Linear)]) == n_hidden + 1
    torch.manual_seed(5)
    with torch.no_grad():
        for p in te.adapter.parameters():
            if p.ndim == 2:
                p.normal_(0, 2.0 * p.shape[1] ** -0.5)
    prompts = ["forest in green, lake in blue", "swiss map style with buildings", "road"]
    got = te.encode(prompts)
    assert got.shape == (3, 77, 768) and got.dtype == torch.float32 and got.is_cuda
    enc = tok(prompts, return_tensors="pt", padding="max_length", truncation=True, max_length=77)
    b32, a32 = bert.float().cpu(), te.adapter.float().cpu()
    with torch.no_grad():
        x = b32(input_ids=enc["input_ids"], attention_mask=enc["attention_mask"], return_dict=True).last_hidden_state
        ref = a32.final_ln(a32.ff(x))            # …_ctrlora.py:160-163
    c = _cos(got, ref)
    print(f"[parity] encode() with {n_hidden}-hidden-layer adapter: cosine={c:.6f}")
    assert c >= 0.999
    assert_close(got.cpu(), ref, 3e-2, f"adapter(bert(prompts)) n_hidden={n_hidden}")


def test_mask_must_be_right_padded():
    from topo_b200.textenc import BERTTextEncoder
    from topo_b200.capi import TopoError
    te = BERTTextEncoder(_bert(1)).cuda()
    ids, mask = _ids(1, 16, [8])
    mask[0, 2] = 0
    with pytest.raises(TopoError):
        te.encode_ids(ids, mask)


def test_bert_align_and_clip_bert_fusion():
    """token_utils.py:95-158: BertAlign (LN -> zero-init proj -> sigmoid gate) and cond/uc = w_clip*clip (+ gamma*tail)."""
    import torch.nn.functional as F
    from topo_b200.textenc import BertAlign, build_cond_uc_safe_pad
    torch.manual_seed(0)
    B, Lb, C = 3, 20, 768
    al = BertAlign(C).cuda()
    torch.nn.init.normal_(al.proj.weight, std=0.03)
    torch.nn.init.normal_(al.proj.bias, std=0.1)
    x = torch.randn(B, Lb, C, device="cuda") * 2 + 0.5
    got = al(x)
    ref = torch.sigmoid(al.beta) * F.linear(F.layer_norm(x, (C,)), al.proj.weight, al.proj.bias)
    rel = (got - ref).abs().max().item() / ref.abs().max().item()
    print(f"[parity] BertAlign: rel={rel:.3e}")
    assert rel <= 2e-2

    class Clip:
        def get_learned_conditioning(self, prompts):
            g = torch.Generator().manual_seed(1)
            return torch.randn(len(prompts), 77, C, generator=g).cuda()

    class Bert:
        def encode(self, prompts):
            return x[:len(prompts)]

    seg = torch.rand(B, 3, 64, 64)
    cond, uc = build_cond_uc_safe_pad(seg, ["a", "b", "c"], None, Clip(), Bert(), "cuda", gamma=0.7, w_clip=0.4,
                                      bert_align=al, start_idx=8)
    clip = Clip().get_learned_conditioning(["x"] * B)
    tail = torch.zeros(B, 77, C, device="cuda")
    tail[:, 8:8 + Lb] = F.layer_norm(ref, (C,))
    want = 0.4 * clip + 0.7 * tail
    rel2 = (cond["c_crossattn"][0] - want).abs().max().item() / want.abs().max().item()
    print(f"[parity] build_cond_uc_safe_pad: rel={rel2:.3e}")
    assert rel2 <= 3e-2
    assert torch.equal(uc["c_crossattn"][0], 0.4 * clip) and cond["c_concat"][0] is uc["c_concat"][0]
    assert cond["c_crossattn"][0].shape == (B, 77, C)
