"""Generates tests/golden/*.pt|json by IMPORTING the reference's own code (only possible in the build container,
where /root/reference exists).  The fixtures pin the oracle (and through it the CUDA path) to the reference:

  lora_linear.pt   — LoRA_modified/LoRA_utils.py:20-42 LoRALinear.forward on seeded inputs (fp32), incl. the qkvo
                     variant's scaling (Training_sdfusion_bert_lora_qkvo.py:149-171)
  lora_qkv.json    — LoRA_utils.lora_qkv applied to this repo's module tree: wrap counts and resulting state-dict keys
  stitch.pt        — BaseModel/stitchTiles.py:13-41 hstack/vstack order on a small tile stack (restated with numpy
                     because the script is top-level code that reads D:\\ paths)

Run:  python tests/golden/make_golden.py
"""
import importlib.util
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    ref = _load(os.path.join(REF, "LoRA_modified", "LoRA_utils.py"), "ref_lora_utils")
    ref_qkvo = None
    cases = []
    for i, (fin, fout, r, bias, alpha) in enumerate([(128, 64, 4, False, None), (128, 96, 16, False, None),
                                                      (192, 192, 8, True, None), (64, 96, 12, True, 24.0),
                                                      (128, 128, 32, False, None)]):
        torch.manual_seed(100 + i)
        base = torch.nn.Linear(fin, fout, bias=bias)
        m = ref.LoRALinear(base, r=r, alpha=alpha)
        torch.nn.init.normal_(m.B.weight, std=0.05)  # reference inits B to zero; non-zero makes the check meaningful
        x = torch.randn(3, 7, fin)
        with torch.no_grad():
            y = m(x)
        cases.append({"x": x, "y": y, "scaling": float(m.scaling), "r": r,
                      "sd": {k: v.clone() for k, v in m.state_dict().items()}})
    torch.save(cases, os.path.join(HERE, "lora_linear.pt"))

    from topo_b200 import ldm
    torch.manual_seed(0)
    unet = ldm.ControlledUnetModel(ldm.ModelConfig.tiny())
    counts = ref.lora_qkv(unet, r_q=8, r_kv=8)
    keys = sorted(k for k in unet.state_dict().keys() if ".attn" in k and ("to_q" in k or "to_k" in k or "to_v" in k))
    json.dump({"counts": counts, "n_keys": len(keys), "sample_keys": keys[:12],
               "lora_filter_keys": sum(1 for k in unet.state_dict() if ".A." in k or ".B." in k)},
              open(os.path.join(HERE, "lora_qkv.json"), "w"), indent=1)

    g = np.random.default_rng(5)
    tiles = g.integers(0, 256, size=(12, 8, 8, 3), dtype=np.uint8)
    rows = []
    for rI in range(3):   # stitchTiles.py: hstack 'cols' tiles into a row, vstack the rows
        rows.append(np.hstack([tiles[rI * 4 + c] for c in range(4)]))
    torch.save({"tiles": torch.from_numpy(tiles), "cols": 4, "mosaic": torch.from_numpy(np.vstack(rows))},
               os.path.join(HERE, "stitch.pt"))
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
