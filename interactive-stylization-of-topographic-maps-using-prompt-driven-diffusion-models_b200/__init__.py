"""topo_b200 — B200-native hot path of the prompt-driven topographic-map stylization reference:
the ControlNet + LoRA SD-1.5 UNet denoising step over 512x512 map tiles.

Importable as ``topo_b200`` (see ``topo_b200/__init__.py`` for why the directory name differs).
Sub-modules: ``ldm`` (module tree / state-dict surface), ``lora`` (the LoRA_modified API), ``sampling``
(DDIMSampler, sample_preview, validate_step), ``engine`` (execution plan on the sm_100a kernels),
``ops`` / ``capi`` (C-ABI binding), ``tiles`` (tile sharding, stitching, multi-GPU gather).
"""
__version__ = "0.1.0"
