"""Test-side alias of the package's synthetic workload factory (``topo_b200.synth``)."""
from topo_b200.synth import *  # noqa: F401,F403
from topo_b200.synth import PALETTE, build_model, lora_scales, make_ctx, make_hint, make_xT  # noqa: F401
