"""-m gpu, collected LAST (file name): ``topo_groupnorm_silu_static`` — GroupNorm(+SiLU) whose statistics scratch is a
file-scope ``static __device__`` array of the library instead of a caller workspace, for callers that cannot allocate.
No counterpart in the reference; checked against the existing entry
point ``topo_groupnorm_silu`` on the same inputs, through the C ABI, on two calls in a row.

All device memory is allocated by the test BEFORE the calls under test; between the two ``cudaMemGetInfo`` readings
only the new entry point runs, so free device memory must not move.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = 5e-3          # the tolerance of the other GroupNorm tests; the two entry points share kernels, so 0.0 is measured


def _call(lib, h, fn_static, x, g, b, y, B, HW, Cn, G, silu, ws=None):
    from topo_b200 import capi
    st = capi.stream()
    if fn_static:
        rc = lib.topo_groupnorm_silu_static(h, x.data_ptr(), Cn, g.data_ptr(), b.data_ptr(), y.data_ptr(), Cn, B, HW, Cn, G, 1e-5,
                                            int(silu), st)
        capi.check(rc, "topo_groupnorm_silu_static")
    else:
        rc = lib.topo_groupnorm_silu(h, x.data_ptr(), Cn, g.data_ptr(), b.data_ptr(), y.data_ptr(), Cn, B, HW, Cn, G, 1e-5,
                                     int(silu), ws.data_ptr(), st)
        capi.check(rc, "topo_groupnorm_silu")


# (B, HW, C): the two-pass shapes of a denoising step (partial sums go through the scratch) and one small shape of the
# single-kernel path (scratch untouched)
@pytest.mark.parametrize("B,HW,Cn", [(2, 4096, 320), (4, 1024, 640), (16, 4096, 320), (2, 64, 1280)])
def test_static_scratch_groupnorm_matches_workspace_entry_point(B, HW, Cn):
    from topo_b200 import capi
    lib, h = capi.load_library(), capi.handle()
    G = 32
    assert lib.topo_groupnorm_ws_floats(B, G) <= lib.topo_groupnorm_static_max_floats()
    gen = torch.Generator(device="cuda").manual_seed(B * 1000 + Cn)
    gamma = torch.randn(Cn, device="cuda", generator=gen)
    beta = torch.randn(Cn, device="cuda", generator=gen)
    ws = torch.empty(lib.topo_groupnorm_ws_floats(B, G), device="cuda", dtype=torch.float32)
    xs = [(torch.randn(B, HW, Cn, device="cuda", generator=gen) * (1.0 + 3.0 * k) + 0.5 * k).to(torch.bfloat16) for k in range(2)]
    y_ref = [torch.empty_like(x) for x in xs]
    y_new = [torch.full_like(x, float("nan")) for x in xs]
    for k in range(2):
        _call(lib, h, False, xs[k], gamma, beta, y_ref[k], B, HW, Cn, G, k == 0, ws)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for k in range(2):                     # two calls in a row on DIFFERENT data: the second must not see the first's sums
        _call(lib, h, True, xs[k], gamma, beta, y_new[k], B, HW, Cn, G, k == 0)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0, "the static-scratch entry point changed free device memory"
    for k in range(2):
        assert torch.isfinite(y_new[k].float()).all()
        err = (y_new[k].float() - y_ref[k].float()).abs().max().item()
        assert err <= TOL, (k, err)
        ref32 = torch.nn.functional.group_norm(xs[k].float().transpose(1, 2), G, gamma, beta, 1e-5).transpose(1, 2)
        if k == 0:
            ref32 = torch.nn.functional.silu(ref32)
        rel = (y_new[k].float() - ref32).abs().max().item() / ref32.abs().max().item()
        assert rel <= 1e-2, (k, rel)


def test_static_scratch_groupnorm_rejects_sizes_beyond_the_maximum():
    from topo_b200 import capi
    lib, h = capi.load_library(), capi.handle()
    G = 32
    B = lib.topo_groupnorm_static_max_floats() // lib.topo_groupnorm_ws_floats(1, G) + 1      # 17
    x = torch.zeros(B, 64, 320, device="cuda", dtype=torch.bfloat16)
    g = torch.ones(320, device="cuda")
    y = torch.full_like(x, 7.0)
    rc = lib.topo_groupnorm_silu_static(h, x.data_ptr(), 320, g.data_ptr(), g.data_ptr(), y.data_ptr(), 320, B, 64, 320, G, 1e-5, 0,
                                        capi.stream())
    assert rc == -1
    msg = capi.last_error()
    assert "static scratch" in msg and "topo_groupnorm_silu" in msg, msg
    torch.cuda.synchronize()
    assert (y == 7.0).all()                # nothing was launched
    rc = lib.topo_groupnorm_silu_static(h, None, 320, g.data_ptr(), g.data_ptr(), y.data_ptr(), 320, 1, 64, 320, G, 1e-5, 0, capi.stream())
    assert rc == -1
