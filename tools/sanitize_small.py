"""One launch of every kernel family at its smallest shape that still selects that kernel (incl. the cta_group::2 pair
kernels and the LoRA pre-pass path), for `compute-sanitizer --tool memcheck|racecheck|synccheck`.
Each result is also checked against fp32 torch math so a silent corruption shows up as a mismatch, not only as a
sanitizer report.  Prints one line per kernel with the kernel name the C ABI reports."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import torch
import torch.nn.functional as F
from topo_b200 import capi, ops
from topo_b200 import engine as E
BF = torch.bfloat16
g = torch.Generator(device="cuda").manual_seed(0)
bad = []


def rnd(*s, scale=1.0):
    return (torch.randn(*s, device="cuda", generator=g) * scale)


def check(name, got, ref, tol=3e-2):
    torch.cuda.synchronize()
    rel = (got.float() - ref.float()).abs().max().item() / (ref.float().abs().max().item() + 1e-12)
    print(f"{name:46s} rel={rel:.2e}  {capi.last_kernel()}", flush=True)
    if not rel <= tol:
        bad.append((name, rel))


def gemm_case(M, N, K, r=0, res=False, act=0, nseg=1):
    x = rnd(M, K).to(BF)
    lins, refs = [], []
    for _ in range(nseg):
        W = rnd(N, K, scale=K ** -0.5).to(BF)
        A = Bm = None
        ref = x.float() @ W.float().t()
        if r:
            A0, B0 = rnd(r, K, scale=K ** -0.5), rnd(N, r, scale=0.2)
            A, Bm, rp = ops.prep_lora(A0, B0)
            ref = ref + (x.float() @ A.float().t()) @ Bm.float().t()
            lins.append(E.Lin(W, None, A, Bm, rp, 1.0, r))
        else:
            lins.append(E.Lin(W, None))
        refs.append(ref)
    if nseg > 1:
        out = torch.empty(M, nseg * N, device="cuda", dtype=BF)
        E.gemm_multi(x, lins, out, [i * N for i in range(nseg)])
        check(f"gemm M{M} N{nseg}x{N} K{K} r{r}", out, torch.cat(refs, 1))
        return
    rs = rnd(M, N).to(BF) if res else None
    if act == 2:
        idx = ops.geglu_index(N, "cuda")
        lin = E.Lin(lins[0].W[idx].contiguous(), None)
        out = E.gemm(x, lin, act=2)
        a, gte = refs[0].chunk(2, -1)
        check(f"gemm GEGLU M{M} N{N} K{K}", out, a * F.gelu(gte))
        return
    out = E.gemm(x, lins[0], res=rs)
    check(f"gemm M{M} N{N} K{K} r{r} res{int(res)}", out, refs[0] + (rs.float() if res else 0))


def conv_case(B, H, C, N):
    x = rnd(B, H, H, C).to(BF)
    w = rnd(N, C, 3, 3, scale=(9 * C) ** -0.5)
    y = ops.conv3x3(x, w)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.to(BF).float(), padding=1).permute(0, 2, 3, 1)
    check(f"conv3x3 B{B} H{H} C{C} N{N}", y, ref)


def attn_case(B, Nq, Nk, d, lcol=False):
    h, dp = 2, ops.head_pad(d)
    def mk(n):
        t = torch.zeros(B, n, h, dp, device="cuda")
        t[..., :d] = rnd(B, n, h, d)
        return t
    q, k, v = mk(Nq), mk(Nk), mk(Nk)
    vv = v.clone()
    if lcol:
        vv[..., d] = 1.0
    o = ops.attention(q.reshape(B, Nq, -1).to(BF), k.reshape(B, Nk, -1).to(BF), vv.reshape(B, Nk, -1).to(BF), h, dp, d ** -0.5,
                      l_col=d if lcol else -1)
    qf, kf, vf = (t.to(BF).float()[..., :d].permute(0, 2, 1, 3) for t in (q, k, v))
    ref = (torch.softmax(qf @ kf.transpose(-1, -2) * d ** -0.5, -1) @ vf).permute(0, 2, 1, 3)
    check(f"attention B{B} Nq{Nq} Nk{Nk} d{d} lcol{int(lcol)}", o.reshape(B, Nq, h, dp)[..., :d], ref)


gemm_case(256, 128, 64)                      # <128,0,6>
gemm_case(148 * 128, 320, 320, res=True)     # <160,0,5>
gemm_case(148 * 128 * 2, 256, 320)           # <256,0,3>
gemm_case(148 * 128 * 2, 256, 1024)          # <256,0,4> or pair <256,0,6,0,1>
gemm_case(148 * 128, 512, 320, act=2)        # GEGLU <256,0,3>
gemm_case(148 * 128, 384, 320, nseg=3)       # <192,0,4> multi
gemm_case(512, 384, 320, r=16, nseg=3)       # LoRA <192,16,4>
gemm_case(512, 320, 384, r=16, res=True)     # LoRA <160,16,5>
gemm_case(512, 640, 640, r=32, res=True)     # LoRA <160,32,4>
gemm_case(384, 200, 128, r=4)                # LoRA <128,16,5>, ragged N
gemm_case(384, 384, 128, r=32)               # LoRA <192,32,4>
conv_case(1, 16, 64, 128)                    # <128,0,6,1,0>
conv_case(2, 64, 64, 320)                    # pair <320,0,5,1,1>
conv_case(2, 64, 64, 512)                    # pair <256,0,6,1,1>
attn_case(1, 128, 128, 40)
attn_case(1, 512, 512, 40, lcol=True)
attn_case(1, 256, 256, 80)
attn_case(1, 128, 192, 160)
attn_case(2, 384, 77, 40)                    # cross_attention_kernel<48,2,1> (TMA store), ragged K/V segments
attn_case(2, 256, 77, 80)
attn_case(1, 128, 77, 160)
attn_case(1, 64, 77, 160)                    # flash alias mode
x = rnd(2, 64, 64, 320).to(BF)
gm, bt = rnd(320).float(), rnd(320).float()
y = ops.groupnorm_silu(x, gm, bt, 32, 1e-5, True)
check("groupnorm_silu 64x64x320", y, F.silu(F.group_norm(x.float().permute(0, 3, 1, 2), 32, gm, bt, 1e-5)).permute(0, 2, 3, 1))
x = rnd(2, 8, 8, 1280).to(BF)
gm, bt = rnd(1280).float(), rnd(1280).float()
y = ops.groupnorm_silu(x, gm, bt, 32, 1e-5, True)
check("groupnorm_silu 8x8x1280 (fused small)", y, F.silu(F.group_norm(x.float().permute(0, 3, 1, 2), 32, gm, bt, 1e-5)).permute(0, 2, 3, 1))
x = rnd(300, 640).to(BF)
gm, bt = rnd(640).float(), rnd(640).float()
check("layernorm 300x640", ops.layernorm(x, gm, bt), F.layer_norm(x.float(), (640,), gm, bt))
xx, ee = rnd(2, 4, 8, 8), rnd(4, 4, 8, 8)
px0 = torch.empty_like(xx)
xo = ops.ddim_step(xx, ee, True, 7.5, 0.5, 0.6, pred_x0=px0)
e = ee[:2] + 7.5 * (ee[2:] - ee[:2])
x0 = (xx - (0.5 ** 0.5) * e) / 0.5 ** 0.5
check("ddim_step_x0", xo, 0.6 ** 0.5 * x0 + 0.4 ** 0.5 * e, 1e-5)
check("ddim_step_x0 pred", px0, x0, 1e-5)
print("MISMATCHES:", bad)
sys.exit(1 if bad else 0)
