"""K1 in isolation at the hot-path shapes (Be = 38): the LoRA-fused projection (q/k/v fused launch and the out-projection with
residual) next to the same projection without LoRA.  CUDA events, L2 flushed between iterations.  TOPO_LORA_BN=128|160|192
overrides the LoRA tile width (A/B)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import torch
from topo_b200 import capi, ops
from topo_b200 import engine as E
BF = torch.bfloat16
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def lin(N, K, r):
    W = (torch.randn(N, K, device="cuda") * K ** -0.5).to(BF)
    if not r:
        return E.Lin(W, None)
    A, Bm, rp = ops.prep_lora(torch.randn(r, K, device="cuda") * K ** -0.5, torch.randn(N, r, device="cuda") * 0.1)
    return E.Lin(W, None, A, Bm, rp, 1.0, r)


def timeit(fn, flops, nbytes, name):
    for _ in range(2):
        fn()
    ts = []
    for _ in range(7):
        flush.zero_()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[len(ts) // 2]
    print(f"{name:44s} {ms * 1e3:8.1f} us  {flops / 1e9 / ms:7.1f} TF/s  {nbytes / 1e6 / ms:7.1f} GB/s  {capi.last_kernel()}", flush=True)


for M, C, hdp in ((38 * 4096, 320, 384), (38 * 1024, 640, 640), (38 * 256, 1280, 1280)):
    x = torch.randn(M, C, device="cuda").to(BF)
    o = torch.randn(M, hdp, device="cuda").to(BF)
    h = torch.randn(M, C, device="cuda").to(BF)
    for r in (0, 16, 32, 8):
        qkv = [lin(hdp, C, r) for _ in range(3)]
        out3 = torch.empty(M, 3 * hdp, device="cuda", dtype=BF)
        fl = 3 * (2.0 * M * hdp * C + 2.0 * M * r * (C + hdp))
        timeit(lambda: E.gemm_multi(x, qkv, out3, [0, hdp, 2 * hdp]), fl, 2.0 * (M * C + 3 * hdp * C + 3 * M * hdp),
               f"qkv  M{M} N3x{hdp} K{C} r{r}")
        lo = lin(C, hdp, r)
        out = torch.empty(M, C, device="cuda", dtype=BF)
        timeit(lambda: E.gemm(o, lo, out=out, res=h), 2.0 * M * C * hdp + 2.0 * M * r * (C + hdp), 2.0 * (M * hdp + C * hdp + 2 * M * C),
               f"out  M{M} N{C} K{hdp} res r{r}")
        q1 = lin(hdp, C, r)
        outq = torch.empty(M, hdp, device="cuda", dtype=BF)
        timeit(lambda: E.gemm(x, q1, out=outq), 2.0 * M * hdp * C + 2.0 * M * r * (C + hdp), 2.0 * (M * C + C * hdp + M * hdp),
               f"q    M{M} N{hdp} K{C} r{r}")
