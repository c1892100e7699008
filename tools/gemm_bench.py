"""Micro-benchmark of the GEMM / conv / attention kernels at the hot-path shapes (CUDA events, L2 flushed between
iterations by writing a 256 MB buffer).  Usage: python tools/gemm_bench.py [filter]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
from topo_b200 import ops, capi
from topo_b200 import engine as E
BF = torch.bfloat16
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
filt = sys.argv[1] if len(sys.argv) > 1 else ""
iters = int(os.environ.get("ITERS", "5"))


def timeit(fn, flops, nbytes, name):
    if filt and filt not in name:
        return
    for _ in range(2):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[len(ts) // 2]
    print(f"{name:52s} {ms*1e3:8.1f} us  {flops/1e9/ms:7.1f} TF/s  {nbytes/1e6/ms:7.1f} GB/s  {capi.last_kernel()}", flush=True)


def gemm_case(M, N, K, res=False, act=0, r=0):
    x = torch.randn(M, K, device="cuda").to(BF)
    W = (torch.randn(N, K, device="cuda") * K ** -0.5).to(BF)
    b = torch.randn(N, device="cuda")
    lin = E.Lin(W, b)
    if r:
        A, Bm, rp = ops.prep_lora(torch.randn(r, K, device="cuda") * K ** -0.5, torch.randn(N, r, device="cuda") * 0.1)
        lin = E.Lin(W, b, A, Bm, rp, 1.0, r)
    if act == 2:
        lin = E.Lin(W, b)
    n_out = N // 2 if act == 2 else N
    out = torch.empty(M, n_out, device="cuda", dtype=BF)
    rs = torch.randn(M, n_out, device="cuda").to(BF) if res else None
    timeit(lambda: E.gemm(x, lin, out=out, res=rs, act=act), 2.0 * M * N * K, 2.0 * (M * K + N * K + M * n_out * (2 if res else 1)),
           f"gemm M{M} N{N} K{K} res{int(res)} act{act} r{r}")


def conv_case(B, H, C, N, res=False):
    x = torch.randn(B * H * H, C, device="cuda").to(BF)
    W = (torch.randn(N, 9 * C, device="cuda") * (9 * C) ** -0.5).to(BF)
    lin = E.Lin(W, torch.randn(N, device="cuda"))
    out = torch.empty(B * H * H, N, device="cuda", dtype=BF)
    rs = torch.randn(B * H * H, N, device="cuda").to(BF) if res else None
    M = B * H * H
    timeit(lambda: E.conv3(x, B, H, H, lin, out=out, res=rs), 2.0 * M * N * 9 * C, 2.0 * (M * C + N * 9 * C + M * N * (2 if res else 1)),
           f"conv B{B} H{H} C{C} N{N} res{int(res)}")


def attn_case(B, N, Nk, d, heads=8):
    dp = ops.head_pad(d)
    q = torch.randn(B, N, heads * dp, device="cuda").to(BF)
    k = torch.randn(B, Nk, heads * dp, device="cuda").to(BF)
    v = torch.randn(B, Nk, heads * dp, device="cuda").to(BF)
    timeit(lambda: ops.attention(q, k, v, heads, dp, d ** -0.5), 4.0 * B * heads * N * Nk * d, 2.0 * (2 * B * N + 2 * B * Nk) * heads * dp,
           f"attn B{B} N{N} Nk{Nk} d{d}")


def qkv_case(M, N, K):
    x = torch.randn(M, K, device="cuda").to(BF)
    Ws = [(torch.randn(N, K, device="cuda") * K ** -0.5).to(BF) for _ in range(3)]
    out = torch.empty(M, 3 * N, device="cuda", dtype=BF)
    segs = [capi.make_linear(W, out, N, out_ld=3 * N, out_col=i * N) for i, W in enumerate(Ws)]
    timeit(lambda: ops.lora_linear_raw(x, M, K, K, segs), 6.0 * M * N * K, 2.0 * (M * K + 3 * N * K + 3 * M * N), f"qkv  M{M} N3x{N} K{K}")


Be = 38
gemm_case(Be * 4096, 384, 320)
qkv_case(Be * 4096, 384, 320)
qkv_case(Be * 1024, 640, 640)
qkv_case(Be * 256, 1280, 1280)
gemm_case(Be * 4096, 320, 320, res=True)
gemm_case(Be * 4096, 320, 320)
gemm_case(Be * 4096, 320, 384, res=True)
gemm_case(Be * 4096, 2560, 320, act=2)
gemm_case(Be * 4096, 320, 1280, res=True)
gemm_case(Be * 64, 1280, 1280, res=True)
gemm_case(Be * 64, 1280, 5120, res=True)
gemm_case(Be * 1024, 5120, 640, act=2)
gemm_case(Be * 1024, 640, 640, res=True)
gemm_case(Be * 1024, 640, 2560, res=True)
gemm_case(Be * 256, 10240, 1280, act=2)
gemm_case(Be * 256, 1280, 1280, res=True)
gemm_case(Be * 256, 1280, 5120, res=True)
gemm_case(Be * 4096, 320, 384, res=True, r=16)
conv_case(Be, 64, 320, 320, True)
conv_case(Be, 32, 640, 640, True)
conv_case(Be, 16, 1280, 1280, True)
conv_case(Be, 8, 1280, 1280, True)
attn_case(Be, 4096, 4096, 40)
attn_case(Be, 1024, 1024, 80)
attn_case(Be, 256, 256, 160)
attn_case(Be, 4096, 77, 40)
