"""Stand-alone launch time of the fused decode kernels (CUDA events, back-to-back launches)."""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ops = importlib.import_module("distributed-real-time-chat-and-collaboration-tool_b200.ops")
dev = torch.device("cuda:0")
torch.manual_seed(0)


def timeit(fn, reps=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


for rows, d in ((64, 768), (16, 1024), (32, 1280)):
    x = torch.randn(64, d, device=dev)
    gamma, beta = torch.ones(d, device=dev), torch.zeros(d, device=dev)
    for feats, gelu in ((3 * d, False), (4 * d, True)):
        w = (torch.randn(feats, d, device=dev) * 0.05).to(torch.bfloat16)
        bias = torch.zeros(feats, device=dev)
        t = timeit(lambda: ops.ln_gemm(x[:rows], gamma, beta, w, bias, gelu=gelu))
        h = x[:rows].to(torch.bfloat16).contiguous()
        out = torch.zeros(rows, feats, dtype=torch.bfloat16, device=dev)
        t2 = timeit(lambda: ops.gemm(h, w, bias, out, ops.EPI_BIAS_GELU_BF16 if gelu else ops.EPI_BIAS_BF16, orientation=1))
        print(f"rows {rows} d {d}: ln_gemm -> {feats}: {t:7.1f} us   (plain swap-AB gemm, no split: {t2:7.1f} us)")
    for k in (d, 4 * d):
        act = (torch.randn(64, k, device=dev) * 0.5).to(torch.bfloat16)
        w = (torch.randn(d, k, device=dev) * 0.05).to(torch.bfloat16)
        xx = torch.zeros(64, d, device=dev)
        for split in (0,):
            t = timeit(lambda: ops.gemm_redadd(act, w, None, xx, rows, split))
            print(f"rows {rows} d {d}: redadd k={k} split auto: {t:7.1f} us")
