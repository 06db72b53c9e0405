"""ncu target: a few stand-alone launches of the fused decode kernels (GPT-2 small shapes)."""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ops = importlib.import_module("distributed-real-time-chat-and-collaboration-tool_b200.ops")
dev = torch.device("cuda:0")
torch.manual_seed(0)
rows, d = 64, 768
x = torch.randn(64, d, device=dev)
gamma, beta = torch.ones(d, device=dev), torch.zeros(d, device=dev)
w = (torch.randn(3 * d, d, device=dev) * 0.05).to(torch.bfloat16)
bias = torch.zeros(3 * d, device=dev)
act = (torch.randn(64, 4 * d, device=dev) * 0.5).to(torch.bfloat16)
w2 = (torch.randn(d, 4 * d, device=dev) * 0.05).to(torch.bfloat16)
xx = torch.zeros(64, d, device=dev)
h = x.to(torch.bfloat16).contiguous()
out = torch.zeros(rows, 3 * d, dtype=torch.bfloat16, device=dev)
for _ in range(4):
    ops.ln_gemm(x, gamma, beta, w, bias)
    ops.gemm_redadd(act, w2, None, xx, rows, 0)
    ops.gemm(h, w, bias, out, ops.EPI_BIAS_BF16, orientation=1)
torch.cuda.synchronize()
print("ok")
