"""LM-head-shaped GEMM (64 tokens, K = 768, fp32 logits out) for vocabulary sizes that give every
SM the same number of 128-row tiles (2 or 3) and for GPT-2's 50257 (393 tiles on 148 SMs): is the
393-tile case slower per byte, i.e. would balancing the tiles pay?
    python profiles/lmhead_tiles.py"""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ops = importlib.import_module("distributed-real-time-chat-and-collaboration-tool_b200.ops")

dev = torch.device("cuda:0")
B, K = int(os.environ.get("LMHEAD_TOKENS", "64")), 768
act = (torch.randn(B, K, device=dev) * 0.5).to(torch.bfloat16)
for V in (148 * 2 * 128, 50257, 148 * 3 * 128, 148 * 4 * 128):
    copies = [(torch.randn(V, K, device=dev) * 0.02).to(torch.bfloat16) for _ in range(4)]  # > L2 in total
    out = torch.empty(B, V, dtype=torch.float32, device=dev)
    for w in copies:
        ops.gemm(act, w, None, out, 3, 1, 1)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 40
    e0.record()
    for i in range(n):
        ops.gemm(act, copies[i % 4], None, out, 3, 1, 1)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / n
    byt = V * K * 2 + B * K * 2 + B * V * 4
    print(f"V = {V:6d} ({V / 128 / 148:.2f} tiles per SM): {us:6.2f} us per launch, {byt / us / 1e3:7.1f} GB/s "
          f"({V * K * 2 / us / 1e3:7.1f} GB/s of weights)")
