"""ncu target: the LM head of one decode step exactly as the engine launches it -- fused greedy
argmax, no logits stored (b200llm_op_lm_head_argmax; the kernel bench.py's `roofline` reports),
64 tokens x 50257 vocabulary x 768, weights rotated through 4 copies so each launch reads HBM."""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ops = importlib.import_module("distributed-real-time-chat-and-collaboration-tool_b200.ops")
dev = torch.device("cuda:0")
V, d, B = 50257, 768, 64
w = [(torch.randn(V, d, device=dev) * 0.02).to(torch.bfloat16) for _ in range(4)]
act = torch.randn(B, d, device=dev).to(torch.bfloat16)
work = None
for i in range(8):
    work = ops.lm_head_argmax(act, w[i % 4], V - 1, None, work)
torch.cuda.synchronize()
print("ok")
