"""Small driver for ncu: one prefill + a few decode steps of a BASELINE workload.

    python profiles/prof_step.py [workload] [decode_steps] [graphs 0|1]
"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "distributed-real-time-chat-and-collaboration-tool_b200"
from bench import WORKLOADS  # noqa: E402

E = importlib.import_module(f"{PKG}.engine")
W = importlib.import_module(f"{PKG}.weights")

workload = sys.argv[1] if len(sys.argv) > 1 else "smart_reply_b64"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
graphs = bool(int(sys.argv[3])) if len(sys.argv) > 3 else False
model, B, T, N = WORKLOADS[workload]
shape = E.MODEL_SHAPES[model]
if T + N > shape.n_positions:
    shape = E.ModelShape(**{**shape.__dict__, "n_positions": 2048})
eng = E.Engine(shape, W.synthetic_blob(shape), max_seqs=B, max_seq_len=T + N,
               max_batch_tokens=min(B * T, 16384), max_new_tokens=N, use_cuda_graphs=graphs)
rng = np.random.default_rng(0)
prompts = rng.integers(0, shape.vocab_size - 1, size=(B, T), dtype=np.int32)
slots = [eng.seq_alloc(T + N) for _ in range(B)]
chunk = max(1, min(B, 16384 // T))
for c0 in range(0, B, chunk):
    eng.prefill(list(prompts[c0:c0 + chunk]), slots[c0:c0 + chunk],
                [E.SamplingParams(N, N)] * len(slots[c0:c0 + chunk]))
eng.decode(slots, steps)
eng.sync()
print("launches", eng.kernel_launches)
eng.close()
