"""Decode-step timing (CUDA events on the engine stream) under different engine switches.

    python profiles/time_decode.py [workload]
"""
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "distributed-real-time-chat-and-collaboration-tool_b200"
from bench import WORKLOADS  # noqa: E402

E = importlib.import_module(f"{PKG}.engine")
W = importlib.import_module(f"{PKG}.weights")

workload = sys.argv[1] if len(sys.argv) > 1 else "smart_reply_b64"
model, B, T, N = WORKLOADS[workload]
shape = E.MODEL_SHAPES[model]
if T + N > shape.n_positions:
    shape = E.ModelShape(**{**shape.__dict__, "n_positions": 2048})
blob = W.synthetic_blob(shape)
rng = np.random.default_rng(0)
prompts = rng.integers(0, shape.vocab_size - 1, size=(B, T), dtype=np.int32)
stream = torch.cuda.Stream()
chunk = max(1, min(B, 16384 // T))
# TIME_DECODE_SAMPLING=1: every row samples (temperature 0.7, top-k 40) like the service's ask-AI path
SP = (E.SamplingParams(N, N, temperature=0.7, top_k=40, seed=1) if os.environ.get("TIME_DECODE_SAMPLING")
      else E.SamplingParams(N, N))

for graphs, pdl, split, reduce in [(True, False, 1, 0), (True, False, 0, 0), (True, True, 0, 0), (True, True, 0, 2)]:
    eng = E.Engine(shape, blob, max_seqs=B, max_seq_len=T + N, max_batch_tokens=min(B * T, 16384),
                   max_new_tokens=N, use_cuda_graphs=graphs, use_pdl=pdl, decode_split_k=split,
                   decode_reduce=reduce)
    eng.set_stream(stream.cuda_stream)
    best = 1e9
    pf = 1e9
    for rep in range(4):
        slots = [eng.seq_alloc(T + N) for _ in range(B)]
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record(stream)
        for c0 in range(0, B, chunk):
            eng.prefill(list(prompts[c0:c0 + chunk]), slots[c0:c0 + chunk],
                        [SP] * len(slots[c0:c0 + chunk]))
        e1.record(stream)
        eng.decode(slots, N - 1)
        e2.record(stream)
        eng.sync()
        best = min(best, e1.elapsed_time(e2) / (N - 1))
        pf = min(pf, e0.elapsed_time(e1))
        for s in slots:
            eng.seq_free(s)
    print(f"graphs={graphs!s:5} pdl={pdl!s:5} split={split} reduce={reduce}: decode {best * 1e3:8.1f} us/step   prefill {pf:7.3f} ms")
    eng.close()
