"""Would k engine replicas on ONE GPU (batch 64/k each, own stream + CUDA graphs) finish a decode
step sooner than one replica with the whole batch? The step is latency-bound, so independent
chains can overlap.   python profiles/two_replicas.py [k ...]"""
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "distributed-real-time-chat-and-collaboration-tool_b200"
E = importlib.import_module(f"{PKG}.engine")
W = importlib.import_module(f"{PKG}.weights")

B, T, N = 64, 128, 32
shape = E.MODEL_SHAPES["small"]
blob = W.synthetic_blob(shape)
prompts = np.random.default_rng(0).integers(0, shape.vocab_size - 1, size=(B, T), dtype=np.int32)
for k in [int(a) for a in sys.argv[1:]] or [1, 2, 4]:
    b = B // k
    engines, streams, slots = [], [], []
    for r in range(k):
        st = torch.cuda.Stream()
        eng = E.Engine(shape, blob, max_seqs=b, max_seq_len=T + N, max_batch_tokens=b * T, max_new_tokens=N)
        eng.set_stream(st.cuda_stream)
        engines.append(eng)
        streams.append(st)
    best = 1e9
    best_total = 1e9
    for rep in range(4):
        slots = []
        start = torch.cuda.Event(enable_timing=True)
        mid = torch.cuda.Event(enable_timing=True)
        ends = [torch.cuda.Event(enable_timing=True) for _ in range(k)]
        torch.cuda.synchronize()
        start.record(torch.cuda.current_stream())
        for r, eng in enumerate(engines):
            streams[r].wait_event(start)
            s = [eng.seq_alloc(T + N) for _ in range(b)]
            slots.append(s)
            eng.prefill(list(prompts[r * b:(r + 1) * b]), s, [E.SamplingParams(N, N)] * b)
        for r in range(k):
            ev = torch.cuda.Event()
            ev.record(streams[r])
            torch.cuda.current_stream().wait_event(ev)
        mid.record(torch.cuda.current_stream())
        for r, eng in enumerate(engines):
            streams[r].wait_event(mid)
        for r, eng in enumerate(engines):
            eng.decode(slots[r], N - 1)
            ends[r].record(streams[r])
        for eng in engines:
            eng.sync()
        dec = max(mid.elapsed_time(e) for e in ends)
        tot = max(start.elapsed_time(e) for e in ends)
        best = min(best, dec / (N - 1))
        best_total = min(best_total, tot)
        for r, eng in enumerate(engines):
            for s in slots[r]:
                eng.seq_free(s)
    print(f"replicas on one GPU: {k} x batch {b}: decode {best * 1e3:7.1f} us per step of all {B} rows; "
          f"prefill+decode {best_total:6.2f} ms -> {B * N / best_total:7.1f} k tok/s")
    for eng in engines:
        eng.close()
