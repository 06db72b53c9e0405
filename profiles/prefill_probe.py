"""Prefill time of a workload under the tune build's switches: B200LLM_PREFILL_ATTN (0 tcgen05, 1 mma.sync,
-1 auto) or, with PROBE=pair_split, B200LLM_PAIR_SPLIT (K splits of the residual pair GEMMs: 1 none, -1 by
the wave model, n forced)."""
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
for workload in sys.argv[1:] or ["smart_reply_b64"]:
    model, B, T, N = WORKLOADS[workload]
    shape = E.MODEL_SHAPES[model]
    if T + N > shape.n_positions:
        shape = E.ModelShape(**{**shape.__dict__, "n_positions": 2048})
    blob = W.synthetic_blob(shape)
    prompts = np.random.default_rng(0).integers(0, shape.vocab_size - 1, size=(B, T), dtype=np.int32)
    stream = torch.cuda.Stream()
    chunk = int(os.environ.get("CHUNK", 0)) or max(1, min(B, 16384 // T))  # sequences per prefill call
    knob, modes = ("B200LLM_PAIR_SPLIT", ("1", "-1", "2", "3", "4")) if os.environ.get("PROBE") == "pair_split" \
        else ("B200LLM_PREFILL_ATTN", ("0", "1", "-1"))
    for mode in modes:
        os.environ[knob] = mode
        eng = E.Engine(shape, blob, max_seqs=B, max_seq_len=T + N, max_batch_tokens=min(B * T, 16384), max_new_tokens=N)
        eng.set_stream(stream.cuda_stream)
        best = 1e9
        for _ in range(5):
            slots = [eng.seq_alloc(T + N) for _ in range(B)]
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for c0 in range(0, B, chunk):
                eng.prefill(list(prompts[c0:c0 + chunk]), slots[c0:c0 + chunk], [E.SamplingParams(N, N)] * len(slots[c0:c0 + chunk]))
            e1.record(stream)
            eng.sync()
            best = min(best, e0.elapsed_time(e1))
            for s in slots:
                eng.seq_free(s)
        eng.close()
        print(f"{workload}: prefill attention mode {mode:>2}: {best:7.3f} ms", flush=True)
