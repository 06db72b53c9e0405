"""Decode-step time with the two decode-attention kernels (tune build, one process, several engines).

    B200LLM_LIB=<pkg>/libb200llm_tune.so python profiles/attn_stream_probe.py workload [K=V;K=V ...]
Without extra arguments: B200LLM_ATTN_STREAM=0 (one CTA per unit, cp.async passes) against =9 (persistent
warps, TMA tiles; 1..5 force an instance) and the engine's own choice, each with the engine's split rule; further arguments are environments to add,
e.g. "B200LLM_ATTN_STREAM=1;B200LLM_ATTN_SPLITS=4"."""
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

workload = sys.argv[1] if len(sys.argv) > 1 else "summarize_medium"
model, B, T, N = WORKLOADS[workload]
shape = E.MODEL_SHAPES[model]
if T + N > shape.n_positions:
    shape = E.ModelShape(**{**shape.__dict__, "n_positions": 2048})
blob = W.synthetic_blob(shape)
rng = np.random.default_rng(0)
prompts = rng.integers(0, shape.vocab_size - 1, size=(B, T), dtype=np.int32)
stream = torch.cuda.Stream()
chunk = max(1, min(B, 16384 // T))
SP = E.SamplingParams(N, N)
steps = min(N - 1, 63)
KNOBS = ("B200LLM_ATTN_STREAM", "B200LLM_ATTN_SPLITS", "B200LLM_ATTN_MERGE", "B200LLM_SKIP", "B200LLM_FUSED_QKV",
         "B200LLM_LNQKV_SPLIT", "B200LLM_LNUP_SPLIT", "B200LLM_FUSED_SPLITS", "REDUCE")


def run(env: dict):
    for k in KNOBS:
        os.environ.pop(k, None)
    os.environ.update(env)
    eng = E.Engine(shape, blob, max_seqs=B, max_seq_len=T + N, max_batch_tokens=min(B * T, 16384), max_new_tokens=N,
                   decode_reduce=int(os.environ.get("REDUCE", "0")))  # REDUCE=3 in a case: the fused decode layer forced
    eng.set_stream(stream.cuda_stream)
    best = 1e9
    toks = None
    for _ in range(3):
        slots = [eng.seq_alloc(T + N) for _ in range(B)]
        for c0 in range(0, B, chunk):
            eng.prefill(list(prompts[c0:c0 + chunk]), slots[c0:c0 + chunk], [SP] * len(slots[c0:c0 + chunk]))
        e1, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e1.record(stream)
        eng.decode(slots, steps)
        e2.record(stream)
        eng.sync()
        best = min(best, e1.elapsed_time(e2) / steps * 1e3)
        toks = [eng.read_tokens(s)[0][:steps + 1] for s in slots]
        for s in slots:
            eng.seq_free(s)
    eng.close()
    return best, toks


cases = [{"B200LLM_ATTN_STREAM": "0"}, {"B200LLM_ATTN_STREAM": "9"}, {}]
for sp in sys.argv[2:]:
    cases.append(dict(kv.split("=") for kv in sp.split(";")))
ref = None
for env in cases:
    try:
        t, toks = run(env)
        if ref is None:
            ref = toks
        same = sum(int(list(a) == list(b)) for a, b in zip(ref, toks))
        print(f"{workload} {env}: {t:8.1f} us/step   rows with the first case's tokens: {same}/{len(toks)}", flush=True)
    except Exception as exc:  # noqa: BLE001
        print(f"{workload} {env}: FAILED {exc}", flush=True)
