"""Decode-step time of the fused path under tune-build switches (one process, several engines).

    B200LLM_LIB=<pkg>/libb200llm_tune.so python profiles/fused_probe.py [workload]
Each line: the environment of the engine and the step time. B200LLM_SKIP bits 64 / 128 leave out
the LayerNorm+GEMM kernels / the reduce-add GEMMs (timing only); B200LLM_FUSED_SPLITS=o,p sets the
K splits of the out-projection and the MLP down-projection."""
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
SP = E.SamplingParams(N, N)
steps = min(N - 1, 63)


def run(env: dict, reduce: int) -> float:
    for k in ("B200LLM_SKIP", "B200LLM_FUSED_SPLITS", "B200LLM_SPLITS", "B200LLM_FUSED_QKV", "B200LLM_LNQKV_SPLIT"):
        os.environ.pop(k, None)
    os.environ.update(env)
    eng = E.Engine(shape, blob, max_seqs=B, max_seq_len=T + N, max_batch_tokens=min(B * T, 16384), max_new_tokens=N,
                   decode_reduce=reduce)
    eng.set_stream(stream.cuda_stream)
    best = 1e9
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
        for s in slots:
            eng.seq_free(s)
    eng.close()
    return best


cases = [({}, 1), ({}, 3)]
for sp in sys.argv[2:]:
    if "=" in sp:
        cases.append((dict(kv.split("=") for kv in sp.split(";")), 3))
    elif sp == "ablate":
        cases += [({"B200LLM_SKIP": m}, 3) for m in ("64", "128", "192", "4", "196")]
    else:
        cases.append(({"B200LLM_FUSED_SPLITS": sp}, 3))
for env, reduce in cases:
    try:
        print(f"reduce={reduce} {env}: {run(env, reduce):8.1f} us/step", flush=True)
    except Exception as exc:  # noqa: BLE001  an invalid split for this model: report and go on
        print(f"reduce={reduce} {env}: {exc}", flush=True)
