"""In-graph cost of every kernel class of the decode step, by leaving the class out.

    B200LLM_LIB=<pkg>/libb200llm_tune.so python profiles/decode_classes.py [workload] [hbm_gbs]

B200LLM_SKIP (a bit mask read at engine creation, honoured only by the tune build) drops one
class of kernels from the captured step; the difference to the full step is what that class
costs INSIDE the graph, where launches overlap through programmatic dependent launch (ncu's
serialised per-launch times over-count). Results with a class left out are garbage; only the
timing is used. Prints one JSON object: per class the in-graph microseconds, the algorithmic HBM
bytes of the class and bytes/us over the measured HBM peak.
"""
import importlib
import json
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
L = importlib.import_module(f"{PKG}._lib")

workload = sys.argv[1] if len(sys.argv) > 1 else "smart_reply_b64"
hbm_gbs = float(sys.argv[2]) if len(sys.argv) > 2 else 6454.0
assert "tune" in str(L.LIB_PATH), "needs B200LLM_LIB=<pkg>/libb200llm_tune.so (build.py --tune)"
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

# bits 64 / 128 exist only on the fused decode path (d_model <= 768, <= 64 rows), where the layer's
# LayerNorms live inside the QKV / MLP-up GEMMs and there is no gelu-reduce kernel
CLASSES = [("layernorm_reduce", 1), ("layer_gemms", 2), ("ln_gemm_qkv_and_mlp_up", 64), ("gemm_reduce_add_o_and_mlp_down", 128),
           ("attention", 4), ("lm_head", 8), ("select", 16), ("gelu_reduce", 32)]


def step_us(mask: int) -> float:
    os.environ["B200LLM_SKIP"] = str(mask)
    eng = E.Engine(shape, blob, max_seqs=B, max_seq_len=T + N, max_batch_tokens=min(B * T, 16384), max_new_tokens=N)
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


d, Lr, V, H = shape.d_model, shape.n_layer, shape.vocab_size, shape.n_head
ctx = T + steps / 2.0
alg = {
    "layer_gemms": 2 * Lr * 12 * d * d,                      # bf16 weights of the 4 GEMMs per layer, streamed once
    "attention": int(B * ctx * Lr * 2 * d * 2 + B * Lr * 2 * d * 2),   # KV read for ctx tokens + one token written
    "lm_head": V * d * 2 + B * d * 2,
    "layernorm_reduce": 0, "gelu_reduce": 0, "select": 0,    # activations only: L2-resident, no HBM bytes by design
    "ln_gemm_qkv_and_mlp_up": 2 * Lr * 7 * d * d,            # c_attn (3 d^2) + c_fc (4 d^2) weights
    "gemm_reduce_add_o_and_mlp_down": 2 * Lr * 5 * d * d,    # attn c_proj (d^2) + mlp c_proj (4 d^2) weights
}
full = step_us(0)
out = {"workload": workload, "step_us": full, "method": "B200LLM_SKIP differences inside the CUDA graph (libb200llm_tune.so)",
       "classes": {}}
acc = 0.0
for name, bit in CLASSES:
    us = max(full - step_us(bit), 0.0)
    if bit in (64, 128) and us < 0.03 * full:
        continue  # not the fused path: these classes do not exist in this step
    if bit != 2:
        acc += us  # (bit 2 = all layer GEMMs: on the fused path that is the sum of the two classes below it)
    entry = {"in_graph_us": us, "share": us / full, "algorithmic_bytes": alg[name]}
    if alg[name] > 0 and us > 0:
        entry["gb_per_s"] = alg[name] / us / 1e3
        entry["frac_of_hbm_peak"] = alg[name] / us / 1e3 / hbm_gbs
    out["classes"][name] = entry
if "ln_gemm_qkv_and_mlp_up" not in out["classes"]:
    acc += out["classes"]["layer_gemms"]["in_graph_us"]
out["unattributed_us"] = full - acc   # embed + overlap between classes (differences do not add up exactly)
out["hbm_gbs"] = hbm_gbs
print(json.dumps(out))
