"""In-pipeline timeline of the decode attention kernel (CTA 0 of the last layer, last step):
needs the trace build:  python <pkg>/build.py --trace ; B200LLM_LIB=<pkg>/libb200llm_trace.so python profiles/attn_in_graph_trace.py"""
import ctypes as C
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
L = importlib.import_module(f"{PKG}._lib")
workload = sys.argv[1] if len(sys.argv) > 1 else "smart_reply_b64"
model, B, T, N = WORKLOADS[workload]
shape = E.MODEL_SHAPES[model]
eng = E.Engine(shape, W.synthetic_blob(shape), max_seqs=B, max_seq_len=T + N, max_batch_tokens=min(B * T, 16384),
               max_new_tokens=N)
rng = np.random.default_rng(0)
prompts = rng.integers(0, shape.vocab_size - 1, size=(B, T), dtype=np.int32)
slots = [eng.seq_alloc(T + N) for _ in range(B)]
eng.prefill(list(prompts), slots, [E.SamplingParams(N, N)] * B)
for steps in (4, 12):
    eng.decode(slots, steps)
    eng.sync()
    t = (C.c_ulonglong * 16)()
    L.lib.b200llm_debug_attn_trace.argtypes = [C.c_void_p]
    assert L.lib.b200llm_debug_attn_trace(t) == 0
    names = ["start", "prewait done", "pos known", "first loads issued", "qkv staged", "loop done", "end"]
    d = {n: int(t[i] - t[0]) for i, n in enumerate(names)}
    d["reduce loads done (t0)"] = int(t[7] - t[0])
    d["cp.async drained (t0)"] = int(t[8] - t[0])
    print("cycles from CTA start:", d)
    nb = B * shape.n_head
    ct = (C.c_ulonglong * (4 * nb))()
    L.lib.b200llm_debug_attn_cta_trace.argtypes = [C.c_void_p, C.c_int]
    assert L.lib.b200llm_debug_attn_cta_trace(ct, nb) == 0
    r = np.array(ct[:], dtype=np.int64).reshape(nb, 4)
    t0 = r[:, 0].min()
    ent, dep, ext, sm = (r[:, 0] - t0) / 1e3, (r[:, 1] - t0) / 1e3, (r[:, 2] - t0) / 1e3, r[:, 3]
    q = lambda x: "min %.2f p50 %.2f p90 %.2f max %.2f" % (x.min(), np.percentile(x, 50), np.percentile(x, 90), x.max())
    print("all CTAs of the last launch, us from first entry:\n  entry  ", q(ent), "\n  dep ok ", q(dep), "\n  exit   ", q(ext))
    late = ent > dep.min()
    print("  CTAs entering after the dependency resolved: %d of %d; per-SM CTA count min/max %d/%d on %d SMs" % (
        late.sum(), nb, np.bincount(sm).min(), np.bincount(sm).max(), len(np.unique(sm))))
    print("  post-dependency time (exit - dep ok):", q(ext - dep))
eng.close()
