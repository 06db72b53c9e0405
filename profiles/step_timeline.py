"""Timeline of one decode step inside the CUDA-graph + PDL pipeline (trace build):
    python <pkg>/build.py --trace
    B200LLM_LIB=<pkg>/libb200llm_trace.so python profiles/step_timeline.py [workload]
Each record is (entry, exit, tag, milestones...) of block 0 of a kernel, %globaltimer ns, in launch order."""
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
TAGS = {1: "layernorm", 2: "gelu_reduce", 3: "select", 4: "embed", 5: "attention", 100: "gemm bias", 101: "gemm gelu",
        102: "gemm resid", 103: "gemm f32", 104: "gemm partial", 105: "gemm lm-head+argmax",
        300: "ln+gemm (qkv)", 301: "ln+gemm+gelu (mlp up)", 310: "gemm reduce-add"}
workload = sys.argv[1] if len(sys.argv) > 1 else "smart_reply_b64"
model, B, T, N = WORKLOADS[workload]
shape = E.MODEL_SHAPES[model]
eng = E.Engine(shape, W.synthetic_blob(shape), max_seqs=B, max_seq_len=T + N, max_batch_tokens=min(B * T, 16384),
               max_new_tokens=N)
prompts = np.random.default_rng(0).integers(0, shape.vocab_size - 1, size=(B, T), dtype=np.int32)
slots = [eng.seq_alloc(T + N) for _ in range(B)]
eng.prefill(list(prompts), slots, [E.SamplingParams(N, N)] * B)
eng.decode(slots, 6)
eng.sync()
fn = L.lib.b200llm_debug_step_trace
fn.argtypes = [C.c_int, C.c_void_p, C.c_int]
cap = 4096
assert fn(0, None, cap) == 0
eng.decode(slots, 3)
eng.sync()
buf = (C.c_ulonglong * (16 * cap))()
n = fn(1, buf, cap)
full = np.array(buf[: 16 * n], dtype=np.int64).reshape(n, 16)
rec = full[:, :3]
t0 = rec[0, 0]
per_step = n // 3
print(f"{n} kernel records, {per_step} per step; step = {(rec[per_step, 0] - rec[0, 0]) / 1e3:.1f} us")
step = rec[per_step: 2 * per_step]  # the middle step
base = step[0, 0]
agg = {}
prev_end = None
marks = full[per_step: 2 * per_step, 4:]
MARK_NAMES = ["entry", "triggered", "weights issued", "dep ok", "1st k-block landed", "mma committed", "epi sees acc",
              "stores done", "dealloc", "epi ld done", "epi 2nd half", "last k-block landed"]
for i, (b, e, tag) in enumerate(step):
    nxt = step[i + 1, 0] if i + 1 < len(step) else rec[2 * per_step, 0]
    name = TAGS.get(int(tag), str(tag))
    a = agg.setdefault(name, [0, 0.0, 0.0])
    a[0] += 1
    a[1] += (e - b) / 1e3          # entry -> exit of CTA 0
    a[2] += (nxt - b) / 1e3        # entry -> next kernel's entry (pipeline pitch)
    if 8 <= i < 26 or i >= len(step) - 4:
        print(f"{i:3d} {name:22s} entry {(b - base) / 1e3:8.2f} us  exit {(e - base) / 1e3:8.2f} us  dur {(e - b) / 1e3:6.2f}")
        if name.startswith("gemm") and "reduce-add" not in name:
            print("      " + "  ".join(f"{MARK_NAMES[j]} {(m - base) / 1e3:.2f}" for j, m in enumerate(marks[i]) if m > 0))
        if "reduce-add" in name:
            names = {3: "dep ok", 4: "operands landed", 6: "epi sees acc", 5: "tile staged", 7: "reduce-add complete"}
            print("      " + "  ".join(f"{names[j]} {(marks[i][j] - base) / 1e3:.2f}" for j in (3, 4, 6, 5, 7) if marks[i][j] > 0))
        if name.startswith("ln+gemm"):
            names = {3: "dep ok", 4: "rows in registers", 5: "B operand written", 8: "mma sees B", 9: "mma issued",
                     6: "epi sees acc", 7: "stores done"}
            print("      " + "  ".join(f"{names[j]} {(marks[i][j] - base) / 1e3:.2f}" for j in (3, 4, 5, 8, 9, 6, 7) if marks[i][j] > 0))
print("per kernel class in one step: count, sum CTA0 duration us, sum entry-to-next-entry us")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][2]):
    print(f"  {k:22s} {a[0]:4d} {a[1]:9.1f} {a[2]:9.1f}")
eng.close()
