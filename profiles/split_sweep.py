import importlib, os, sys
import numpy as np, torch
sys.path.insert(0, "/root/repo")
PKG = "distributed-real-time-chat-and-collaboration-tool_b200"
E = importlib.import_module(f"{PKG}.engine"); W = importlib.import_module(f"{PKG}.weights")
shape = E.GPT2_SMALL; B, T, N = 64, 128, 32
blob = W.synthetic_blob(shape)
prompts = np.random.default_rng(0).integers(0, 50256, size=(B, T), dtype=np.int32)
stream = torch.cuda.Stream()
for cap in [2, 4, 6, 8, 12, 16, 0]:
    eng = E.Engine(shape, blob, max_seqs=B, max_seq_len=T + N, max_batch_tokens=8192, max_new_tokens=N, decode_split_k=cap)
    eng.set_stream(stream.cuda_stream)
    best = 1e9
    for rep in range(3):
        slots = [eng.seq_alloc(T + N) for _ in range(B)]
        eng.prefill(list(prompts), slots, [E.SamplingParams(N, N)] * B)
        e1, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e1.record(stream); eng.decode(slots, N - 1); e2.record(stream); eng.sync()
        best = min(best, e1.elapsed_time(e2) / (N - 1))
        [eng.seq_free(s) for s in slots]
    print(f"split cap {cap:2d}: decode {best*1e3:7.1f} us/step")
    eng.close()
