import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.getcwd())
PKG = "distributed-real-time-chat-and-collaboration-tool_b200"
E = importlib.import_module(f"{PKG}.engine"); W = importlib.import_module(f"{PKG}.weights")
shape = E.MODEL_SHAPES["small"]
eng = E.Engine(shape, W.synthetic_blob(shape), max_seqs=64, max_seq_len=600, max_batch_tokens=1024, max_new_tokens=32)
ids = np.random.default_rng(0).integers(0, 50000, size=(1, 496), dtype=np.int32)
for rep in range(5):
    s = eng.seq_alloc(560); eng.sync()
    t0 = time.perf_counter(); eng.prefill(list(ids), [s], [E.SamplingParams(32, 32)]); t1 = time.perf_counter(); eng.sync(); t2 = time.perf_counter()
    t3 = time.perf_counter(); eng.decode([s], 31); t4 = time.perf_counter(); eng.sync(); t5 = time.perf_counter()
    print(f"prefill: host enqueue {1e3*(t1-t0):.3f} ms, until done {1e3*(t2-t0):.3f} ms; decode 31: enqueue {1e3*(t4-t3):.3f} ms, done {1e3*(t5-t3):.3f} ms")
    eng.seq_free(s)
