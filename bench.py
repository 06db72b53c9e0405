#!/usr/bin/env python
"""Headline benchmark: generated tokens/s for batched GPT-2 smart replies on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

Workload (BASELINE.json configs[1], the configuration the metric is quoted on): GPT-2 small
(124M, random-init bf16 weights, seed 0), batch 64 per GPU, 128-token synthetic prompts, 32 new
tokens, greedy, min_new_tokens == max_new_tokens. One "step" = one whole batch generated
(prefill + 31 decode steps -> 2048 tokens per GPU). Under torchrun every rank owns one GPU and
one shard of the request stream (request-sharded replicas, no collective on the data path);
the only torch.distributed traffic is the barrier and the max-over-ranks of the timings.

Prints ONE JSON line (rank 0). See DESIGN.md "Measurement" for how each field is produced.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "distributed-real-time-chat-and-collaboration-tool_b200"

WORKLOADS = {
    # name: (model, batch per GPU, prompt tokens, new tokens)
    "smart_reply_b64": ("small", 64, 128, 32),        # BASELINE configs[1] (headline)
    "smart_reply_b1": ("small", 1, 64, 32),           # configs[0]
    "summarize_medium": ("medium", 16, 1024, 128),    # configs[2] (n_positions 2048)
    "ask_large_b32": ("large", 32, 64, 256),          # configs[3] per-GPU shard (256 / 8 GPUs)
    "ask_large_b256": ("large", 256, 64, 256),        # configs[3] on one GPU
    "probe_short_ctx": ("small", 64, 16, 32),         # profiling probe: negligible KV traffic
    "probe_b1_long": ("small", 1, 496, 32),           # profiling probe: one SmartReply request (byte-level prompt)
    "probe_b16": ("small", 16, 128, 32),              # profiling probes: fewer live rows (stream-like buckets)
    "probe_b32": ("small", 32, 128, 32),
    "probe_b48": ("small", 48, 128, 32),
    "probe_long_ctx": ("small", 64, 512, 64),         # profiling probe: contexts like the byte-level gRPC stream
    "probe_ctx96": ("small", 64, 96, 32),             # one 128-row pass of the decode attention kernel
    "probe_ctx256": ("small", 64, 256, 64),           # profiling probes: where the streaming decode attention takes over
    "probe_ctx192": ("small", 64, 192, 64),
    "probe_large_b32_ctx256": ("large", 32, 256, 64),
    "probe_b16_ctx512": ("small", 16, 512, 64),
}


BASELINE_CONFIG = {"smart_reply_b64": "BASELINE configs[1]", "smart_reply_b1": "BASELINE configs[0]",
                   "summarize_medium": "BASELINE configs[2]", "ask_large_b32": "BASELINE configs[3], one GPU's shard of 8",
                   "ask_large_b256": "BASELINE configs[3] on one GPU"}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines: list[str] = []
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                 "-lms", "100", "-i", str(self.gpu)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._pump, daemon=True)
        self.thread.start()

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                smax.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": float(max(smax)) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks() -> tuple[dict, str]:
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


LMHEAD_PROFILES = ["r2_lmhead_argmax_ncu_full.csv", "r1_lmhead_ncu_full.csv"]


def lmhead_traffic_from_profile() -> tuple[float | None, str | None]:
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the LM-head GEMM, from the newest
    committed `ncu --set full` capture under profiles/."""
    import csv

    path = next((os.path.join(ROOT, "profiles", n) for n in LMHEAD_PROFILES
                 if os.path.exists(os.path.join(ROOT, "profiles", n))), None)
    if path is None:
        return None, None
    rows = list(csv.reader(open(path)))
    hdr = rows[0]
    try:
        ri, wi = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        unit = rows[1][ri].lower() if len(rows) > 1 else "mbyte"
        scale = {"byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(unit, 1e6)
        vals = [(float(r[ri]) + float(r[wi])) * scale for r in rows[2:] if len(r) > max(ri, wi)]
        return (sum(vals) / len(vals) if vals else None), "profiles/" + os.path.basename(path)
    except (ValueError, IndexError):
        return None, None


def decode_step_bytes(shape, batch: int, ctx: int) -> int:
    """Algorithmic HBM bytes of one decode step: bf16 weights streamed once +
    KV read for `ctx` cached tokens + KV written for one token, per sequence."""
    d, L, V = shape.d_model, shape.n_layer, shape.vocab_size
    p_blk = L * (12 * d * d + 13 * d)
    w = 2 * (p_blk + V * d + 2 * d)
    kv = L * 2 * d * 2
    return w + batch * ctx * kv + batch * kv


# ------------------------------------------------------------------------------------------
# p50 SmartReply (second half of BASELINE.json's metric)
# ------------------------------------------------------------------------------------------
SMART_REPLY_CONVERSATION = [
    ("alice", "Are we still on for the design review at 3pm?"),
    ("bob", "Yes, I booked room 4B. Bring the latency numbers."),
    ("charlie", "I can't make it before 3:30, start without me."),
    ("alice", "OK. Bob, can you share the doc beforehand?"),
    ("bob", "Sure, sending it now."),
]


def smart_reply_latency(E, shape, blob, device: int, n_calls: int = 200, cpu_calls: int = 3,
                        prefix_cache: bool = False) -> dict:
    """LLMServiceStub.GetSmartReply round trips on 127.0.0.1 through the whole drop-in stack
    (gRPC -> LLMServicer prompts/parsers -> LocalGenerativeModel -> scheduler -> CUDA engine),
    sequential calls, fixed five-message conversation, 32 new tokens."""
    from concurrent import futures as _f

    import grpc

    server_mod = importlib.import_module(f"{PKG}.server")
    svc = importlib.import_module(f"{PKG}.llm_service")
    pb = importlib.import_module(f"{PKG}.llm_proto")
    sched = importlib.import_module(f"{PKG}.scheduler")
    lm = importlib.import_module(f"{PKG}.local_model")
    tok = importlib.import_module(f"{PKG}.tokenizer")

    eng = E.Engine(shape, blob, device=device, max_seqs=8, max_seq_len=1024, max_batch_tokens=4096,
                   max_new_tokens=64, prefix_cache=prefix_cache)
    router = sched.Router([sched.EngineWorker(eng, decode_burst=31)])
    tokenizer = tok.ByteLevelTokenizer()
    model = lm.LocalGenerativeModel(router, tokenizer, default_max_new_tokens=32, max_prompt_tokens=900)
    server = grpc.server(_f.ThreadPoolExecutor(max_workers=8))
    server_mod.add_llm_service(svc.LLMServicer(model), server)
    port = server.add_insecure_port("127.0.0.1:0")
    server.start()
    msgs = [pb.Message(sender=a, content=b) for a, b in SMART_REPLY_CONVERSATION]
    prompt = svc.PROMPTS["smart_reply"].format(conversation="\n".join(f"{a}: {b}" for a, b in SMART_REPLY_CONVERSATION))
    prompt_tokens = len(tokenizer.encode(prompt))
    lat = []
    try:
        with grpc.insecure_channel(f"127.0.0.1:{port}") as ch:
            stub = server_mod.LLMServiceStub(ch)
            req = pb.SmartReplyRequest(request_id="bench", recent_messages=msgs, user_id="bob")
            for _ in range(10):
                stub.GetSmartReply(req, timeout=20)
            for _ in range(n_calls):
                t0 = time.perf_counter()
                r = stub.GetSmartReply(req, timeout=20)
                lat.append((time.perf_counter() - t0) * 1e3)
            assert len(r.suggestions) >= 2
    finally:
        server.stop(0)
        router.shutdown()
        eng.close()
    out = {"p50_ms": float(np.percentile(lat, 50)), "p95_ms": float(np.percentile(lat, 95)), "calls": n_calls,
           "prompt_tokens": prompt_tokens, "new_tokens": 32,
           "path": "gRPC 127.0.0.1 -> LLMServicer -> LocalGenerativeModel -> scheduler -> CUDA engine"}
    if cpu_calls > 0:
        torch, H, hf, _, _ = _cpu_setup("smart_reply_b1")
        torch.set_num_threads(os.cpu_count() or 1)
        ids = torch.tensor([tokenizer.encode(prompt)])
        cpu = []
        for i in range(cpu_calls + 1):
            t0 = time.perf_counter()
            H.hf_generate(hf, ids, 32, min_new_tokens=32)
            if i:
                cpu.append((time.perf_counter() - t0) * 1e3)
        out["cpu_p50_ms"] = float(np.percentile(cpu, 50))
        out["cpu_note"] = "transformers generate on the host cores, same prompt tokens, direct call (no gRPC)"
    return out


# ------------------------------------------------------------------------------------------
# Our arm
# ------------------------------------------------------------------------------------------
def run_ours(args):
    import torch

    rank, world, local = dist_env()
    if world > 1:
        import torch.distributed as dist

        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    else:
        dist = None
        torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")

    E = importlib.import_module(f"{PKG}.engine")
    ops = importlib.import_module(f"{PKG}.ops")
    W = importlib.import_module(f"{PKG}.weights")

    model, B, T, N = WORKLOADS[args.workload]
    shape = E.MODEL_SHAPES[model]
    if T + N > shape.n_positions:
        shape = E.ModelShape(**{**shape.__dict__, "n_positions": 2048})
    blob = W.synthetic_blob(shape, seed=0, hf_like=True)
    eng = E.Engine(shape, blob, device=local, max_seqs=B, max_seq_len=T + N,
                   max_batch_tokens=min(B * T, 16384), max_new_tokens=N)
    stream = torch.cuda.Stream(device=dev)
    eng.set_stream(stream.cuda_stream)
    sp = [E.SamplingParams(max_new_tokens=N, min_new_tokens=N)] * B

    # one global request stream (B sequences per replica), sharded by request across the ranks
    sharding = importlib.import_module(f"{PKG}.sharding")
    rng = np.random.default_rng(1234)
    request_stream = rng.integers(0, shape.vocab_size - 1, size=(B * world, T), dtype=np.int32)
    prompts_np = np.ascontiguousarray(sharding.shard_requests(request_stream, rank, world))
    assert prompts_np.shape[0] == B
    prompts = [row for row in prompts_np]
    ids_dev = torch.from_numpy(prompts_np.reshape(-1)).to(dev)
    chunk = max(1, min(B, 16384 // T))  # sequences per prefill call

    def one_batch_device_resident(ev=None):
        """Hot path with inputs already in HBM; nothing is copied back inside the region."""
        slots = [eng.seq_alloc(T + N) for _ in range(B)]
        if ev:
            ev[0].record(stream)
        for c0 in range(0, B, chunk):
            c1 = min(B, c0 + chunk)
            eng.prefill(prompts[c0:c1], slots[c0:c1], sp[c0:c1],
                        ids_device_ptr=ids_dev.data_ptr() + c0 * T * 4)
        if ev:
            ev[1].record(stream)
        eng.decode(slots, N - 1)
        if ev:
            ev[2].record(stream)
        return slots

    def free(slots):
        for s in slots:
            eng.seq_free(s)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up (also captures the decode CUDA graph) ----
    for _ in range(max(args.warmup, 3)):
        free(one_batch_device_resident())
        eng.sync()

    # ---- device-resident timing: EXACTLY K steps between barriers, CUDA events on the stream ----
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    launches0 = eng.kernel_launches
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    t_begin = torch.cuda.Event(enable_timing=True)
    t_end = torch.cuda.Event(enable_timing=True)
    t_begin.record(stream)
    pending = []
    for k in range(args.steps):
        pending.append(one_batch_device_resident(evs[k]))
        eng.sync()  # slots are recycled per step; the sync is inside the timed region
        free(pending.pop())
    t_end.record(stream)
    barrier()
    clocks = sampler.stop()
    launches = eng.kernel_launches - launches0
    total_ms = t_begin.elapsed_time(t_end)
    prefill_ms = float(np.mean([e[0].elapsed_time(e[1]) for e in evs]))
    decode_ms = float(np.mean([e[1].elapsed_time(e[2]) for e in evs]))

    # ---- end to end through the public API with HOST buffers (H2D + D2H inside) ----
    host_prompts = [p.tolist() for p in prompts]
    for _ in range(2):
        eng.generate(host_prompts, sp)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = eng.generate(host_prompts, sp)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    assert len(out) == B and all(len(o) == N for o in out)

    # ---- longest single kernel of the decode step in isolation: the LM head exactly as the step
    # launches it (fused greedy argmax, no logits stored; b200llm_op_lm_head_argmax) ----
    V, d = shape.vocab_size, shape.d_model
    n_copies = 4  # 4 x 77 MB > 126 MB L2: every launch reads its weights from HBM
    Bh = min(max(B, 1), 256)
    with torch.cuda.stream(stream):
        wcopies = [torch.randn(V, d, device=dev).to(torch.bfloat16) * 0.02 for _ in range(n_copies)]
        act = torch.randn(Bh, d, device=dev).to(torch.bfloat16)
        work = None
        for i in range(4):
            work = ops.lm_head_argmax(act, wcopies[i % n_copies], shape.eos_token_id, None, work)
        reps = 40
        k0 = torch.cuda.Event(enable_timing=True)
        k1 = torch.cuda.Event(enable_timing=True)
        stream.synchronize()
        k0.record(stream)
        for i in range(reps):
            ops.lm_head_argmax(act, wcopies[i % n_copies], shape.eos_token_id, None, work)
        k1.record(stream)
        stream.synchronize()
    head_ms = k0.elapsed_time(k1) / reps
    head_bytes = V * d * 2 + Bh * d * 2
    del wcopies

    # ---- reduce over ranks ----
    times = torch.tensor([total_ms, e2e_s * 1e3, prefill_ms, decode_ms, head_ms], device=dev, dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
        lt = torch.tensor([launches], device=dev, dtype=torch.int64)
        dist.all_reduce(lt)
        launches = int(lt.item())
    total_ms, e2e_ms, prefill_ms, decode_ms, head_ms = times.tolist()

    peaks, peak_src = measured_peaks()
    tokens_per_step = B * N * world
    value = tokens_per_step * args.steps / (total_ms / 1e3)
    e2e_value = tokens_per_step * args.steps / (e2e_ms / 1e3)
    avg_ctx = T + (N - 1) / 2.0
    step_bytes = decode_step_bytes(shape, B, int(round(avg_ctx)))
    decode_step_ms = decode_ms / (N - 1)
    result = {
        "metric": "generated tokens/sec (GPT-2 small, batch 64) at 1/2/4/8 B200; p50 SmartReply ms",
        "value": value,
        "unit": "tokens/s",
        "n_gpus": world,
        "steps": args.steps,
        "warmup": max(args.warmup, 3),
        "ms_per_step": total_ms / args.steps,
        "higher_is_better": True,
        "scaling": "weak",
        "vs_baseline": None,
        "dtype": "bf16",
        "data": "synthetic",
        "config": {
            "workload": f"{args.workload}: GPT-2 {model} ({shape.n_layer}L d={shape.d_model}), batch {B} per GPU, "
                        f"{T}-token prompts, {N} new tokens, greedy ({BASELINE_CONFIG.get(args.workload, 'profiling probe')})",
            "weights": "random init seed 0 (HF init scheme), bf16 GEMM weights, fp32 residual stream",
            "parallelism": f"request-sharded replicas x{world}, no data-path collective",
            "l2": f"no flush: each decode step streams {step_bytes / 1e6:.0f} MB of weights + KV (> 126 MB L2)",
            "prefill_ms": prefill_ms,
            "decode_ms_per_token_step": decode_step_ms,
        },
        "e2e": {
            "value": e2e_value,
            "unit": "tokens/s",
            "h2d_bytes_per_step": int(B * T * 4 + (2 * B + 1) * 4 + B * 32),
            "d2h_bytes_per_step": int(B * N * 4 + 2 * B * 4 * 2),
            "api": "Engine.generate -> b200llm_generate (host ids in, host tokens out)",
        },
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": {
            "bound": "hbm",
            "kernel": "gemm_kernel<kEpiLogitsArgmax, swap-AB, balanced rows> -- the LM head with fused greedy argmax "
                      "as the decode step launches it (b200llm_op_lm_head_argmax); the longest single kernel of the "
                      "step but only ~5 % of it: decode_step_roofline below is the whole-step figure",
            "achieved": head_bytes / (head_ms * 1e-3) / 1e9,
            "peak": peaks["hbm_gbs"],
            "unit": "GB/s",
            "frac": head_bytes / (head_ms * 1e-3) / 1e9 / peaks["hbm_gbs"],
            "peak_source": f"{peak_src} (MEASURED_PEAKS.json hbm_gbs)",
            "bytes_per_launch": head_bytes,
            "us_per_launch": head_ms * 1e3,
            "traffic": lmhead_traffic_from_profile()[0],
            "traffic_source": f"{lmhead_traffic_from_profile()[1]} (ncu --set full, B = 64)",
            "algorithmic_bytes": "V*d*2 (tied head weights, bf16) + B*d*2 (activations); no logits are stored",
        },
        "decode_step_roofline": {
            "bound": "hbm",
            "bytes_per_step": step_bytes,
            "us_per_step": decode_step_ms * 1e3,
            "achieved": step_bytes / (decode_step_ms * 1e-3) / 1e9,
            "peak": peaks["hbm_gbs"],
            "unit": "GB/s",
            "frac": step_bytes / (decode_step_ms * 1e-3) / 1e9 / peaks["hbm_gbs"],
        },
    }
    eng.close()
    if rank == 0 and world == 1 and not args.no_decode_classes:
        result["decode_classes"] = decode_classes(args.workload, peaks["hbm_gbs"])
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # what was timed is also what is right: the first batch's tokens against the CPU oracle
        result["parity_check"] = parity_check(shape, blob, prompts_np, np.array(out), N)
    if rank == 0 and model == "small" and not args.no_smart_reply:
        result["smart_reply"] = smart_reply_latency(E, shape, blob, local,
                                                    cpu_calls=0 if (args.no_cpu_baseline or world > 1) else 3)
        # the served configuration (server.build_local_model) keeps prompt pages in the prefix cache;
        # with ONE fixed conversation every call after the first finds all but its last page cached,
        # so this is the best case of the cache, reported beside the cold-prefill number above
        warm = smart_reply_latency(E, shape, blob, local, n_calls=100, cpu_calls=0, prefix_cache=True)
        result["smart_reply"]["p50_ms_prefix_cache_same_conversation"] = warm["p50_ms"]
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        result["cpu_baseline"] = cpu_baseline(args.workload, budget_s=25.0)
    if rank == 0:
        print(json.dumps(result))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def decode_classes(workload: str, hbm_gbs: float) -> dict | None:
    """In-graph cost of every kernel class of the decode step, read by leaving the class out
    (B200LLM_SKIP, honoured only by libb200llm_tune.so) -- profiles/decode_classes.py in a
    subprocess so that the product library of this process never sees a tuning switch."""
    tune = os.path.join(ROOT, PKG, "libb200llm_tune.so")
    if not os.path.exists(tune):
        return {"unavailable": "libb200llm_tune.so not built (python __graft_entry__.py build)"}
    env = dict(os.environ, B200LLM_LIB=tune)
    try:
        r = subprocess.run([sys.executable, os.path.join(ROOT, "profiles", "decode_classes.py"), workload,
                            str(hbm_gbs)], env=env, capture_output=True, text=True, timeout=600)
        line = [l for l in r.stdout.splitlines() if l.startswith("{")]
        return json.loads(line[-1]) if r.returncode == 0 and line else {"unavailable": (r.stderr or r.stdout)[-300:]}
    except Exception as exc:  # noqa: BLE001
        return {"unavailable": repr(exc)}


# ------------------------------------------------------------------------------------------
# CPU legs (the only place bench.py touches oracle/)
# ------------------------------------------------------------------------------------------
def parity_check(shape, blob, prompts_np, tokens, N: int, rows: int = 4, margin: float = 0.1) -> dict:
    """Checker only (never timed): the oracle, teacher-forced with the engine's own tokens of the
    first batch, must pick the same token wherever its top-1/top-2 margin is decisive."""
    import torch

    from oracle import gpt2_oracle as O

    oshape = O.GPT2Shape(n_layer=shape.n_layer, n_head=shape.n_head, d_model=shape.d_model,
                         vocab_size=shape.vocab_size, n_positions=shape.n_positions, ln_eps=shape.ln_eps,
                         eos_token_id=shape.eos_token_id)
    w = O.unpack_blob(oshape, blob)
    sel = np.linspace(0, prompts_np.shape[0] - 1, min(rows, prompts_np.shape[0])).astype(int)
    steps = min(N, 32)
    teacher = torch.from_numpy(tokens[sel][:, :steps]).long()
    ref, _, margins = O.greedy_generate(oshape, w, torch.from_numpy(prompts_np[sel]).long(), steps,
                                        min_new_tokens=steps, teacher=teacher)
    agree = (ref == teacher).numpy()
    decisive = margins.numpy() > margin
    ok = bool(agree[decisive].all())
    res = {"rows": sel.tolist(), "steps": steps, "agreement": float(agree.mean()),
           "decisive_fraction": float(decisive.mean()), "decisive_margin": margin, "ok": ok,
           "checker": "oracle.greedy_generate teacher-forced with the engine's tokens"}
    if not ok:
        raise SystemExit(f"bench.py: the timed batch disagrees with the oracle at a decisive margin: {res}")
    return res


def _cpu_setup(workload: str):
    import torch

    from oracle import gpt2_oracle as O
    from oracle import hf_reference as H

    model, B, T, N = WORKLOADS[workload]
    shape = O.SHAPES[model]
    if T + N > shape.n_positions:
        shape = O.GPT2Shape(**{**shape.__dict__, "n_positions": 2048})
    torch.manual_seed(0)
    hf = H.build_hf_model(shape, None, attn="sdpa")  # HF's own random init; timing only
    ids = O.synthetic_prompts(B, T)
    return torch, H, hf, ids, (model, B, T, N)


def cpu_baseline(workload: str, budget_s: float) -> dict:
    """transformers GPT2LMHeadModel.generate, fp32, all host cores, on a bounded sample."""
    torch, H, hf, ids, (model, B, T, N) = _cpu_setup(workload)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    # probe with 1/8 of the batch, then take the largest batch slice that fits the budget
    probe_b = max(1, B // 8)
    r = H.time_hf_generate(hf, ids[:probe_b], N, threads=cores)
    per_seq = r["seconds"] / probe_b
    sample_b = int(min(B, max(probe_b, budget_s / max(per_seq, 1e-6))))
    if sample_b > probe_b:
        r = H.time_hf_generate(hf, ids[:sample_b], N, threads=cores)
    else:
        sample_b = probe_b
    return {"value": r["tokens_per_s"], "unit": "tokens/s", "cores": cores, "kind": "port",
            "sample": f"transformers {__import__('transformers').__version__} GPT2LMHeadModel.generate fp32 (sdpa) on "
                      f"{sample_b} of {B} sequences, {T}-token prompts, {N} new tokens, {r['seconds']:.1f} s, "
                      "1 run (library pinned by reference requirements.txt:6; the reference has no local path)"}


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    torch, H, hf, ids, (model, B, T, N) = _cpu_setup(args.workload)
    shape = hf.config
    shape.n_layer, shape.d_model = hf.config.n_layer, hf.config.n_embd
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    # size the per-step sample so that (K + W) steps end within ~150 s
    probe_b = max(1, B // 8)
    r = H.time_hf_generate(hf, ids[:probe_b], N, threads=cores)  # doubles as extra warm-up
    per_seq = r["seconds"] / probe_b
    total_steps = args.steps + args.warmup
    sample_b = int(min(B, max(1, 150.0 / max(total_steps, 1) / max(per_seq, 1e-6))))
    sample = ids[:sample_b]
    for _ in range(args.warmup):
        H.hf_generate(hf, sample, N, min_new_tokens=N)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        H.hf_generate(hf, sample, N, min_new_tokens=N)
    dt = time.perf_counter() - t0
    value = sample_b * N * args.steps / dt
    desc = (f"transformers GPT2LMHeadModel.generate fp32 (sdpa), {sample_b} of {B} sequences per step, "
            f"{T}-token prompts, {N} new tokens")
    print(json.dumps({
        "impl": "reference",
        "metric": "generated tokens/sec (GPT-2 small, batch 64) at 1/2/4/8 B200; p50 SmartReply ms",
        "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: GPT-2 {model} ({shape.n_layer}L d={shape.d_model}), batch {B} per GPU, "
                               f"{T}-token prompts, {N} new tokens, greedy ({BASELINE_CONFIG.get(args.workload, 'profiling probe')})",
                   "note": "CPU reference arm (transformers on the host cores); rank 0 only"},
        "cpu_baseline": {"value": value, "unit": "tokens/s", "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": value, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="smart_reply_b64", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-smart-reply", action="store_true", help="skip the p50 SmartReply gRPC measurement")
    ap.add_argument("--no-decode-classes", action="store_true", help="skip the per-class decode ablation")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
