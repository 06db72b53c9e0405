#!/usr/bin/env python
"""BASELINE.json configs[4]: a mixed-length SmartReply / Summarize / Ask request stream through the
gRPC llm.LLMService, served by the local B200 engine on 1..N GPUs of one box.

    python bench_stream.py [--gpus N] [--requests R] [--rate REQ_PER_S] [--client-procs P] [--cpu-sample K]

Stream: request kinds in ratio 6:1:3 (SmartReply : Summarize : Ask), Poisson arrivals
(seed 99), conversations synthesised from a fixed sentence pool. Every request travels the whole
drop-in path: gRPC stub -> LLMServicer (reference prompts/parsers) -> LocalGenerativeModel ->
router -> per-GPU continuous-batching worker -> CUDA engine. Replicas only: the router shards
requests, there is no collective. Prints ONE JSON line: p50/p95 latency per RPC, aggregate
generated tokens/s, and (optionally) the same RPCs timed on the host CPU with transformers for a
small sample. The callers run in separate processes (`--client-procs`, like the reference's remote
Raft nodes) so that their Python does not share an interpreter lock with the server.
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
from concurrent import futures

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "distributed-real-time-chat-and-collaboration-tool_b200"

SENTENCES = [
    "Are we still on for the design review at 3pm?", "Yes, I booked room 4B.", "Bring the latency numbers.",
    "I can't make it before 3:30, start without me.", "Can you share the doc beforehand?", "Sure, sending it now.",
    "The benchmark finished overnight.", "Throughput is up but p95 regressed a little.", "Let's look at the traces.",
    "Who owns the rollout plan?", "I will draft it today.", "Please add the rollback steps.", "Sounds good to me.",
    "The cache hit rate dropped after the deploy.", "Could be the new key format.", "I'll check the dashboards.",
]
USERS = ["alice", "bob", "charlie", "dave"]


def make_requests(pb, n: int, seed: int = 99):
    rng = np.random.default_rng(seed)
    kinds = rng.choice(["smart", "summarize", "ask"], size=n, p=[0.6, 0.1, 0.3])
    reqs = []
    for i, kind in enumerate(kinds):
        def msgs(k):
            return [pb.Message(sender=USERS[int(rng.integers(len(USERS)))],
                               content=SENTENCES[int(rng.integers(len(SENTENCES)))]) for _ in range(k)]
        if kind == "smart":
            reqs.append(("GetSmartReply", pb.SmartReplyRequest(request_id=str(i), recent_messages=msgs(5), user_id="bob")))
        elif kind == "summarize":
            reqs.append(("SummarizeConversation",
                         pb.SummarizeRequest(request_id=str(i), messages=msgs(int(rng.integers(8, 14))), max_length=200)))
        else:
            ctx = [f"{m.sender}: {m.content}" for m in msgs(3)]
            reqs.append(("GetLLMAnswer", pb.LLMRequest(request_id=str(i), query="What did we decide about the rollout?",
                                                        context=ctx)))
    return reqs


class CountingModel:
    """Wraps the generator to total the generated tokens (the RPC responses do not carry them)."""

    def __init__(self, inner):
        self.inner = inner
        self.lock = threading.Lock()
        self.generated = 0
        self.prompt = 0

    def generate_content(self, prompt, generation_config=None):
        resp = self.inner.generate_content(prompt, generation_config=generation_config)
        with self.lock:
            self.generated += resp.usage["generated_tokens"]
            self.prompt += resp.usage["prompt_tokens"]
        return resp


def run_calls(stub_factory, reqs, indices, arrivals, threads, t_zero):
    """Issue reqs[i] for i in indices from `threads` caller threads; returns (latencies, t_first, t_last)
    with wall-clock (time.time) stamps that are comparable across processes."""
    lat: dict[str, list[float]] = {"GetSmartReply": [], "SummarizeConversation": [], "GetLLMAnswer": []}
    lock = threading.Lock()
    stub = stub_factory()
    stamps = []

    def call(i):
        rpc, req = reqs[i]
        delay = t_zero + arrivals[i] - time.time()
        if delay > 0:
            time.sleep(delay)
        t0 = time.time()
        getattr(stub, rpc)(req, timeout=60)
        t1 = time.time()
        with lock:
            lat[rpc].append((t1 - t0) * 1e3)
            stamps.append((t0, t1))

    with futures.ThreadPoolExecutor(max_workers=threads) as pool:
        list(pool.map(call, indices))
    return lat, min(s[0] for s in stamps), max(s[1] for s in stamps)


def client_main(args):
    """--as-client: one caller process. Prints one JSON line with its latencies and time stamps."""
    import grpc

    server_mod = importlib.import_module(f"{PKG}.server")
    pb = importlib.import_module(f"{PKG}.llm_proto")
    reqs = make_requests(pb, args.requests)
    rng = np.random.default_rng(99)
    arrivals = np.cumsum(rng.exponential(1.0 / args.rate, size=len(reqs))) if args.rate > 0 else np.zeros(len(reqs))
    indices = list(range(args.client_index, len(reqs), args.client_procs))
    # several TCP connections per caller process: with more than one serving process the kernel
    # balances CONNECTIONS over the listeners (SO_REUSEPORT), not requests
    chans = [grpc.insecure_channel(f"127.0.0.1:{args.port}", options=(("grpc.use_local_subchannel_pool", 1),))
             for _ in range(max(1, args.connections))]
    stubs = [server_mod.LLMServiceStub(ch) for ch in chans]

    class RoundRobin:
        def __init__(self):
            self.n = 0

        def __getattr__(self, rpc):
            self.n += 1
            return getattr(stubs[self.n % len(stubs)], rpc)

    lat, t_first, t_last = run_calls(RoundRobin, reqs, indices, arrivals,
                                     max(1, args.clients // args.client_procs), args.t_zero)
    for ch in chans:
        ch.close()
    print(json.dumps({"lat": lat, "t_first": t_first, "t_last": t_last}))


def growing_channels(pb, server_mod, port: int, channels: int, first: int, last: int) -> dict:
    """Summaries of growing channels (the call pattern of reference server/raft_node.py:2032-2084: the
    Raft node sends the channel's recent messages, in order, on every summarise request). Each of
    `channels` callers owns one channel, appends one message at a time and asks for a summary after
    each, so consecutive prompts of a channel share everything but their last message."""
    import grpc

    rng = np.random.default_rng(7)
    lat: list[float] = []
    lock = threading.Lock()

    def one_channel(c: int):
        r = np.random.default_rng(1000 + c)
        msgs = [pb.Message(sender=USERS[int(r.integers(len(USERS)))], content=f"[ch{c}] " + SENTENCES[int(r.integers(len(SENTENCES)))])
                for _ in range(last)]
        with grpc.insecure_channel(f"127.0.0.1:{port}") as ch:
            stub = server_mod.LLMServiceStub(ch)
            for n in range(first, last + 1):
                req = pb.SummarizeRequest(request_id=f"{c}-{n}", messages=msgs[:n], max_length=200)
                t0 = time.perf_counter()
                stub.SummarizeConversation(req, timeout=60)
                with lock:
                    lat.append((time.perf_counter() - t0) * 1e3)

    del rng
    t0 = time.perf_counter()
    with futures.ThreadPoolExecutor(max_workers=channels) as pool:
        list(pool.map(one_channel, range(channels)))
    wall = time.perf_counter() - t0
    return {"channels": channels, "messages_per_channel": [first, last], "requests": len(lat), "wall_s": wall,
            "p50_ms": float(np.percentile(lat, 50)), "p95_ms": float(np.percentile(lat, 95))}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--requests", type=int, default=400)
    ap.add_argument("--rate", type=float, default=0.0, help="Poisson arrival rate in requests/s (0 = closed loop)")
    ap.add_argument("--clients", type=int, default=64, help="concurrent callers (threads over all caller processes)")
    ap.add_argument("--client-procs", type=int, default=4, help="caller processes (0 = callers share this process)")
    ap.add_argument("--cpu-sample", type=int, default=0, help="also time this many requests per RPC on the host CPU")
    ap.add_argument("--serving-procs", type=int, default=1,
                    help="serving processes sharing the port (server.serve_many), each with gpus / serving-procs GPUs")
    ap.add_argument("--connections", type=int, default=1, help="TCP connections per caller process")
    ap.add_argument("--scenario", default="mixed", choices=["mixed", "growing_channel"],
                    help="mixed: BASELINE configs[4]; growing_channel: consecutive summaries of growing channels")
    ap.add_argument("--channels", type=int, default=8)
    ap.add_argument("--process-per-gpu", type=int, default=-1,
                    help="1 = one worker process per GPU behind the router, 0 = worker threads, -1 = server default")
    ap.add_argument("--prefix-cache", type=int, default=1, help="shared-prefix KV reuse in the engines (1 = as served)")
    ap.add_argument("--switch-interval", type=float, default=0.0, help="sys.setswitchinterval for the server process (0 = leave)")
    ap.add_argument("--as-client", action="store_true", help=argparse.SUPPRESS)
    ap.add_argument("--port", type=int, default=0, help=argparse.SUPPRESS)
    ap.add_argument("--client-index", type=int, default=0, help=argparse.SUPPRESS)
    ap.add_argument("--t-zero", type=float, default=0.0, help=argparse.SUPPRESS)
    args = ap.parse_args()
    if args.as_client:
        client_main(args)
        return

    import grpc

    if args.serving_procs > 1:
        many_main(args)
        return
    if args.switch_interval > 0:
        sys.setswitchinterval(args.switch_interval)
    server_mod = importlib.import_module(f"{PKG}.server")
    svc = importlib.import_module(f"{PKG}.llm_service")
    pb = importlib.import_module(f"{PKG}.llm_proto")

    model = CountingModel(server_mod.build_local_model("small", args.gpus, max_seqs=64, max_seq_len=1024,
                                                       max_new_tokens=160, prefix_cache=bool(args.prefix_cache),
                                                       process_per_gpu=None if args.process_per_gpu < 0
                                                       else bool(args.process_per_gpu)))
    model.inner.default_max_new_tokens = 32        # SmartReply 32; Summarize overrides below; Ask passes 150
    servicer = svc.LLMServicer(model)
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max(args.clients, 8) + 8))
    server_mod.add_llm_service(servicer, server)
    port = server.add_insecure_port("127.0.0.1:0")
    server.start()

    reqs = make_requests(pb, args.requests)
    if args.scenario == "growing_channel":
        with grpc.insecure_channel(f"127.0.0.1:{port}") as ch:
            stub = server_mod.LLMServiceStub(ch)
            for rpc, req in reqs[:8]:
                getattr(stub, rpc)(req, timeout=60)
        base = [w.snapshot()["prefix_cache"] or {"lookup_tokens": 0, "hit_tokens": 0} for w in model.inner.router.workers]
        res = growing_channels(pb, server_mod, port, args.channels, 3, 12)
        server.stop(0)
        cache = [w.snapshot()["prefix_cache"] or {"lookup_tokens": 0, "hit_tokens": 0} for w in model.inner.router.workers]
        model.inner.router.shutdown()
        look = sum(c["lookup_tokens"] for c in cache) - sum(c["lookup_tokens"] for c in base)
        hit = sum(c["hit_tokens"] for c in cache) - sum(c["hit_tokens"] for c in base)
        print(json.dumps({"metric": "SummarizeConversation latency for growing channels (shared-prefix KV reuse)",
                          "n_gpus": args.gpus, "prefix_cache": {"enabled": bool(args.prefix_cache), "lookup_tokens": look,
                                                                 "hit_tokens": hit, "hit_rate": hit / max(1, look)},
                          **res, "new_tokens": 32,
                          "tokenizer": "byte-level GPT-2 ids: prompts of 350-900 tokens", "data": "synthetic conversations, random-init GPT-2 small"}))
        return
    rng = np.random.default_rng(99)
    arrivals = np.cumsum(rng.exponential(1.0 / args.rate, size=len(reqs))) if args.rate > 0 else np.zeros(len(reqs))
    with grpc.insecure_channel(f"127.0.0.1:{port}") as ch:
        stub = server_mod.LLMServiceStub(ch)
        for rpc, req in reqs[:8]:  # warm-up (graph capture, tokenizer caches)
            getattr(stub, rpc)(req, timeout=60)
        model.generated = model.prompt = 0
        if args.client_procs <= 0:
            lat, t_first, t_last = run_calls(lambda: stub, reqs, range(len(reqs)), arrivals, args.clients, time.time())
        else:
            t_zero = time.time() + 10.0  # callers import grpc / build the descriptors first, then start together
            procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--as-client", "--port", str(port),
                                       "--requests", str(args.requests), "--rate", str(args.rate),
                                       "--clients", str(args.clients), "--client-procs", str(args.client_procs),
                                       "--client-index", str(i), "--t-zero", repr(t_zero)],
                                      stdout=subprocess.PIPE, text=True) for i in range(args.client_procs)]
            lat = {"GetSmartReply": [], "SummarizeConversation": [], "GetLLMAnswer": []}
            t_first, t_last = float("inf"), 0.0
            for pr in procs:
                out, _ = pr.communicate(timeout=600)
                if pr.returncode != 0:
                    raise RuntimeError(f"caller process failed with code {pr.returncode}")
                d = json.loads(out.strip().splitlines()[-1])
                for k, v in d["lat"].items():
                    lat[k].extend(v)
                t_first, t_last = min(t_first, d["t_first"]), max(t_last, d["t_last"])
        wall = t_last - t_first
    server.stop(0)
    workers = model.inner.router.workers
    snaps = [w.snapshot() for w in workers]
    worker_phases = [{**s["phase_s"], "iterations": s["iterations"], "completed": s["completed"],
                      "prefix_hit_tokens": s["prefix_hit_tokens"]} for s in snaps]
    cache = [s["prefix_cache"] or {"lookup_tokens": 0, "hit_tokens": 0} for s in snaps]
    per_process = type(workers[0]).__name__ == "RemoteEngineWorker"
    model.inner.router.shutdown()

    out = {
        "metric": "mixed SmartReply/Summarize/Ask stream through gRPC LLMService (BASELINE configs[4])",
        "n_gpus": args.gpus, "requests": len(reqs), "arrival_rate_rps": args.rate or None, "clients": args.clients,
        "client_processes": args.client_procs,
        "wall_s": wall, "requests_per_s": len(reqs) / wall,
        "generated_tokens_per_s": model.generated / wall, "prompt_tokens_per_s": model.prompt / wall,
        "latency_ms": {k: {"n": len(v), "p50": float(np.percentile(v, 50)), "p95": float(np.percentile(v, 95))}
                       for k, v in lat.items() if v},
        "worker_loop_s": worker_phases,
        "process_model": (f"one serving process (gRPC, prompts, tokenizer, router) + {args.gpus} engine worker "
                          f"{'process(es), one per GPU' if per_process else 'thread(s) in the same process'}"),
        "prefix_cache": {"enabled": bool(args.prefix_cache),
                         "lookup_tokens": sum(c["lookup_tokens"] for c in cache),
                         "hit_tokens": sum(c["hit_tokens"] for c in cache),
                         "hit_rate": sum(c["hit_tokens"] for c in cache) / max(1, sum(c["lookup_tokens"] for c in cache))},
        "tokenizer": "byte-level GPT-2 ids (no vocab files offline): prompts are ~4x longer than with BPE",
        "data": "synthetic conversations, random-init GPT-2 small",
    }
    if args.cpu_sample > 0:
        out["cpu_latency_ms"] = cpu_sample(pb, svc, reqs, args.cpu_sample)
    print(json.dumps(out))


def many_main(args):
    """Several serving processes on one port (server.serve_many): the callers are the same caller
    processes as in the single-server run, each with `--connections` TCP connections."""
    import signal
    import socket
    import tempfile

    import grpc

    server_mod = importlib.import_module(f"{PKG}.server")
    pb = importlib.import_module(f"{PKG}.llm_proto")
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    stats_dir = tempfile.mkdtemp(prefix="b200llm_stream_")
    extra = ["--workers", str(max(args.clients, 8) + 8), "--max-seqs", "64", "--max-new-tokens", "160",
             "--default-new-tokens", "32", "--prefix-cache", str(args.prefix_cache)]
    servers = server_mod.serve_many("small", args.gpus, args.serving_procs, port, extra, stats_dir)
    try:
        deadline = time.time() + 600
        ready = [os.path.join(stats_dir, f"serving_{i}.json.ready") for i in range(args.serving_procs)]
        while not all(os.path.exists(r) for r in ready):
            if time.time() > deadline or any(p.poll() is not None for p in servers):
                raise RuntimeError("a serving process did not come up")
            time.sleep(0.5)
        reqs = make_requests(pb, args.requests)
        # warm-up over many connections so that every listener captures its graphs
        for _ in range(4 * args.serving_procs):
            with grpc.insecure_channel(f"127.0.0.1:{port}", options=(("grpc.use_local_subchannel_pool", 1),)) as ch:
                stub = server_mod.LLMServiceStub(ch)
                for rpc, req in reqs[:6]:
                    getattr(stub, rpc)(req, timeout=120)
        t_zero = time.time() + 10.0
        procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--as-client", "--port", str(port),
                                   "--requests", str(args.requests), "--rate", str(args.rate),
                                   "--clients", str(args.clients), "--client-procs", str(args.client_procs),
                                   "--connections", str(args.connections),
                                   "--client-index", str(i), "--t-zero", repr(t_zero)],
                                  stdout=subprocess.PIPE, text=True) for i in range(args.client_procs)]
        lat = {"GetSmartReply": [], "SummarizeConversation": [], "GetLLMAnswer": []}
        t_first, t_last = float("inf"), 0.0
        for pr in procs:
            out, _ = pr.communicate(timeout=900)
            if pr.returncode != 0:
                raise RuntimeError(f"caller process failed with code {pr.returncode}")
            d = json.loads(out.strip().splitlines()[-1])
            for k, v in d["lat"].items():
                lat[k].extend(v)
            t_first, t_last = min(t_first, d["t_first"]), max(t_last, d["t_last"])
    finally:
        for p in servers:
            p.send_signal(signal.SIGTERM)
        for p in servers:
            try:
                p.wait(timeout=60)
            except subprocess.TimeoutExpired:
                p.kill()
    wall = t_last - t_first
    stats = []
    for i in range(args.serving_procs):
        path = os.path.join(stats_dir, f"serving_{i}.json")
        if os.path.exists(path):
            stats.append(json.load(open(path)))
    gen = sum(s_["generated_tokens"] for s_ in stats)
    warm = 4 * args.serving_procs * 6
    print(json.dumps({
        "metric": "mixed SmartReply/Summarize/Ask stream through gRPC LLMService (BASELINE configs[4])",
        "n_gpus": args.gpus, "requests": len(reqs), "clients": args.clients, "client_processes": args.client_procs,
        "connections_per_client_process": args.connections,
        "process_model": f"{args.serving_procs} serving processes on one port (SO_REUSEPORT), each with "
                         f"{args.gpus // args.serving_procs} GPU worker process(es) behind its router",
        "wall_s": wall, "requests_per_s": len(reqs) / wall,
        "generated_tokens_per_s_incl_warmup_tokens": gen / wall,
        "requests_per_serving_process": [s_["requests"] for s_ in stats], "warmup_requests": warm,
        "latency_ms": {k: {"n": len(v), "p50": float(np.percentile(v, 50)), "p95": float(np.percentile(v, 95))}
                       for k, v in lat.items() if v},
        "completed_per_gpu": [w["completed"] for s_ in stats for w in s_["workers"]],
        "tokenizer": "byte-level GPT-2 ids (no vocab files offline): prompts are ~4x longer than with BPE",
        "data": "synthetic conversations, random-init GPT-2 small"}))


def cpu_sample(pb, svc, reqs, k: int) -> dict:
    """The same RPC bodies on the host cores: transformers GPT2LMHeadModel.generate behind the same
    servicer code (checker-side only; uses oracle/)."""
    import torch

    from oracle import gpt2_oracle as O
    from oracle import hf_reference as H

    tok = importlib.import_module(f"{PKG}.tokenizer").ByteLevelTokenizer()
    lm = importlib.import_module(f"{PKG}.local_model")
    torch.manual_seed(0)
    torch.set_num_threads(os.cpu_count() or 1)
    shape = O.GPT2Shape(**{**O.SMALL.__dict__, "n_positions": 2048})
    hf = H.build_hf_model(shape, None, attn="sdpa")

    class HfModel:
        def generate_content(self, prompt, generation_config=None):
            n = int((generation_config or {}).get("max_output_tokens", 32))
            ids = torch.tensor([tok.encode(prompt)[-860:]])
            toks = H.hf_generate(hf, ids, n, min_new_tokens=n)[0].tolist()
            return lm.GenerateContentResponse(candidates=[lm.Candidate(2, lm.Content([lm.Part(tok.decode(toks))]))])

    servicer = svc.LLMServicer(HfModel())
    res: dict[str, list[float]] = {}
    for rpc, req in reqs:
        if len(res.setdefault(rpc, [])) >= k:
            continue
        t0 = time.perf_counter()
        getattr(servicer, rpc)(req, None)
        res[rpc].append((time.perf_counter() - t0) * 1e3)
        if all(len(v) >= k for v in res.values()) and len(res) == 3:
            break
    return {r: {"n": len(v), "p50": float(np.percentile(v, 50))} for r, v in res.items()}


if __name__ == "__main__":
    main()
