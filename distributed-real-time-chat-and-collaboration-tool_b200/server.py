"""gRPC bootstrap for llm.LLMService backed by the local B200 engine.

Replaces the reference's serve() (llm_server.py:476-522). Differences forced by the reference's
own defects: all FOUR RPCs are registered (the shipped
add_LLMServiceServicer_to_server registers two), and the handler pool is larger than 4 threads so
that batches can form. Port 50055 and the insecure channel are kept.

    python -m b200llm.server --model small --gpus 1            
"""
from __future__ import annotations

import argparse
import logging
from concurrent import futures

import grpc

from . import llm_proto as pb
from .llm_service import LLMServicer

logger = logging.getLogger(__name__)
DEFAULT_PORT = 50055  # llm_server.py:507


def add_llm_service(servicer, server: grpc.Server) -> None:
    """Register every llm.LLMService method on `server` (generic handlers, no generated code)."""
    handlers = {}
    for rpc, (req_name, resp_name) in pb.METHODS.items():
        handlers[rpc] = grpc.unary_unary_rpc_method_handler(
            getattr(servicer, rpc),
            request_deserializer=pb.MESSAGE_CLASSES[req_name].FromString,
            response_serializer=pb.MESSAGE_CLASSES[resp_name].SerializeToString)
    server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(f"{pb.PACKAGE}.{pb.SERVICE}", handlers),))


class LLMServiceStub:
    """Client stub with the reference stub's surface (generated/llm_service_pb2_grpc.py:59-78)."""

    def __init__(self, channel: grpc.Channel):
        for rpc, (req_name, resp_name) in pb.METHODS.items():
            setattr(self, rpc, channel.unary_unary(
                pb.method_path(rpc),
                request_serializer=pb.MESSAGE_CLASSES[req_name].SerializeToString,
                response_deserializer=pb.MESSAGE_CLASSES[resp_name].FromString))


def build_local_model(model: str = "small", gpus: int = 1, *, max_seqs: int = 64, max_seq_len: int = 1024,
                      max_new_tokens: int = 256, weights: str | None = None, vocab_json: str | None = None,
                      merges_txt: str | None = None, seed: int = 0, prefix_cache: bool = True,
                      process_per_gpu: bool | None = None):
    """Engines (one per GPU) + workers + router + tokenizer -> LocalGenerativeModel.

    process_per_gpu: every GPU's engine and scheduler loop in a process of its own behind the router
    (remote_worker.py) instead of a thread of this process. Default: threads for one GPU, processes
    for several (with threads the workers and the gRPC handlers share one interpreter lock; measured
    in profiles/r2_stream_*)."""
    from .engine import MODEL_SHAPES, Engine
    from .local_model import LocalGenerativeModel
    from .scheduler import EngineWorker, Router
    from .tokenizer import load_tokenizer
    from .weights import blob_from_safetensors, synthetic_blob

    shape = MODEL_SHAPES[model]
    engine_kwargs = dict(max_seqs=max_seqs, max_seq_len=min(max_seq_len, shape.n_positions), max_batch_tokens=8192,
                         max_new_tokens=max_new_tokens, prefix_cache=prefix_cache)
    if process_per_gpu is None:
        process_per_gpu = gpus > 1
    workers = []
    if process_per_gpu:
        from .remote_worker import RemoteEngineWorker

        workers = [RemoteEngineWorker(dev, model, engine_kwargs, weights=weights, seed=seed) for dev in range(gpus)]
    else:
        blob = blob_from_safetensors(shape, weights) if weights else synthetic_blob(shape, seed=seed)
        for dev in range(gpus):
            eng = Engine(shape, blob, device=dev, **engine_kwargs)
            workers.append(EngineWorker(eng, name=f"b200llm-gpu{dev}"))
    router = Router(workers)
    return LocalGenerativeModel(router, load_tokenizer(vocab_json, merges_txt),
                                max_prompt_tokens=min(max_seq_len, shape.n_positions) - max_new_tokens - 1)


def create_server(servicer, port: int = DEFAULT_PORT, max_workers: int = 128, host: str = "[::]"):
    # SO_REUSEPORT (gRPC's default on Linux, stated here because serve_many relies on it): several
    # serving processes may listen on the same port and the kernel spreads the callers' connections
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers), options=(("grpc.so_reuseport", 1),))
    add_llm_service(servicer, server)
    bound = server.add_insecure_port(f"{host}:{port}")
    return server, bound


def serve_many(model: str, gpus: int, serving_procs: int, port: int = DEFAULT_PORT, extra_args: list[str] | None = None,
               stats_dir: str | None = None):
    """One serving process per `gpus / serving_procs` GPUs, all listening on `port`.

    A single Python serving process (gRPC handlers, prompts, tokenizer, router, futures) saturates at
    about 2.5 k requests/s -- four GPUs' worth of the mixed stream (profiles/r2_stream_*). Each child
    is a complete server (`serve`) that sees only its own GPUs through CUDA_VISIBLE_DEVICES; the kernel
    balances the callers' connections over the listeners (SO_REUSEPORT), so callers should hold
    several connections. Returns the child processes (terminate them to stop serving)."""
    import os
    import subprocess
    import sys

    if gpus % serving_procs != 0:
        raise ValueError("gpus must be a multiple of serving_procs")
    per = gpus // serving_procs
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    procs = []
    for i in range(serving_procs):
        env = dict(os.environ, CUDA_VISIBLE_DEVICES=",".join(str(d) for d in range(i * per, (i + 1) * per)))
        code = ("import sys, importlib; sys.path.insert(0, %r); "
                "importlib.import_module(%r).serve(sys.argv[1:])" % (root, __name__))
        cmd = [sys.executable, "-c", code, "--model", model, "--gpus", str(per), "--port", str(port)] + (extra_args or [])
        if stats_dir:
            cmd += ["--stats-file", os.path.join(stats_dir, f"serving_{i}.json")]
        procs.append(subprocess.Popen(cmd, env=env))
    return procs


def serve(argv=None) -> None:
    ap = argparse.ArgumentParser(description="llm.LLMService on the local B200 engine")
    ap.add_argument("--model", default="small", choices=["small", "medium", "large"])
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--port", type=int, default=DEFAULT_PORT)
    ap.add_argument("--workers", type=int, default=128)
    ap.add_argument("--weights", default=None, help=".safetensors of a GPT-2 state dict (default: random init)")
    ap.add_argument("--vocab-json", default=None)
    ap.add_argument("--merges-txt", default=None)
    ap.add_argument("--serving-procs", type=int, default=0,
                    help="serving processes sharing the port, each with gpus / serving-procs GPUs (see serve_many); "
                         "0 = one for up to 2 GPUs, one per 2 GPUs beyond (measured: 8 GPUs 6.9 k requests/s with 4 "
                         "serving processes, 2.3 k with one)")
    ap.add_argument("--max-seqs", type=int, default=64)
    ap.add_argument("--max-new-tokens", type=int, default=256)
    ap.add_argument("--default-new-tokens", type=int, default=64)
    ap.add_argument("--prefix-cache", type=int, default=1)
    ap.add_argument("--stats-file", default=None, help="write request / token counters and worker snapshots here on exit")
    args = ap.parse_args(argv)
    logging.basicConfig(level=logging.INFO)
    if args.serving_procs <= 0:
        args.serving_procs = 1 if args.gpus <= 2 or args.gpus % 2 else args.gpus // 2
    if args.serving_procs > 1:
        import signal

        extra = ["--workers", str(args.workers), "--max-seqs", str(args.max_seqs), "--max-new-tokens",
                 str(args.max_new_tokens), "--default-new-tokens", str(args.default_new_tokens),
                 "--prefix-cache", str(args.prefix_cache)]
        for flag, val in (("--weights", args.weights), ("--vocab-json", args.vocab_json), ("--merges-txt", args.merges_txt)):
            if val:
                extra += [flag, val]
        procs = serve_many(args.model, args.gpus, args.serving_procs, args.port, extra)
        try:
            for p in procs:
                p.wait()
        except KeyboardInterrupt:
            for p in procs:
                p.send_signal(signal.SIGTERM)
        return
    model = build_local_model(args.model, args.gpus, max_seqs=args.max_seqs, max_new_tokens=args.max_new_tokens,
                              weights=args.weights, vocab_json=args.vocab_json, merges_txt=args.merges_txt,
                              prefix_cache=bool(args.prefix_cache))
    model.default_max_new_tokens = args.default_new_tokens
    server, port = create_server(LLMServicer(model), args.port, args.workers)
    server.start()
    logger.info("llm.LLMService (B200 local engine, GPT-2 %s x%d GPU) on port %d", args.model, args.gpus, port)
    if args.stats_file:
        open(args.stats_file + ".ready", "w").close()

    def finish(*_):
        server.stop(0)
        if args.stats_file:
            import json

            snaps = [w.snapshot() for w in model.router.workers]
            with open(args.stats_file, "w") as f:
                json.dump({"requests": model.requests, "prompt_tokens": model.prompt_tokens,
                           "generated_tokens": model.generated_tokens, "workers": snaps}, f)
        model.router.shutdown()
        raise SystemExit(0)

    import signal

    signal.signal(signal.SIGTERM, finish)
    try:
        server.wait_for_termination()
    except KeyboardInterrupt:
        finish()


if __name__ == "__main__":
    serve()
