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
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers))
    add_llm_service(servicer, server)
    bound = server.add_insecure_port(f"{host}:{port}")
    return server, bound


def serve(argv=None) -> None:
    ap = argparse.ArgumentParser(description="llm.LLMService on the local B200 engine")
    ap.add_argument("--model", default="small", choices=["small", "medium", "large"])
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--port", type=int, default=DEFAULT_PORT)
    ap.add_argument("--workers", type=int, default=128)
    ap.add_argument("--weights", default=None, help=".safetensors of a GPT-2 state dict (default: random init)")
    ap.add_argument("--vocab-json", default=None)
    ap.add_argument("--merges-txt", default=None)
    args = ap.parse_args(argv)
    logging.basicConfig(level=logging.INFO)
    model = build_local_model(args.model, args.gpus, weights=args.weights, vocab_json=args.vocab_json,
                              merges_txt=args.merges_txt)
    server, port = create_server(LLMServicer(model), args.port, args.workers)
    server.start()
    logger.info("llm.LLMService (B200 local engine, GPT-2 %s x%d GPU) on port %d", args.model, args.gpus, port)
    try:
        server.wait_for_termination()
    except KeyboardInterrupt:
        server.stop(0)
        model.router.shutdown()


if __name__ == "__main__":
    serve()
