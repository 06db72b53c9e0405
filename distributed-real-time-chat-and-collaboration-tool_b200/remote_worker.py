"""One process per GPU behind the router ("one process, one scheduler thread + CUDA
stream per GPU, or one process per GPU behind the router. Either is fine").

With every GPU's EngineWorker as a THREAD of the serving process, the workers' Python (building
batches, ctypes marshalling, resolving futures) and the gRPC handlers share one interpreter lock;
measured on 2 GPUs the stream reached 1.66x of one GPU, not 2x (profiles/r2_stream_*). Here each
GPU gets its own process that owns the Engine and runs the very same EngineWorker loop; the serving
process keeps the gRPC handlers, prompt building, tokenisation and the Router, and talks to the
workers over pipes carrying token ids. RemoteEngineWorker offers the interface the Router uses
(submit / inflight_tokens / healthy / due_for_reprobe / shutdown), so Router and
LocalGenerativeModel are unchanged. Replicas only: nothing crosses between GPUs.
"""
from __future__ import annotations

import itertools
import multiprocessing as mp
import threading
import time
from concurrent.futures import Future
from dataclasses import asdict
from typing import Sequence

import numpy as np

from .engine import SamplingParams
from .scheduler import EngineWorker, GenerationResult


def _worker_main(conn, spec: dict) -> None:
    """Child process: build the engine on spec['device'], run an EngineWorker, serve the pipe."""
    from .engine import MODEL_SHAPES, Engine, ModelShape
    from .weights import blob_from_safetensors, synthetic_blob

    try:
        if spec.get("engine_factory"):  # "module:callable" -- host-logic tests run the loop on a fake engine
            import importlib

            mod, _, attr = spec["engine_factory"].partition(":")
            eng = getattr(importlib.import_module(mod), attr)(**spec["engine_kwargs"])
        else:
            shape = spec["shape"] if isinstance(spec["shape"], ModelShape) else MODEL_SHAPES[spec["shape"]]
            blob = blob_from_safetensors(shape, spec["weights"]) if spec.get("weights") else \
                synthetic_blob(shape, seed=spec.get("seed", 0))
            eng = Engine(shape, blob, device=spec["device"], **spec["engine_kwargs"])
            del blob
        worker = EngineWorker(eng, decode_burst=spec.get("decode_burst", 8), name=f"b200llm-gpu{spec['device']}")
        worker.start()
    except BaseException as exc:  # noqa: BLE001
        conn.send(("init_failed", repr(exc)))
        return
    send_lock = threading.Lock()

    def send(msg):
        with send_lock:
            conn.send(msg)

    send(("ready", spec["device"]))
    pending: dict[int, Future] = {}

    def on_done(rid: int, fut: Future):
        pending.pop(rid, None)
        exc = fut.exception()
        if exc is not None:
            send(("fail", rid, type(exc).__name__, str(exc)))
        else:
            r: GenerationResult = fut.result()
            send(("done", rid, np.asarray(r.tokens, dtype=np.int32).tobytes(), r.finish_reason, r.prompt_tokens,
                  r.queue_ms, r.total_ms, r.cached_prompt_tokens))

    while True:
        try:
            msg = conn.recv()
        except (EOFError, OSError):
            break
        kind = msg[0]
        if kind == "gen":
            _, rid, ids, sampling = msg
            fut = worker.submit(np.frombuffer(ids, dtype=np.int32).tolist(), SamplingParams(**sampling))
            pending[rid] = fut
            fut.add_done_callback(lambda f, rid=rid: on_done(rid, f))
        elif kind == "cancel":
            fut = pending.get(msg[1])
            if fut is not None:
                EngineWorker.cancel(fut)
        elif kind == "snapshot":
            send(("snapshot", msg[1], worker.snapshot()))
        elif kind == "stop":
            break
    worker.shutdown()
    worker.join(timeout=5)
    if hasattr(eng, "close"):
        eng.close()
    conn.close()


class RemoteEngineWorker:
    """Serving-process proxy of an EngineWorker that lives in its own process on its own GPU."""

    FAIL_LIMIT = EngineWorker.FAIL_LIMIT
    REPROBE_S = EngineWorker.REPROBE_S

    def __init__(self, device: int, shape, engine_kwargs: dict, *, weights: str | None = None, seed: int = 0,
                 decode_burst: int = 8, start_timeout_s: float = 300.0, engine_factory: str | None = None):
        ctx = mp.get_context("spawn")  # a forked child would inherit the parent's CUDA state
        self.device = device
        self.conn, child = ctx.Pipe()
        spec = {"device": device, "shape": shape, "engine_kwargs": dict(engine_kwargs), "weights": weights,
                "seed": seed, "decode_burst": decode_burst, "engine_factory": engine_factory}
        self.proc = ctx.Process(target=_worker_main, args=(child, spec), daemon=True, name=f"b200llm-gpu{device}")
        self.proc.start()
        child.close()
        self._start_timeout_s = start_timeout_s
        self._ready = False
        self._lock = threading.Lock()
        self._futures: dict[int, tuple[Future, int]] = {}
        self._snapshots: dict[int, Future] = {}
        self._ids = itertools.count(1)
        self._inflight_tokens = 0
        self.completed = 0
        self.consecutive_failures = 0
        self.last_failure_t = 0.0
        self._reader = threading.Thread(target=self._read_loop, daemon=True, name=f"b200llm-gpu{device}-reader")

    # ---- Router interface -------------------------------------------------------------------
    def is_alive(self) -> bool:
        return self._ready and self.proc.is_alive()

    def start(self) -> None:
        """Wait for the child's engine (weights upload, KV pool) and start reading its results."""
        if self._ready:
            return
        if not self.conn.poll(self._start_timeout_s):
            raise RuntimeError(f"GPU {self.device}: worker process did not come up")
        msg = self.conn.recv()
        if msg[0] != "ready":
            raise RuntimeError(f"GPU {self.device}: worker process failed to start: {msg[1:]}")
        self._ready = True
        self._reader.start()

    def submit(self, ids: Sequence[int], sampling: SamplingParams) -> Future:
        fut: Future = Future()
        rid = next(self._ids)
        cost = len(ids) + sampling.max_new_tokens
        with self._lock:
            self._futures[rid] = (fut, cost)
            self._inflight_tokens += cost
            try:
                self.conn.send(("gen", rid, np.asarray(ids, dtype=np.int32).tobytes(), asdict(sampling)))
            except (OSError, ValueError) as exc:
                self._futures.pop(rid, None)
                self._inflight_tokens -= cost
                self._note_failure()
                fut.set_exception(RuntimeError(f"GPU {self.device}: worker process is gone: {exc}"))
                return fut
        fut.cancel_request = lambda: self._send(("cancel", rid))
        return fut

    @property
    def inflight_tokens(self) -> int:
        with self._lock:
            return self._inflight_tokens

    @property
    def healthy(self) -> bool:
        return self.consecutive_failures < self.FAIL_LIMIT and self.proc.is_alive()

    def due_for_reprobe(self, now: float) -> bool:
        return self.proc.is_alive() and self.consecutive_failures >= self.FAIL_LIMIT and \
            now - self.last_failure_t >= self.REPROBE_S

    def snapshot(self, timeout: float = 10.0) -> dict:
        fut: Future = Future()
        rid = next(self._ids)
        self._snapshots[rid] = fut
        self._send(("snapshot", rid))
        return fut.result(timeout=timeout)

    def shutdown(self) -> None:
        try:
            self._send(("stop",))
        except Exception:  # noqa: BLE001
            pass

    def join(self, timeout: float | None = None) -> None:
        self.proc.join(timeout)
        if self.proc.is_alive():
            self.proc.terminate()

    # ---- internals --------------------------------------------------------------------------
    def _send(self, msg) -> None:
        with self._lock:
            try:
                self.conn.send(msg)
            except (OSError, ValueError):
                pass

    def _note_failure(self) -> None:
        self.consecutive_failures += 1
        self.last_failure_t = time.perf_counter()

    def _read_loop(self) -> None:
        while True:
            try:
                msg = self.conn.recv()
            except (EOFError, OSError):
                break
            kind = msg[0]
            if kind == "snapshot":
                fut = self._snapshots.pop(msg[1], None)
                if fut is not None:
                    fut.set_result(msg[2])
                continue
            with self._lock:
                fut, cost = self._futures.pop(msg[1], (None, 0))
                self._inflight_tokens -= cost
            if fut is None:
                continue
            if kind == "done":
                _, _, tok, reason, n_prompt, queue_ms, total_ms, cached = msg
                self.completed += 1
                self.consecutive_failures = 0
                fut.set_result(GenerationResult(tokens=np.frombuffer(tok, dtype=np.int32).tolist(), finish_reason=reason,
                                                prompt_tokens=n_prompt, queue_ms=queue_ms, total_ms=total_ms,
                                                cached_prompt_tokens=cached))
            else:
                _, _, exc_type, text = msg
                if exc_type not in ("ValueError", "TimeoutError"):  # a bad request is not a bad replica
                    self._note_failure()
                exc_cls = {"ValueError": ValueError, "TimeoutError": TimeoutError}.get(exc_type, RuntimeError)
                fut.set_exception(exc_cls(text))
        # the child is gone: nothing in flight will ever come back
        with self._lock:
            orphans = list(self._futures.values())
            self._futures.clear()
            self._inflight_tokens = 0
        for fut, _ in orphans:
            if not fut.done():
                fut.set_exception(RuntimeError(f"GPU {self.device}: worker process exited"))
        self.consecutive_failures = max(self.consecutive_failures, self.FAIL_LIMIT)
