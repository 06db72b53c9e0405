"""Batching front end: turns concurrent generate_content calls into engine batches.

The reference serves requests from a 4-thread pool with no batching at all
(llm_server.py:501) because its generator is a remote service. Here each GPU has one worker
thread that owns an Engine (CUDA is never touched from gRPC handler threads) and runs
continuous batching: between short bursts of decode steps it admits waiting requests
(prefill), retires finished sequences and frees their KV pages. A Router shards independent
requests across the per-GPU workers by in-flight token count -- replicas only, no collective
(SURVEY.md 8e).
"""
from __future__ import annotations

import os
import queue
import threading
import time
from concurrent.futures import Future
from dataclasses import dataclass, field
from typing import Callable, Sequence

from .engine import SamplingParams


@dataclass
class GenerationResult:
    tokens: list[int]
    finish_reason: int          # 1 = EOS (STOP), 2 = MAX_TOKENS  (llm_server.py:184)
    prompt_tokens: int
    queue_ms: float = 0.0
    total_ms: float = 0.0


@dataclass
class _Request:
    ids: list[int]
    sampling: SamplingParams
    future: Future
    t_submit: float = field(default_factory=time.perf_counter)
    t_start: float = 0.0
    slot: int = -1
    n_done: int = 1  # tokens generated so far (prefill selects the first one), refreshed at every status read


class EngineWorker(threading.Thread):
    """One GPU: admits, decodes and retires sequences. `engine` needs the Engine interface:
    cfg.max_seqs/max_seq_len/max_batch_tokens/max_new_tokens, shape.eos_token_id, seq_alloc,
    seq_free, prefill, decode, read_status, read_tokens."""

    def __init__(self, engine, *, decode_burst: int = 8, name: str = "b200llm-worker"):
        super().__init__(name=name, daemon=True)
        self.engine = engine
        # B200LLM_DECODE_BURST overrides the steps per burst (throughput / admission-latency trade-off)
        self.decode_burst = int(os.environ.get("B200LLM_DECODE_BURST", decode_burst))
        self.inbox: "queue.Queue[_Request | None]" = queue.Queue()
        self.active: list[_Request] = []
        self.waiting: list[_Request] = []
        self._inflight_tokens = 0
        self._lock = threading.Lock()
        self.completed = 0
        self.generated_tokens = 0
        self._stopping = False
        # wall time of the worker loop by phase, for the stream benchmark's report
        self.phase_s = {"idle": 0.0, "admit": 0.0, "decode": 0.0, "retire": 0.0}
        self.iterations = 0

    # ---- called from any thread ----
    def submit(self, ids: Sequence[int], sampling: SamplingParams) -> Future:
        fut: Future = Future()
        req = _Request(list(ids), sampling, fut)
        with self._lock:
            self._inflight_tokens += len(req.ids) + sampling.max_new_tokens
        self.inbox.put(req)
        return fut

    @property
    def inflight_tokens(self) -> int:
        with self._lock:
            return self._inflight_tokens

    def shutdown(self) -> None:
        self._stopping = True
        self.inbox.put(None)

    # ---- worker thread ----
    def _drain_inbox(self, block: bool) -> None:
        try:
            item = self.inbox.get(timeout=0.05) if block else self.inbox.get_nowait()
        except queue.Empty:
            return
        while True:
            if item is not None:
                self.waiting.append(item)
            try:
                item = self.inbox.get_nowait()
            except queue.Empty:
                return

    def _fail(self, req: _Request, exc: BaseException) -> None:
        with self._lock:
            self._inflight_tokens -= len(req.ids) + req.sampling.max_new_tokens
        if not req.future.done():
            req.future.set_exception(exc)

    def _admit(self) -> None:
        cfg = self.engine.cfg
        batch: list[_Request] = []
        tokens = 0
        keep: list[_Request] = []
        for req in self.waiting:
            n = len(req.ids)
            max_new = max(1, min(req.sampling.max_new_tokens, cfg.max_new_tokens))
            if n == 0 or n >= cfg.max_seq_len or n > cfg.max_batch_tokens:
                self._fail(req, ValueError(f"prompt of {n} tokens does not fit the engine "
                                           f"(max_seq_len {cfg.max_seq_len}, batch budget {cfg.max_batch_tokens})"))
                continue
            if len(self.active) + len(batch) >= cfg.max_seqs or tokens + n > cfg.max_batch_tokens:
                keep.append(req)
                continue
            try:
                req.slot = self.engine.seq_alloc(min(n + max_new, cfg.max_seq_len))
            except Exception:  # KV pool exhausted: retry after something retires
                keep.append(req)
                continue
            req.t_start = time.perf_counter()
            batch.append(req)
            tokens += n
        self.waiting = keep
        if not batch:
            return
        try:
            self.engine.prefill([r.ids for r in batch], [r.slot for r in batch], [r.sampling for r in batch])
        except Exception as exc:  # noqa: BLE001
            for r in batch:
                try:
                    self.engine.seq_free(r.slot)
                except Exception:  # noqa: BLE001
                    pass
                self._fail(r, exc)
            return
        self.active.extend(batch)

    def _retire(self) -> None:
        if not self.active:
            return
        counts, finished = self.engine.read_status([r.slot for r in self.active])
        still: list[_Request] = []
        eos = self.engine.shape.eos_token_id
        for req, n, fin in zip(self.active, counts, finished):
            if not fin:
                req.n_done = int(n)
                still.append(req)
                continue
            tokens, _ = self.engine.read_tokens(req.slot, int(n))
            self.engine.seq_free(req.slot)
            hit_eos = eos in tokens
            if hit_eos:
                tokens = tokens[: tokens.index(eos)]
            now = time.perf_counter()
            with self._lock:
                self._inflight_tokens -= len(req.ids) + req.sampling.max_new_tokens
            self.completed += 1
            self.generated_tokens += len(tokens)
            req.future.set_result(GenerationResult(
                tokens=tokens, finish_reason=1 if hit_eos else 2, prompt_tokens=len(req.ids),
                queue_ms=(req.t_start - req.t_submit) * 1e3, total_ms=(now - req.t_submit) * 1e3))
        self.active = still

    def run(self) -> None:
        clock = time.perf_counter
        while not self._stopping:
            t0 = clock()
            self._drain_inbox(block=not self.active and not self.waiting)
            t1 = clock()
            self.phase_s["idle"] += t1 - t0
            if self._stopping:
                break
            try:
                # admit first: a new request joins the burst of the same iteration (enqueueing the
                # burst ahead of the admission was measured: the stream is GPU-bound, and each
                # request then waits one more burst -- 476 vs 515 requests/s)
                if self.waiting:
                    self._admit()
                    self._retire()  # max_new_tokens == 1 finishes at prefill
                t2 = clock()
                self.phase_s["admit"] += t2 - t1
                if self.active:
                    # no more steps than the longest remaining generation needs (a lone 32-token
                    # request runs 8 + 8 + 8 + 7 steps, not 4 x 8)
                    cap = self.engine.cfg.max_new_tokens
                    left = max(min(r.sampling.max_new_tokens, cap) - r.n_done for r in self.active)
                    self.engine.decode([r.slot for r in self.active], max(1, min(self.decode_burst, left)))
                    t3 = clock()
                    self._retire()
                    t4 = clock()
                    self.phase_s["decode"] += t3 - t2
                    self.phase_s["retire"] += t4 - t3
                    self.iterations += 1
            except Exception as exc:  # noqa: BLE001  a broken engine fails its requests, not the server
                for req in self.active + self.waiting:
                    self._fail(req, exc)
                self.active, self.waiting = [], []
        for req in self.active + self.waiting:
            self._fail(req, RuntimeError("worker shut down"))


class Router:
    """Shards independent requests over the per-GPU workers (fewest in-flight tokens first)."""

    def __init__(self, workers: Sequence[EngineWorker]):
        if not workers:
            raise ValueError("Router needs at least one worker")
        self.workers = list(workers)
        for w in self.workers:
            if not w.is_alive():
                w.start()

    def submit(self, ids: Sequence[int], sampling: SamplingParams) -> Future:
        worker = min(self.workers, key=lambda w: w.inflight_tokens)
        return worker.submit(ids, sampling)

    def generate(self, ids: Sequence[int], sampling: SamplingParams, timeout: float | None = None) -> GenerationResult:
        return self.submit(ids, sampling).result(timeout=timeout)

    def shutdown(self) -> None:
        for w in self.workers:
            w.shutdown()
        for w in self.workers:
            w.join(timeout=5)
