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
from concurrent.futures import TimeoutError as FutureTimeout
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
    cached_prompt_tokens: int = 0   # leading prompt tokens whose KV came from the prefix cache


@dataclass
class _Request:
    ids: list[int]
    sampling: SamplingParams
    future: Future
    t_submit: float = field(default_factory=time.perf_counter)
    t_start: float = 0.0
    slot: int = -1
    n_done: int = 1  # tokens generated so far (prefill selects the first one), refreshed at every status read
    cancelled: bool = False    # set from any thread when the caller gave up (timeout)
    cancel_sent: bool = False  # worker thread: the device has been told
    cached_tokens: int = 0     # prompt tokens served from the prefix cache


class EngineWorker(threading.Thread):
    """One GPU: admits, decodes and retires sequences. `engine` needs the Engine interface:
    cfg.max_seqs/max_seq_len/max_batch_tokens/max_new_tokens, shape.eos_token_id, seq_alloc,
    seq_free, seq_cancel, prefill, decode, read_tokens_batch."""

    # a replica whose last FAIL_LIMIT engine calls raised is left out of the rotation and
    # re-probed with a single request every REPROBE_S seconds ("a dead GPU
    # replica should be removed from the router, not crash the server")
    FAIL_LIMIT = 3
    REPROBE_S = 5.0

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
        self.cancelled = 0
        self.prefix_hit_tokens = 0
        self.consecutive_failures = 0
        self.last_failure_t = 0.0
        self._stopping = False
        # wall time of the worker loop by phase, for the stream benchmark's report
        self.phase_s = {"idle": 0.0, "admit": 0.0, "decode": 0.0, "retire": 0.0}
        self.iterations = 0

    # ---- called from any thread ----
    def submit(self, ids: Sequence[int], sampling: SamplingParams) -> Future:
        fut: Future = Future()
        req = _Request(list(ids), sampling, fut)
        fut.cancel_request = lambda: setattr(req, "cancelled", True)  # Router.generate on timeout
        with self._lock:
            self._inflight_tokens += len(req.ids) + sampling.max_new_tokens
        self.inbox.put(req)
        return fut

    @staticmethod
    def cancel(fut: Future) -> None:
        """The caller no longer wants the result: a waiting request is dropped at the next
        admission, an active one is stopped on the device and its slot freed at the next retire."""
        stop = getattr(fut, "cancel_request", None)
        if stop is not None:
            stop()

    def snapshot(self) -> dict:
        """Counters for the stream benchmark's report (also served across the process boundary by
        remote_worker.RemoteEngineWorker)."""
        cache = self.engine.prefix_cache_stats() if hasattr(self.engine, "prefix_cache_stats") else {}
        return {"phase_s": {k: round(v, 3) for k, v in self.phase_s.items()}, "iterations": self.iterations,
                "completed": self.completed, "cancelled": self.cancelled, "generated_tokens": self.generated_tokens,
                "prefix_hit_tokens": self.prefix_hit_tokens, "prefix_cache": cache}

    @property
    def healthy(self) -> bool:
        return self.consecutive_failures < self.FAIL_LIMIT

    def due_for_reprobe(self, now: float) -> bool:
        return not self.healthy and now - self.last_failure_t >= self.REPROBE_S

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

    def _release(self, req: _Request) -> None:
        with self._lock:
            self._inflight_tokens -= len(req.ids) + req.sampling.max_new_tokens

    def _fail(self, req: _Request, exc: BaseException) -> None:
        self._release(req)
        if not req.future.done():
            req.future.set_exception(exc)

    def _pages_needed(self, tokens: int) -> int:
        return (tokens + 15) // 16

    def _admit(self) -> bool:
        """Prefill as many waiting requests as fit. Returns True when anything was admitted or
        dropped (progress), False when every waiting request has to wait for a retirement."""
        cfg = self.engine.cfg
        batch: list[_Request] = []
        tokens = 0
        keep: list[_Request] = []
        progress = False
        pool_tokens = int(getattr(cfg, "kv_pool_tokens", 0)) or cfg.max_seqs * cfg.max_seq_len
        for req in self.waiting:
            n = len(req.ids)
            max_new = max(1, min(req.sampling.max_new_tokens, cfg.max_new_tokens))
            if req.cancelled:
                self.cancelled += 1
                self._fail(req, TimeoutError("request cancelled before admission"))
                progress = True
                continue
            if n == 0 or n >= cfg.max_seq_len or n > cfg.max_batch_tokens:
                self._fail(req, ValueError(f"prompt of {n} tokens does not fit the engine "
                                           f"(max_seq_len {cfg.max_seq_len}, batch budget {cfg.max_batch_tokens})"))
                progress = True
                continue
            cap = min(n + max_new, cfg.max_seq_len)
            if self._pages_needed(cap) * 16 > pool_tokens:
                self._fail(req, ValueError(f"request needs {cap} KV tokens, the whole pool holds {pool_tokens}"))
                progress = True
                continue
            if len(self.active) + len(batch) >= cfg.max_seqs or tokens + n > cfg.max_batch_tokens:
                keep.append(req)
                continue
            try:
                req.slot = self.engine.seq_alloc(cap, req.ids)
            except Exception as exc:  # noqa: BLE001  KV pool exhausted: retry after something retires
                if not self.active and not batch:
                    # nothing is running that could ever free pages: waiting would spin forever
                    self._fail(req, exc)
                    progress = True
                else:
                    keep.append(req)
                continue
            req.cached_tokens = int(getattr(self.engine, "last_cached_tokens", 0))
            req.t_start = time.perf_counter()
            batch.append(req)
            tokens += n - req.cached_tokens
        self.waiting = keep
        if not batch:
            return progress
        try:
            self.engine.prefill([r.ids for r in batch], [r.slot for r in batch], [r.sampling for r in batch])
        except Exception as exc:  # noqa: BLE001
            self._note_failure()
            for r in batch:
                try:
                    self.engine.seq_free(r.slot)
                except Exception:  # noqa: BLE001
                    pass
                self._fail(r, exc)
            return True
        self.prefix_hit_tokens += sum(r.cached_tokens for r in batch)
        self.active.extend(batch)
        return True

    def _note_failure(self) -> None:
        self.consecutive_failures += 1
        self.last_failure_t = time.perf_counter()

    def _retire(self) -> None:
        if not self.active:
            return
        for req in self.active:
            if req.cancelled and not req.cancel_sent:
                # enqueued ahead of the status gather below, so this very read already sees it
                self.engine.seq_cancel(req.slot)
                req.cancel_sent = True
        # ONE stream synchronisation for the status of every active slot and the tokens of the
        # finished ones (b200llm_read_tokens_batch)
        counts, finished, rows = self.engine.read_tokens_batch([r.slot for r in self.active])
        still: list[_Request] = []
        eos = self.engine.shape.eos_token_id
        for req, n, fin, tokens in zip(self.active, counts, finished, rows):
            if not fin:
                req.n_done = int(n)
                still.append(req)
                continue
            self.engine.seq_free(req.slot)
            req.slot = -1
            if req.cancelled:
                self.cancelled += 1
                self._fail(req, TimeoutError("request cancelled"))
                continue
            tokens = list(tokens[: int(n)])
            hit_eos = eos in tokens
            if hit_eos:
                tokens = tokens[: tokens.index(eos)]
            now = time.perf_counter()
            self._release(req)
            self.completed += 1
            self.generated_tokens += len(tokens)
            req.future.set_result(GenerationResult(
                tokens=tokens, finish_reason=1 if hit_eos else 2, prompt_tokens=len(req.ids),
                queue_ms=(req.t_start - req.t_submit) * 1e3, total_ms=(now - req.t_submit) * 1e3,
                cached_prompt_tokens=req.cached_tokens))
        self.active = still
        self.consecutive_failures = 0  # a full admit/decode/retire round trip worked

    def run(self) -> None:
        clock = time.perf_counter
        stalled = False  # waiting requests exist but none can be admitted and nothing is running
        while not self._stopping:
            t0 = clock()
            self._drain_inbox(block=(not self.active and not self.waiting) or stalled)
            t1 = clock()
            self.phase_s["idle"] += t1 - t0
            if self._stopping:
                break
            try:
                # admit first: a new request joins the burst of the same iteration (enqueueing the
                # burst ahead of the admission was measured: the stream is GPU-bound, and each
                # request then waits one more burst -- 476 vs 515 requests/s)
                stalled = False
                if self.waiting:
                    progress = self._admit()
                    self._retire()  # max_new_tokens == 1 finishes at prefill
                    stalled = not progress and not self.active
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
                self._note_failure()
                # give the slots (and their KV pages) back before failing the requests: one
                # non-fatal error must not leak capacity for the rest of the process
                for req in self.active:
                    if req.slot >= 0:
                        try:
                            self.engine.seq_free(req.slot)
                        except Exception:  # noqa: BLE001
                            pass
                        req.slot = -1
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
        for w in self.workers:  # thread workers start running; process workers finish coming up
            if not w.is_alive():
                w.start()

    def _pick(self) -> EngineWorker:
        now = time.perf_counter()
        for w in self.workers:  # one probe request at a time for a replica that was taken out
            if w.due_for_reprobe(now) and w.inflight_tokens == 0:
                w.last_failure_t = now
                return w
        healthy = [w for w in self.workers if w.healthy]
        # with every replica failing there is nothing better than the least loaded one: its
        # exception reaches the caller, whose catch-all answers with the canned reply
        return min(healthy or self.workers, key=lambda w: w.inflight_tokens)

    def submit(self, ids: Sequence[int], sampling: SamplingParams) -> Future:
        return self._pick().submit(ids, sampling)

    def generate(self, ids: Sequence[int], sampling: SamplingParams, timeout: float | None = None) -> GenerationResult:
        fut = self.submit(ids, sampling)
        try:
            return fut.result(timeout=timeout)
        except FutureTimeout:
            EngineWorker.cancel(fut)  # do not keep decoding for a caller that has already answered
            raise

    def shutdown(self) -> None:
        for w in self.workers:
            w.shutdown()
        for w in self.workers:
            w.join(timeout=5)
