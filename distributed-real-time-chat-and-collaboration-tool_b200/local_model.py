"""The drop-in for ``LLMServicer.model``: a local generator with Gemini's call shape.

The reference creates ``genai.GenerativeModel('gemini-2.5-flash')`` (llm_server.py:32-33) and
calls ``self.model.generate_content(prompt, generation_config=...)`` at four sites
(:167, :231, :287, :403), then reads ``.candidates[0].finish_reason``, ``.text`` and
``.candidates[0].content.parts[0].text`` (:176-194). ``LocalGenerativeModel`` offers exactly
that surface, backed by the B200 engine through the batching Router. It is thread safe (the
reference calls it from a 4-thread pool without a lock, :501) and fails fast with an exception
rather than hanging, which the callers turn into their canned replies.
"""
from __future__ import annotations

import itertools
import threading
from dataclasses import dataclass, field

from .engine import SamplingParams
from .scheduler import GenerationResult, Router


@dataclass
class Part:
    text: str


@dataclass
class Content:
    parts: list[Part]


@dataclass
class Candidate:
    finish_reason: int  # 1 = STOP (EOS), 2 = MAX_TOKENS
    content: Content


@dataclass
class GenerateContentResponse:
    """Gemini-shaped result. ``usage`` is extra (tokens and timings for the service metrics)."""
    candidates: list[Candidate]
    usage: dict = field(default_factory=dict)

    @property
    def text(self) -> str:
        return self.candidates[0].content.parts[0].text

    def __bool__(self) -> bool:
        return True


class LocalGenerativeModel:
    def __init__(self, router: Router, tokenizer, *, default_max_new_tokens: int = 64,
                 default_temperature: float = 0.0, default_top_k: int = 40, timeout_s: float = 9.0,
                 max_prompt_tokens: int | None = None):
        self.router = router
        self.tokenizer = tokenizer
        self.default_max_new_tokens = default_max_new_tokens
        self.default_temperature = default_temperature
        self.default_top_k = default_top_k
        self.timeout_s = timeout_s  # below the Raft node's tightest LLM deadline (10 s, raft_node.py:2084)
        self.max_prompt_tokens = max_prompt_tokens
        self._seed = itertools.count(1)  # next() is atomic under the GIL: 128 handler threads share it
        self._meter = threading.Lock()
        self.requests = 0
        self.prompt_tokens = 0
        self.generated_tokens = 0

    def _sampling(self, generation_config: dict | None) -> SamplingParams:
        cfg = generation_config or {}
        temperature = float(cfg.get("temperature", self.default_temperature))
        max_new = int(cfg.get("max_output_tokens", self.default_max_new_tokens))
        top_k = int(cfg.get("top_k", self.default_top_k)) if temperature > 0 else 0
        return SamplingParams(max_new_tokens=max(1, max_new), min_new_tokens=0, temperature=temperature,
                              top_k=top_k, seed=int(cfg.get("seed", next(self._seed))))

    def generate_content(self, prompt: str, generation_config: dict | None = None) -> GenerateContentResponse:
        ids = self.tokenizer.encode(prompt)
        if not ids:
            raise ValueError("empty prompt")
        if self.max_prompt_tokens is not None and len(ids) > self.max_prompt_tokens:
            ids = ids[-self.max_prompt_tokens:]  # keep the instructions at the end of the template
        result: GenerationResult = self.router.generate(ids, self._sampling(generation_config), timeout=self.timeout_s)
        text = self.tokenizer.decode(result.tokens)
        with self._meter:
            self.requests += 1
            self.prompt_tokens += result.prompt_tokens
            self.generated_tokens += len(result.tokens)
        return GenerateContentResponse(
            candidates=[Candidate(finish_reason=result.finish_reason, content=Content(parts=[Part(text=text)]))],
            usage={"prompt_tokens": result.prompt_tokens, "generated_tokens": len(result.tokens),
                   "queue_ms": result.queue_ms, "total_ms": result.total_ms,
                   "cached_prompt_tokens": result.cached_prompt_tokens})
