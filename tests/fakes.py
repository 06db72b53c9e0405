"""Test doubles (tests only; the product has no CPU path)."""
from __future__ import annotations

import threading
import types


class FakeEngine:
    """Implements the Engine interface the scheduler uses, with a toy deterministic 'model':
    token t+1 = (sum(prompt) + step) % 50000, EOS when that value is divisible by `eos_every`."""

    def __init__(self, max_seqs=4, max_seq_len=64, max_batch_tokens=128, max_new_tokens=16, eos_every=0,
                 fail_prefill=False, fail_decode_times=0, kv_pool_tokens=0, step_sleep=0.0):
        self.cfg = types.SimpleNamespace(max_seqs=max_seqs, max_seq_len=max_seq_len,
                                         max_batch_tokens=max_batch_tokens, max_new_tokens=max_new_tokens,
                                         kv_pool_tokens=kv_pool_tokens)
        self.fail_decode_times = fail_decode_times   # the next n decode calls raise
        self.step_sleep = step_sleep                 # seconds per decode step (slow replica)
        self.pool_left = kv_pool_tokens or max_seqs * max_seq_len
        self.cancels: list[int] = []
        self.last_cached_tokens = 0
        self.shape = types.SimpleNamespace(eos_token_id=50256, vocab_size=50257)
        self.eos_every = eos_every
        self.fail_prefill = fail_prefill
        self.slots: dict[int, dict] = {}
        self.free = list(range(max_seqs))
        self.calls: list[tuple] = []
        self.max_active_seen = 0
        self.lock = threading.Lock()

    def seq_alloc(self, capacity, prompt=None):
        if not self.free:
            raise RuntimeError("no free slot")
        need = (capacity + 15) // 16 * 16
        if need > self.pool_left:
            raise RuntimeError("KV pool exhausted")
        self.pool_left -= need
        s = self.free.pop(0)
        self.slots[s] = {"cap": capacity, "tokens": [], "fin": False, "base": 0, "max_new": 0, "pool": need}
        return s

    def seq_free(self, slot):
        self.pool_left += self.slots[slot]["pool"]
        del self.slots[slot]
        self.free.append(slot)

    def seq_cancel(self, slot):
        self.cancels.append(slot)
        self.slots[slot]["fin"] = True

    def _emit(self, st):
        if st["fin"]:
            return
        step = len(st["tokens"])
        tok = (st["base"] + step) % 50000
        if self.eos_every and step > 0 and (st["base"] + step) % self.eos_every == 0:
            tok = self.shape.eos_token_id
        st["tokens"].append(tok)
        if tok == self.shape.eos_token_id or len(st["tokens"]) >= st["max_new"]:
            st["fin"] = True

    def prefill(self, prompts, slots, sampling=None, **kw):
        if self.fail_prefill:
            raise RuntimeError("prefill exploded")
        self.calls.append(("prefill", [len(p) for p in prompts]))
        assert sum(len(p) for p in prompts) <= self.cfg.max_batch_tokens
        for p, s, sp in zip(prompts, slots, sampling):
            st = self.slots[s]
            assert len(p) <= st["cap"]
            st["base"] = sum(p)
            st["max_new"] = min(sp.max_new_tokens, self.cfg.max_new_tokens)
            self._emit(st)

    def decode(self, slots, n_steps):
        if self.fail_decode_times > 0:
            self.fail_decode_times -= 1
            raise RuntimeError("decode exploded")
        if self.step_sleep:
            import time
            time.sleep(self.step_sleep * n_steps)
        self.calls.append(("decode", len(slots), n_steps))
        self.max_active_seen = max(self.max_active_seen, len(slots))
        for _ in range(n_steps):
            for s in slots:
                self._emit(self.slots[s])

    def read_status(self, slots):
        return [len(self.slots[s]["tokens"]) for s in slots], [int(self.slots[s]["fin"]) for s in slots]

    def read_tokens_batch(self, slots, only_finished=True):
        counts, fin = self.read_status(slots)
        rows = [list(self.slots[s]["tokens"]) if (f or not only_finished) else None for s, f in zip(slots, fin)]
        return counts, fin, rows

    def read_tokens(self, slot, max_tokens=None):
        st = self.slots[slot]
        return list(st["tokens"][:max_tokens]), st["fin"]
