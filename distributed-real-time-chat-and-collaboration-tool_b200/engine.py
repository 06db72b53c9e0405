"""Python face of the CUDA generation engine (C ABI: include/b200llm.h).

``Engine`` owns one GPU: weights, the paged KV pool and per-sequence generation state all live
in HBM inside libb200llm.so. Python only moves token ids in and out. The method names mirror
the C entry points (seq_alloc / prefill / decode / generate) which in turn mirror the phases of
transformers' GenerationMixin.generate that the engine reproduces.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Sequence

import numpy as np

from ._lib import Config, Sampling, check, lib


@dataclass(frozen=True)
class ModelShape:
    """GPT2Config fields the engine needs (names follow transformers' GPT2Config)."""
    n_layer: int = 12
    n_head: int = 12
    d_model: int = 768
    vocab_size: int = 50257
    n_positions: int = 1024
    ln_eps: float = 1e-5
    eos_token_id: int = 50256


GPT2_SMALL = ModelShape()
GPT2_MEDIUM = ModelShape(n_layer=24, n_head=16, d_model=1024)
GPT2_LARGE = ModelShape(n_layer=36, n_head=20, d_model=1280)
MODEL_SHAPES = {"small": GPT2_SMALL, "medium": GPT2_MEDIUM, "large": GPT2_LARGE}


@dataclass
class SamplingParams:
    """Per-request generation controls (the reference passes temperature / max_output_tokens
    through generation_config, llm_server.py:169-172)."""
    max_new_tokens: int = 32
    min_new_tokens: int = 0
    temperature: float = 0.0  # <= 0: greedy
    top_k: int = 0
    seed: int = 0


def _i32(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a, dtype=np.int32))


class Engine:
    def __init__(self, shape: ModelShape, weights_blob: np.ndarray, *, device: int = 0,
                 max_seqs: int = 64, max_seq_len: int = 256, max_batch_tokens: int = 8192,
                 max_new_tokens: int = 256, kv_pool_tokens: int = 0, use_cuda_graphs: bool = True,
                 use_pdl: bool = True, decode_split_k: int = 0, keep_logits: bool = False,
                 decode_reduce: int = 0, pad_token_id: int | None = None, prefix_cache: bool = False):
        blob = np.ascontiguousarray(weights_blob, dtype=np.float32).reshape(-1)
        self.shape = shape
        self.cfg = Config(
            n_layer=shape.n_layer, n_head=shape.n_head, d_model=shape.d_model,
            vocab_size=shape.vocab_size, n_positions=shape.n_positions, ln_eps=shape.ln_eps,
            eos_token_id=shape.eos_token_id,
            pad_token_id=shape.eos_token_id if pad_token_id is None else pad_token_id,
            max_seqs=max_seqs, max_seq_len=max_seq_len, max_batch_tokens=max_batch_tokens,
            max_new_tokens=max_new_tokens, kv_pool_tokens=kv_pool_tokens, device=device,
            use_cuda_graphs=int(use_cuda_graphs), use_pdl=int(use_pdl), decode_split_k=decode_split_k,
            keep_logits=int(keep_logits), decode_reduce=decode_reduce, prefix_cache=int(prefix_cache))
        expect = lib.b200llm_weight_blob_floats(C.byref(self.cfg))
        if blob.size != expect:
            raise ValueError(f"weight blob has {blob.size} floats, config needs {expect}")
        self._h = C.c_void_p()
        check(lib.b200llm_engine_create(C.byref(self.cfg), blob.ctypes.data, blob.size, C.byref(self._h)))
        self.last_cached_tokens = 0

    # -- lifecycle ------------------------------------------------------------------------
    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h.value:
            lib.b200llm_engine_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def set_stream(self, cuda_stream: int | None) -> None:
        check(lib.b200llm_engine_set_stream(self._h, cuda_stream))

    def sync(self) -> None:
        check(lib.b200llm_sync(self._h))

    @property
    def kernel_launches(self) -> int:
        return int(lib.b200llm_kernel_launches(self._h))

    # -- KV-cache manager -------------------------------------------------------------------
    def seq_alloc(self, capacity_tokens: int, prompt: Sequence[int] | None = None) -> int:
        """A KV slot of `capacity_tokens`. With `prompt` (and an engine created with
        prefix_cache=True) the longest cached run of the prompt's leading pages is shared with the
        slot and the next prefill of it skips those tokens; `last_cached_tokens` says how many."""
        slot = C.c_int32(-1)
        if prompt is None:
            check(lib.b200llm_seq_alloc(self._h, capacity_tokens, C.byref(slot)))
            self.last_cached_tokens = 0
        else:
            ids = _i32(prompt)
            cached = C.c_int32(0)
            check(lib.b200llm_seq_alloc_with_prefix(self._h, ids.ctypes.data, len(ids), capacity_tokens,
                                                    C.byref(slot), C.byref(cached)))
            self.last_cached_tokens = cached.value
        return slot.value

    def seq_cancel(self, slot: int) -> None:
        """Mark the slot finished on the device (its caller gave up); free it after the next
        status read."""
        check(lib.b200llm_seq_cancel(self._h, slot))

    def prefix_cache_stats(self) -> dict:
        a, b, c = C.c_int64(), C.c_int64(), C.c_int64()
        check(lib.b200llm_prefix_cache_stats(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return {"lookup_tokens": a.value, "hit_tokens": b.value, "cached_pages": c.value}

    def seq_free(self, slot: int) -> None:
        check(lib.b200llm_seq_free(self._h, slot))

    def kv_pages(self) -> tuple[int, int]:
        free, total = C.c_int64(), C.c_int64()
        check(lib.b200llm_kv_pages_free(self._h, C.byref(free), C.byref(total)))
        return free.value, total.value

    # -- model executor ---------------------------------------------------------------------
    @staticmethod
    def _sampling_array(params: Sequence[SamplingParams] | None, n: int):
        if params is None:
            return None
        arr = (Sampling * n)()
        for i, p in enumerate(params):
            arr[i] = Sampling(p.max_new_tokens, p.min_new_tokens, p.temperature, p.top_k, p.seed)
        return arr

    def prefill(self, prompts: Sequence[Sequence[int]], slots: Sequence[int],
                sampling: Sequence[SamplingParams] | None = None, *, ids_device_ptr: int | None = None,
                logits_all_ptr: int | None = None) -> None:
        """Forward the packed prompts, fill their KV pages, select each first new token.
        ``ids_device_ptr``: the packed int32 ids already resident in device memory."""
        lens = [len(p) for p in prompts]
        cu = _i32(np.concatenate([[0], np.cumsum(lens)]))
        sl = _i32(slots)
        samp = self._sampling_array(sampling, len(prompts))
        if ids_device_ptr is None:
            ids = _i32(np.concatenate([np.asarray(p, dtype=np.int32) for p in prompts]))
            ids_ptr, on_dev = ids.ctypes.data, 0
        else:
            ids, ids_ptr, on_dev = None, ids_device_ptr, 1
        check(lib.b200llm_prefill(self._h, ids_ptr, on_dev, cu.ctypes.data, sl.ctypes.data,
                                  C.cast(samp, C.c_void_p) if samp is not None else None,
                                  len(prompts), logits_all_ptr))
        # (host ids are copied into the engine's pinned staging buffer before the call returns)

    def decode(self, slots: Sequence[int], n_steps: int = 1) -> None:
        sl = _i32(slots)
        check(lib.b200llm_decode(self._h, sl.ctypes.data, len(sl), n_steps))

    def read_tokens(self, slot: int, max_tokens: int | None = None) -> tuple[list[int], bool]:
        n = max_tokens or self.cfg.max_new_tokens
        buf = np.empty(n, dtype=np.int32)
        cnt, fin = C.c_int32(), C.c_int32()
        check(lib.b200llm_read_tokens(self._h, slot, buf.ctypes.data, n, C.byref(cnt), C.byref(fin)))
        return buf[: cnt.value].tolist(), bool(fin.value)

    def read_status(self, slots: Sequence[int]) -> tuple[np.ndarray, np.ndarray]:
        sl = _i32(slots)
        counts = np.empty(len(sl), dtype=np.int32)
        fin = np.empty(len(sl), dtype=np.int32)
        check(lib.b200llm_read_status(self._h, sl.ctypes.data, len(sl), counts.ctypes.data, fin.ctypes.data))
        return counts, fin

    def read_tokens_batch(self, slots: Sequence[int], only_finished: bool = True):
        """(counts, finished, tokens) of several slots with one stream synchronisation; tokens[i]
        is None for a slot that was not copied (unfinished with only_finished)."""
        sl = _i32(slots)
        n = len(sl)
        stride = self.cfg.max_new_tokens
        buf = np.empty((n, stride), dtype=np.int32)
        counts = np.empty(n, dtype=np.int32)
        fin = np.empty(n, dtype=np.int32)
        check(lib.b200llm_read_tokens_batch(self._h, sl.ctypes.data, n, int(only_finished), buf.ctypes.data,
                                            stride, counts.ctypes.data, fin.ctypes.data))
        toks = [buf[i, : counts[i]].tolist() if (fin[i] or not only_finished) else None for i in range(n)]
        return counts, fin, toks

    def read_last_logits(self, batch: int) -> np.ndarray:
        out = np.empty((batch, self.shape.vocab_size), dtype=np.float32)
        check(lib.b200llm_read_last_logits(self._h, batch, out.ctypes.data))
        return out

    def generate(self, prompts: Sequence[Sequence[int]],
                 sampling: Sequence[SamplingParams] | SamplingParams | None = None) -> list[list[int]]:
        """End-to-end generation with host buffers (b200llm_generate): ids in, tokens out."""
        n = len(prompts)
        if isinstance(sampling, SamplingParams):
            sampling = [sampling] * n
        lens = [len(p) for p in prompts]
        cu = _i32(np.concatenate([[0], np.cumsum(lens)]))
        ids = _i32(np.concatenate([np.asarray(p, dtype=np.int32) for p in prompts]))
        samp = self._sampling_array(sampling, n)
        stride = max((p.max_new_tokens for p in sampling), default=self.cfg.max_new_tokens) \
            if sampling else self.cfg.max_new_tokens
        stride = min(stride, self.cfg.max_new_tokens)
        out = np.empty((n, stride), dtype=np.int32)
        counts = np.empty(n, dtype=np.int32)
        check(lib.b200llm_generate(self._h, ids.ctypes.data, cu.ctypes.data,
                                   C.cast(samp, C.c_void_p) if samp is not None else None, n,
                                   out.ctypes.data, stride, counts.ctypes.data))
        return [out[i, : counts[i]].tolist() for i in range(n)]
