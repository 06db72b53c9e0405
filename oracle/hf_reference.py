"""Hugging Face transformers driver: the library the reference pins (requirements.txt:6-7,
transformers==4.35.0; installed here: 5.5.0) run on CPU.

TEST INFRASTRUCTURE -- used (a) to pin oracle/gpt2_oracle.py and generate tests/golden/,
(b) as the timed CPU baseline of bench.py (`cpu_baseline`, `--impl reference`). The reference
repository itself has no local generation code to run, so this is the
closest thing to "the reference's own CPU implementation" that exists.
"""
from __future__ import annotations

import os
import time

import torch

from .gpt2_oracle import GPT2Shape


def build_hf_model(shape: GPT2Shape, weights: dict[str, torch.Tensor] | None = None, attn: str = "eager"):
    os.environ.setdefault("HF_HUB_OFFLINE", "1")
    from transformers import GPT2Config, GPT2LMHeadModel

    cfg = GPT2Config(vocab_size=shape.vocab_size, n_positions=shape.n_positions, n_embd=shape.d_model,
                     n_layer=shape.n_layer, n_head=shape.n_head, layer_norm_epsilon=shape.ln_eps,
                     bos_token_id=shape.eos_token_id, eos_token_id=shape.eos_token_id,
                     attn_implementation=attn)
    model = GPT2LMHeadModel(cfg).eval()
    if weights is not None:
        sd = dict(weights)
        sd["lm_head.weight"] = sd["transformer.wte.weight"]  # tied (modeling_gpt2.py:646-651)
        missing, unexpected = model.load_state_dict(sd, strict=False)
        bad = [k for k in missing if "attn.bias" not in k and "masked_bias" not in k]
        assert not bad and not unexpected, (bad, unexpected)
    return model


@torch.no_grad()
def hf_logits(model, ids: torch.Tensor) -> torch.Tensor:
    return model(input_ids=ids).logits.float()


@torch.no_grad()
def hf_generate(model, ids: torch.Tensor, max_new_tokens: int, min_new_tokens: int = 0) -> torch.Tensor:
    out = model.generate(input_ids=ids, attention_mask=torch.ones_like(ids), do_sample=False,
                         max_new_tokens=max_new_tokens, min_new_tokens=min_new_tokens,
                         pad_token_id=model.config.eos_token_id)
    return out[:, ids.shape[1]:]


def time_hf_generate(model, ids: torch.Tensor, max_new_tokens: int, threads: int | None = None,
                     repeats: int = 1) -> dict:
    """Wall-clock of .generate on the host cores; returns generated tokens/s (best of repeats)."""
    threads = threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    best = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        hf_generate(model, ids, max_new_tokens, min_new_tokens=max_new_tokens)
        best = min(best, time.perf_counter() - t0)
    n = ids.shape[0] * max_new_tokens
    return {"seconds": best, "tokens": n, "tokens_per_s": n / best, "threads": threads}
