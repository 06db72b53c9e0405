"""Generate tests/golden/gpt2_golden.npz from Hugging Face transformers (CPU).

TEST INFRASTRUCTURE. Weights come from oracle.gpt2_oracle.synthetic_weights (deterministic,
seeded) loaded into transformers' GPT2LMHeadModel; prompts from synthetic_prompts. The file pins
  - a 2-layer "tiny" model: full fp32 logits of every prompt position,
  - GPT-2 small (124M): sampled + top-64 last-position logits, greedy tokens for 8 prompts x 32
    new tokens (min_new_tokens = 32), and the top-1/top-2 margins used by the margin-aware gate.

    python -m oracle.make_golden
"""
from __future__ import annotations

from pathlib import Path

import numpy as np
import torch

from . import gpt2_oracle as O
from . import hf_reference as H

OUT = Path(__file__).resolve().parent.parent / "tests" / "golden" / "gpt2_golden.npz"
TINY = O.GPT2Shape(n_layer=2, n_head=2, d_model=128, vocab_size=1000, n_positions=96, eos_token_id=999)


def main():
    torch.set_num_threads(8)
    out = {}
    # ---- tiny model, eager attention (the strictest comparison) ----
    w = O.synthetic_weights(TINY, seed=3)
    ids = O.synthetic_prompts(3, 24, vocab=TINY.vocab_size, seed=11)
    hf = H.build_hf_model(TINY, w, attn="eager")
    out["tiny_ids"] = ids.numpy()
    out["tiny_logits"] = H.hf_logits(hf, ids).numpy()
    out["tiny_greedy"] = H.hf_generate(hf, ids, 12, min_new_tokens=12).numpy()

    # ---- GPT-2 small ----
    w = O.synthetic_weights(O.SMALL, seed=0)
    hf = H.build_hf_model(O.SMALL, w, attn="eager")
    ids = O.synthetic_prompts(8, 64)
    logits = H.hf_logits(hf, ids[:2])[:, -1, :]
    out["small_ids"] = ids.numpy()
    out["small_last_logits_stride25"] = logits[:, ::25].numpy()
    top = logits.topk(64, dim=-1)
    out["small_top64_idx"] = top.indices.numpy()
    out["small_top64_val"] = top.values.numpy()
    toks = H.hf_generate(hf, ids, 32, min_new_tokens=32)
    out["small_greedy"] = toks.numpy()
    _, _, margins = O.greedy_generate(O.SMALL, w, ids, 32, min_new_tokens=32)
    out["small_margins"] = margins.numpy()
    hf_sdpa = H.build_hf_model(O.SMALL, w, attn="sdpa")
    out["small_last_logits_stride25_sdpa"] = H.hf_logits(hf_sdpa, ids[:2])[:, -1, ::25].numpy()
    OUT.parent.mkdir(parents=True, exist_ok=True)
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
