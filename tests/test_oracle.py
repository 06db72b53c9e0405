"""Pins the CPU oracle (oracle/gpt2_oracle.py) against the golden vectors produced by Hugging
Face transformers (tests/golden/gpt2_golden.npz, generator: oracle/make_golden.py) and, for a
tiny model, against transformers run live. fp32 vs fp32: only summation order differs -> 1e-4."""
import os

import numpy as np
import pytest
import torch

from oracle import gpt2_oracle as O

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "gpt2_golden.npz"))
TINY = O.GPT2Shape(n_layer=2, n_head=2, d_model=128, vocab_size=1000, n_positions=96, eos_token_id=999)


def test_tiny_full_logits_match_golden():
    w = O.synthetic_weights(TINY, seed=3)
    ids = torch.from_numpy(GOLD["tiny_ids"])
    got = O.forward(TINY, w, ids).numpy()
    assert np.abs(got - GOLD["tiny_logits"]).max() <= 1e-4


def test_tiny_greedy_matches_golden_and_teacher_forcing_is_consistent():
    w = O.synthetic_weights(TINY, seed=3)
    ids = torch.from_numpy(GOLD["tiny_ids"])
    tok, logits, margins = O.greedy_generate(TINY, w, ids, 12, min_new_tokens=12)
    assert np.array_equal(tok.numpy(), GOLD["tiny_greedy"])
    tok2, logits2, _ = O.greedy_generate(TINY, w, ids, 12, min_new_tokens=12, teacher=tok)
    assert torch.equal(tok, tok2) and torch.allclose(logits, logits2)
    assert (margins >= 0).all()


@pytest.fixture(scope="module")
def small_weights():
    return O.synthetic_weights(O.SMALL, seed=0)


def test_small_last_position_logits_match_golden(small_weights):
    ids = torch.from_numpy(GOLD["small_ids"])[:2]
    logits = O.forward(O.SMALL, small_weights, ids, last_only=True)[:, -1, :]
    assert np.abs(logits[:, ::25].numpy() - GOLD["small_last_logits_stride25"]).max() <= 2e-4
    # eager vs sdpa attention inside transformers itself differ by less than that too
    assert np.abs(GOLD["small_last_logits_stride25_sdpa"] - GOLD["small_last_logits_stride25"]).max() <= 2e-4
    top = logits.topk(64, dim=-1)
    assert np.array_equal(top.indices.numpy()[:, :8], GOLD["small_top64_idx"][:, :8])
    assert np.abs(top.values.numpy() - GOLD["small_top64_val"]).max() <= 2e-4


def test_small_greedy_first_tokens_match_golden(small_weights):
    ids = torch.from_numpy(GOLD["small_ids"])[:4]
    tok, _, margins = O.greedy_generate(O.SMALL, small_weights, ids, 6, min_new_tokens=6)
    assert np.array_equal(tok.numpy(), GOLD["small_greedy"][:4, :6])
    assert np.abs(margins.numpy() - GOLD["small_margins"][:4, :6]).max() <= 1e-3


def test_oracle_matches_transformers_live_on_a_tiny_model():
    hf_ref = pytest.importorskip("oracle.hf_reference")
    w = O.synthetic_weights(TINY, seed=9)
    ids = O.synthetic_prompts(2, 17, vocab=1000, seed=4)
    hf = hf_ref.build_hf_model(TINY, w, attn="eager")
    assert (hf_ref.hf_logits(hf, ids) - O.forward(TINY, w, ids)).abs().max().item() <= 1e-4
    assert torch.equal(hf_ref.hf_generate(hf, ids, 5, min_new_tokens=5),
                       O.greedy_generate(TINY, w, ids, 5, min_new_tokens=5)[0])


def test_eos_handling_matches_transformers():
    """Rows that emit EOS are padded afterwards; generation stops when every row is finished
    (generation/utils.py:2796-2805)."""
    hf_ref = pytest.importorskip("oracle.hf_reference")
    w = dict(O.synthetic_weights(TINY, seed=3))
    wte = w["transformer.wte.weight"].clone()
    wte[TINY.eos_token_id] = 0.5
    w["transformer.wte.weight"] = wte
    w["transformer.ln_f.weight"] = torch.zeros(128)
    w["transformer.ln_f.bias"] = torch.ones(128)
    ids = torch.tensor([[1, 2, 3], [4, 5, 6]])
    hf = hf_ref.build_hf_model(TINY, w, attn="eager")
    ref = hf_ref.hf_generate(hf, ids, 6, min_new_tokens=2)
    got = O.greedy_generate(TINY, w, ids, 6, min_new_tokens=2)[0]
    assert torch.equal(ref, got) and got.shape[1] == 3 and (got[:, 2] == TINY.eos_token_id).all()


def test_blob_layout_matches_header_order_and_size():
    blob = O.pack_blob(TINY, O.synthetic_weights(TINY, seed=3))
    d, V, P, L = 128, 1000, 96, 2
    per_layer = 2 * d + d * 3 * d + 3 * d + d * d + d + 2 * d + d * 4 * d + 4 * d + 4 * d * d + d
    assert blob.numel() == V * d + P * d + L * per_layer + 2 * d
