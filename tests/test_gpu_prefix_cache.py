"""Shared-prefix KV reuse through the C ABI.

The reference's Summarize prompt is a static head followed by every message of the channel in
order (llm_server.py:273-285) and the Raft node sends the last 20 (raft_node.py:2032-2084), so
consecutive summaries of a growing channel share almost their whole prompt; with the byte-level
tokenizer those prompts are 500-1000 tokens. With cfg.prefix_cache the engine keeps every full
16-token page of a prompt in a ref-counted cache keyed by the token chain, maps the longest cached
page prefix into a new slot and prefills the suffix only (kv_scatter + paged prefill attention).

Gates: the warm run's last-position logits equal the cold run's within 2e-2 and the oracle's
within LOGIT_TOL; greedy tokens equal the oracle's wherever its margin is decisive; page
accounting balances; eviction frees cached pages under pressure.
"""
import numpy as np
import pytest
import torch

from conftest import pkg
from oracle import gpt2_oracle as O

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(600)]

LOGIT_TOL = 0.05
TINY = O.GPT2Shape(n_layer=2, n_head=2, d_model=128, vocab_size=1000, n_positions=256, eos_token_id=999)


def _engine(shape, weights, **kw):
    E = pkg("engine")
    ms = E.ModelShape(n_layer=shape.n_layer, n_head=shape.n_head, d_model=shape.d_model,
                      vocab_size=shape.vocab_size, n_positions=shape.n_positions,
                      ln_eps=shape.ln_eps, eos_token_id=shape.eos_token_id)
    return E.Engine(ms, O.pack_blob(shape, weights).numpy(), **kw), E


def _margin_check(got, ref_tok, margins, tol=2 * LOGIT_TOL):
    got, ref = np.asarray(got), ref_tok.numpy()
    same_prefix = np.cumprod(np.concatenate([np.ones((got.shape[0], 1), bool), got[:, :-1] == ref[:, :-1]], 1), 1).astype(bool)
    m = same_prefix & (margins.numpy() > tol)
    assert (got[m] == ref[m]).all(), (got, ref)


def test_second_summary_of_a_growing_channel_prefills_only_its_suffix(cuda):
    shape = O.SMALL
    w = O.synthetic_weights(shape, seed=0)
    base = O.synthetic_prompts(1, 420, seed=44)[0].tolist()
    p1, p2 = base[:300], base[:300] + base[330:393]          # 300 shared tokens -> 18 full pages = 288 cached
    N = 12
    ref_tok, ref_logits, margins = O.greedy_generate(shape, w, torch.tensor([p2]), N, min_new_tokens=N)
    kw = dict(max_seqs=8, max_seq_len=1024, max_batch_tokens=2048, max_new_tokens=N, keep_logits=True)

    cold, E = _engine(shape, w, **kw)
    sp = E.SamplingParams(max_new_tokens=N, min_new_tokens=N)
    s = cold.seq_alloc(len(p2) + N)
    cold.prefill([p2], [s], [sp])
    cold_logits = cold.read_last_logits(1)[0]
    cold.decode([s], N - 1)
    cold_tok = cold.read_tokens(s)[0]
    cold.close()

    warm, E = _engine(shape, w, prefix_cache=True, **kw)
    s1 = warm.seq_alloc(len(p1) + N, prompt=p1)
    assert warm.last_cached_tokens == 0
    warm.prefill([p1], [s1], [sp])
    warm.decode([s1], 3)
    warm.sync()
    # the first request is still decoding when the second arrives: its full prompt pages are shared
    s2 = warm.seq_alloc(len(p2) + N, prompt=p2)
    assert warm.last_cached_tokens == 288
    warm.prefill([p2], [s2], [sp])                          # 75 rows instead of 363
    warm_logits = warm.read_last_logits(1)[0]
    warm.decode([s1, s2], N - 1)
    warm_tok = warm.read_tokens(s2)[0]
    st = warm.prefix_cache_stats()
    assert st["hit_tokens"] == 288 and st["lookup_tokens"] == len(p1) + len(p2)
    assert st["cached_pages"] == len(p2) // 16              # 18 shared + 4 more full pages of p2
    free_before, total = warm.kv_pages()
    warm.seq_free(s1)
    warm.seq_free(s2)
    free_after, _ = warm.kv_pages()
    assert free_after + st["cached_pages"] == total         # every page is free or held by the cache
    warm.close()

    assert np.abs(warm_logits - cold_logits).max() <= 2e-2
    assert np.abs(warm_logits - ref_logits[0, 0].numpy()).max() <= LOGIT_TOL
    _margin_check([warm_tok], ref_tok, margins)
    _margin_check([cold_tok], ref_tok, margins)
    print(f"prefix cache: warm vs cold max |dlogit| {np.abs(warm_logits - cold_logits).max():.4f}, "
          f"tokens equal {np.mean(np.array(warm_tok) == np.array(cold_tok)):.2f}")


def test_mixed_batch_duplicates_and_generate_api(cuda):
    """One prefill call with a cached-prefix row next to cold rows; identical prompts inside one
    batch; b200llm_generate (host ids) uses the cache by itself."""
    w = O.synthetic_weights(TINY, seed=3)
    N = 6
    a = O.synthetic_prompts(1, 70, vocab=998, seed=1)[0].tolist()
    b = a[:50] + O.synthetic_prompts(1, 30, vocab=998, seed=2)[0].tolist()      # shares 3 pages with a
    c = O.synthetic_prompts(1, 33, vocab=998, seed=3)[0].tolist()
    eng, E = _engine(TINY, w, max_seqs=8, max_seq_len=128, max_batch_tokens=512, max_new_tokens=N, prefix_cache=True)
    sp = E.SamplingParams(max_new_tokens=N, min_new_tokens=N)
    refs = {}
    for name, p in (("a", a), ("b", b), ("c", c)):
        refs[name] = O.greedy_generate(TINY, w, torch.tensor([p]), N, min_new_tokens=N)
    out = eng.generate([a], sp)                             # fills the cache
    _margin_check(out, refs["a"][0], refs["a"][2], 0.04)
    slots, cached = [], []
    for p in (b, c, b):                                     # b twice in one batch
        slots.append(eng.seq_alloc(len(p) + N, prompt=p))
        cached.append(eng.last_cached_tokens)
    assert cached == [48, 0, 48]
    eng.prefill([b, c, b], slots, [sp] * 3)
    eng.decode(slots, N - 1)
    toks = [eng.read_tokens(s)[0] for s in slots]
    _margin_check([toks[2]], refs["b"][0], refs["b"][2], 0.04)   # the same prompt twice in one batch
    _margin_check([toks[0]], refs["b"][0], refs["b"][2], 0.04)
    _margin_check([toks[1]], refs["c"][0], refs["c"][2], 0.04)
    for s in slots:
        eng.seq_free(s)
    hits0 = eng.prefix_cache_stats()["hit_tokens"]
    out = eng.generate([b, a], sp)                          # both fully cached except their tails
    assert eng.prefix_cache_stats()["hit_tokens"] - hits0 == (len(b) - 1) // 16 * 16 + (len(a) - 1) // 16 * 16
    _margin_check([out[0]], refs["b"][0], refs["b"][2], 0.04)
    _margin_check([out[1]], refs["a"][0], refs["a"][2], 0.04)
    eng.close()


def test_cached_pages_are_evicted_under_pressure_and_never_while_in_use(cuda):
    w = O.synthetic_weights(TINY, seed=3)
    N = 4
    # pool of 12 pages; every prompt below fills 4 full pages (64 tokens) + needs 1 more for its tail
    eng, E = _engine(TINY, w, max_seqs=2, max_seq_len=96, max_batch_tokens=256, max_new_tokens=N,
                     kv_pool_tokens=12 * 16, prefix_cache=True)
    sp = E.SamplingParams(max_new_tokens=N, min_new_tokens=N)
    prompts = [O.synthetic_prompts(1, 66, vocab=998, seed=100 + i)[0].tolist() for i in range(5)]
    outs = [eng.generate([p], sp)[0] for p in prompts]      # 5 x 4 cached pages > 12: the oldest must go
    st = eng.prefix_cache_stats()
    free, total = eng.kv_pages()
    assert total == 12 and st["cached_pages"] <= 12 and free + st["cached_pages"] == total
    # the most recent prompt is still cached, the first one is not
    s = eng.seq_alloc(70, prompt=prompts[-1])
    assert eng.last_cached_tokens == 64
    eng.seq_free(s)
    s = eng.seq_alloc(70, prompt=prompts[0])
    assert eng.last_cached_tokens == 0
    # while that slot pins its pages, a request that needs the rest of the pool still fits ...
    s2 = eng.seq_alloc(96, prompt=prompts[1])
    eng.prefill([prompts[0], prompts[1]], [s, s2], [sp, sp])
    eng.decode([s, s2], N - 1)
    again = eng.read_tokens(s)[0]
    ref = O.greedy_generate(TINY, w, torch.tensor([prompts[0]]), N, min_new_tokens=N)
    _margin_check([again], ref[0], ref[2], 0.04)            # re-computed after its pages were evicted
    _margin_check([outs[0]], ref[0], ref[2], 0.04)
    eng.seq_free(s)
    eng.seq_free(s2)
    # ... and a pool in which nothing can be evicted reports exhaustion instead of corrupting pages
    a = eng.seq_alloc(96)
    b = eng.seq_alloc(96)
    B200Error = pkg("_lib").B200Error
    with pytest.raises(B200Error):
        eng.seq_alloc(16)       # no slot left (and no page either)
    eng.seq_free(a)
    eng.seq_free(b)
    free, total = eng.kv_pages()
    assert free + eng.prefix_cache_stats()["cached_pages"] == total
    eng.close()


def test_batched_token_read_and_cancel(cuda):
    """b200llm_read_tokens_batch returns what read_status + read_tokens return, with one
    synchronisation; b200llm_seq_cancel stops a row in stream order."""
    w = O.synthetic_weights(TINY, seed=3)
    eng, E = _engine(TINY, w, max_seqs=8, max_seq_len=128, max_batch_tokens=512, max_new_tokens=40)
    ids = O.synthetic_prompts(5, 9, vocab=998, seed=8).tolist()
    sps = [E.SamplingParams(max_new_tokens=n, min_new_tokens=n) for n in (3, 40, 8, 40, 5)]
    slots = [eng.seq_alloc(64) for _ in ids]
    eng.prefill(ids, slots, sps)
    eng.decode(slots, 7)
    counts, fin, rows = eng.read_tokens_batch(slots)
    assert counts.tolist() == [3, 8, 8, 8, 5] and fin.tolist() == [1, 0, 1, 0, 1]
    for s, n, f, row in zip(slots, counts, fin, rows):
        one, f1 = eng.read_tokens(s)
        assert f1 == bool(f) and len(one) == n
        assert (row == one) if f else (row is None)
    _, _, all_rows = eng.read_tokens_batch(slots, only_finished=False)
    assert all_rows[1] == eng.read_tokens(slots[1])[0]
    # cancel row 1: it stops where it is; row 3 keeps going
    eng.seq_cancel(slots[1])
    eng.decode(slots, 6)
    counts, fin, rows = eng.read_tokens_batch(slots)
    assert fin.tolist() == [1, 1, 1, 0, 1] and counts[3] == 14
    # a finished row emits pad from then on ($TF/generation/utils.py:2796-2797): nothing new after the cancel
    assert rows[1][:8] == all_rows[1] and all(t == TINY.eos_token_id for t in rows[1][8:])
    for s in slots:
        eng.seq_free(s)
    eng.close()
