"""Engine-vs-oracle parity at the shapes BASELINE.json's configs name (through the C ABI).

The decode path changes behaviour with shape: split-K factors depend on d_model, the KV split
count and its combine (inside a thread-block cluster or by a second kernel) on rows x heads x
context, every batch bucket captures its own CUDA graphs, and contexts past 1024 need the
2048-row position table of SURVEY.md 0.5. Each case below runs the engine exactly as bench.py /
the scheduler drive it (default flags: graphs, PDL, 4-step graphs, fused argmax without logits)
and then once more step by step with logits kept, and checks both against the CPU oracle.

The oracle is teacher-forced with the ENGINE's tokens (oracle.greedy_generate(teacher=...)), so
every step of every checked row is compared on the same prefix and a low-margin divergence
cannot hide the steps after it:
  (i)   max |logit - oracle logit| <= LOGIT_TOL at every step;
  (ii)  the engine's token equals the oracle's argmax wherever the oracle's top-1/top-2 margin
        exceeds 2 x LOGIT_TOL, and in >= 97 % of all positions;
  (iii) the default-flag run and the step-by-step run agree wherever the margin is decisive.
Oracle: $TF/models/gpt2/modeling_gpt2.py:522-725 via oracle/gpt2_oracle.py. Callers these
configs serve: reference server/raft_node.py:2084 (summaries), :2126 (ask-AI).
"""
import numpy as np
import pytest
import torch

from conftest import pkg
from oracle import gpt2_oracle as O

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(900)]

LOGIT_TOL = 0.05


def _shape(name, n_positions=None):
    s = O.SHAPES[name]
    if n_positions:
        s = O.GPT2Shape(**{**s.__dict__, "n_positions": n_positions})
    return s


def _engine(shape, weights, **kw):
    E = pkg("engine")
    ms = E.ModelShape(n_layer=shape.n_layer, n_head=shape.n_head, d_model=shape.d_model,
                      vocab_size=shape.vocab_size, n_positions=shape.n_positions,
                      ln_eps=shape.ln_eps, eos_token_id=shape.eos_token_id)
    return E.Engine(ms, O.pack_blob(shape, weights).numpy(), **kw), E


_oracle_memo = {}


def _check_rows(shape, w, ids, rows, tokens, logits, label):
    """tokens [B, N] (engine); logits [len(rows), N, V] of the checked rows, or None. The oracle
    runs on `rows` only (and once per distinct teacher sequence)."""
    rows = list(rows)
    N = tokens.shape[1]
    teacher = torch.from_numpy(tokens[rows]).long()
    key = (label.split(" [")[0], teacher.numpy().tobytes())
    if key not in _oracle_memo:
        _oracle_memo.clear()
        _oracle_memo[key] = O.greedy_generate(shape, w, ids[rows], N, min_new_tokens=N, teacher=teacher)
    ref_tok, ref_logits, margins = _oracle_memo[key]
    assert ref_tok.shape[1] == N
    agree = (teacher == ref_tok).numpy()
    decisive = margins.numpy() > 2 * LOGIT_TOL
    assert agree[decisive].all(), f"{label}: engine token differs from the oracle's argmax at a decisive margin"
    assert agree.mean() >= 0.97, f"{label}: argmax agreement {agree.mean():.3f}"
    err = None
    if logits is not None:
        err = float(np.abs(logits - ref_logits.numpy()).max())
        assert err <= LOGIT_TOL, f"{label}: max |dlogit| {err}"
    print(f"{label}: rows {rows}, agreement {agree.mean():.3f}, decisive {decisive.mean():.2f}, max |dlogit| {err}")
    return margins.numpy()


def _run_both_ways(shape, w, ids, N, check_rows, label, **engine_kw):
    B, T = ids.shape
    cap = T + N
    prompts = ids.tolist()
    # (1) the way bench.py and the scheduler run it
    eng, E = _engine(shape, w, max_seqs=B, max_seq_len=cap, max_batch_tokens=min(B * T, 16384),
                     max_new_tokens=N, **engine_kw)
    sp = E.SamplingParams(max_new_tokens=N, min_new_tokens=N)
    chunk = max(1, min(B, 16384 // T))
    slots = [eng.seq_alloc(cap) for _ in range(B)]
    for c0 in range(0, B, chunk):
        eng.prefill(prompts[c0:c0 + chunk], slots[c0:c0 + chunk], [sp] * len(slots[c0:c0 + chunk]))
    eng.decode(slots, N - 1)
    fast = np.array([eng.read_tokens(s)[0] for s in slots])
    assert fast.shape == (B, N)
    launches = eng.kernel_launches
    assert launches > 0
    eng.close()
    margins = _check_rows(shape, w, ids, check_rows, fast, None, label + " [graphs, fused argmax]")
    # (2) step by step with logits kept
    eng, E = _engine(shape, w, max_seqs=B, max_seq_len=cap, max_batch_tokens=min(B * T, 16384),
                     max_new_tokens=N, keep_logits=True, **engine_kw)
    slots = [eng.seq_alloc(cap) for _ in range(B)]
    V = shape.vocab_size
    logits = np.empty((len(check_rows), N, V), np.float32)
    for c0 in range(0, B, chunk):
        eng.prefill(prompts[c0:c0 + chunk], slots[c0:c0 + chunk], [sp] * len(slots[c0:c0 + chunk]))
        lg = eng.read_last_logits(len(slots[c0:c0 + chunk]))
        for j, r in enumerate(check_rows):
            if c0 <= r < c0 + chunk:
                logits[j, 0] = lg[r - c0]
    for step in range(1, N):
        eng.decode(slots, 1)
        lg = eng.read_last_logits(B)
        logits[:, step] = lg[list(check_rows)]
    slow = np.array([eng.read_tokens(s)[0] for s in slots])
    eng.close()
    _check_rows(shape, w, ids, check_rows, slow, logits, label + " [step by step, logits]")
    # (3) the two runs against each other on the checked rows (same prefix => same decisive tokens)
    rows = list(check_rows)
    same_prefix = np.cumprod(np.concatenate([np.ones((len(rows), 1), bool),
                                             fast[rows][:, :-1] == slow[rows][:, :-1]], 1), 1).astype(bool)
    decisive = margins > 2 * LOGIT_TOL
    m = same_prefix & decisive
    assert (fast[rows][m] == slow[rows][m]).all(), f"{label}: graph run and step-by-step run disagree"
    print(f"{label}: graph run == step-by-step run on {(fast == slow).mean():.3f} of all tokens; {launches} launches")


def test_config2_small_b64_t128_n32(cuda):
    """BASELINE configs[1]: exactly the headline shape with the flags bench.py uses (bucket 64,
    4-step graphs). An 8-row slice is checked against the oracle."""
    shape = O.SMALL
    w = O.synthetic_weights(shape, seed=0)
    ids = O.synthetic_prompts(64, 128)
    _run_both_ways(shape, w, ids, 32, [0, 9, 18, 27, 36, 45, 54, 63], "small B=64 T=128")


@pytest.mark.parametrize("rows,T,check", [(1, 500, [0]), (2, 1024, [0, 1]), (16, 1024, [0, 5, 10, 15])],
                         ids=["1row-ctx500-cluster-combine", "2rows-ctx1024-16splits", "16rows-ctx1024-combine-kernel"])
def test_config3_medium_long_prompts_split_kv(cuda, rows, T, check):
    """BASELINE configs[2]: GPT-2 medium, n_positions 2048, 1024-token prompts whose contexts
    cross position 1024 while decoding; the engine picks split-KV attention (merged inside a
    thread-block cluster for one live row at context 500, by the combine kernel otherwise)."""
    shape = _shape("medium", 2048)
    w = O.synthetic_weights(shape, seed=0)
    ids = O.synthetic_prompts(rows, T, seed=21)
    _run_both_ways(shape, w, ids, 32, check, f"medium rows={rows} T={T}")


def test_config4_large_32_rows(cuda):
    """BASELINE configs[3], one GPU's shard: GPT-2 large, 32 rows (bucket 32, d = 1280 split-K
    picks), 64-token prompts, 36 new tokens through the 4-step graphs plus the remainder graph."""
    shape = O.LARGE
    w = O.synthetic_weights(shape, seed=0)
    ids = O.synthetic_prompts(32, 64, seed=33)
    _run_both_ways(shape, w, ids, 36, [0, 11, 22, 31], "large rows=32 T=64")


def test_streaming_attention_ragged_padded_finished_rows(cuda):
    """The streaming decode attention (the engine picks it by KV volume) in the situations the BASELINE
    shapes above never produce together: 40 live rows in the 64-row bucket (24 padding rows), ragged
    contexts of 300..520 tokens (tiles of 32 with ragged last pages, page lists longer than one tile),
    and rows that reach their own max_new_tokens early and stay in the batch as finished rows (no KV
    append) while the others go on. Checked rows -- finished and unfinished ones -- against the oracle,
    teacher-forced with the engine's tokens, one row at a time (ragged prompts)."""
    shape = O.SMALL
    w = O.synthetic_weights(shape, seed=0)
    B, N = 40, 14
    lens = [300 + (37 * i) % 221 for i in range(B)]
    prompts = [O.synthetic_prompts(1, L, seed=100 + i)[0] for i, L in enumerate(lens)]
    new = [4 if i % 3 == 0 else N for i in range(B)]          # every third row stops after 4 tokens
    cap = max(lens) + N
    eng, E = _engine(shape, w, max_seqs=B, max_seq_len=cap, max_batch_tokens=16384, max_new_tokens=N)
    sps = [E.SamplingParams(max_new_tokens=n, min_new_tokens=n) for n in new]
    slots = [eng.seq_alloc(cap) for _ in range(B)]
    c0 = 0
    while c0 < B:  # prompts packed into calls of at most 16384 tokens
        c1, tok = c0, 0
        while c1 < B and tok + lens[c1] <= 16384:
            tok += lens[c1]
            c1 += 1
        eng.prefill([p.tolist() for p in prompts[c0:c1]], slots[c0:c1], sps[c0:c1])
        c0 = c1
    eng.decode(slots, N - 1)
    got = [eng.read_tokens(s) for s in slots]
    eng.close()
    for i in range(B):
        assert len(got[i][0]) == new[i] and got[i][1], (i, len(got[i][0]), new[i])
    for r in (0, 1, 3, 8, 20, 39):  # rows 0, 3, 39 finish early
        n = new[r]
        teacher = torch.tensor([got[r][0]]).long()
        ref_tok, _, margins = O.greedy_generate(shape, w, prompts[r][None], n, min_new_tokens=n, teacher=teacher)
        agree = (teacher == ref_tok).numpy()
        decisive = margins.numpy() > 2 * LOGIT_TOL
        assert agree[decisive].all(), f"row {r} (context {lens[r]}): token differs from the oracle at a decisive margin"
        assert agree.mean() >= 0.9, f"row {r}: agreement {agree.mean():.2f}"
