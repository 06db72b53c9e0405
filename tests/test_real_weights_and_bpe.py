"""Real-checkpoint plumbing: the packers that `server.py --weights` and
`--vocab-json/--merges-txt` rely on, exercised against transformers' own objects.

No GPT-2 checkpoint or vocabulary can be downloaded here, so the "real" inputs are the ones
transformers itself produces offline: a GPT2LMHeadModel(GPT2Config(...)).state_dict() -- which has
everything a downloaded checkpoint has (`transformer.`-prefixed keys, the tied `lm_head.weight`,
possibly `attn.bias` / `attn.masked_bias` buffers) -- saved and re-read as .safetensors, and a
GPT2Tokenizer built from a vocab.json / merges.txt pair written by this test.
The reference pins transformers at requirements.txt:6-7; the parsers that need real text are
llm_server.py:251-260, :316-330, :435-449.
"""
import json
import os

import numpy as np
import pytest
import torch

from conftest import pkg
from oracle import gpt2_oracle as O

os.environ.setdefault("HF_HUB_OFFLINE", "1")

SHAPE = O.GPT2Shape(n_layer=2, n_head=4, d_model=256, vocab_size=3000, n_positions=128, eos_token_id=2999)


def _hf_model(seed=0):
    from transformers import GPT2Config, GPT2LMHeadModel

    torch.manual_seed(seed)
    cfg = GPT2Config(vocab_size=SHAPE.vocab_size, n_positions=SHAPE.n_positions, n_embd=SHAPE.d_model,
                     n_layer=SHAPE.n_layer, n_head=SHAPE.n_head, bos_token_id=SHAPE.eos_token_id,
                     eos_token_id=SHAPE.eos_token_id, attn_implementation="eager")
    model = GPT2LMHeadModel(cfg).eval()
    with torch.no_grad():  # HF initialises biases to 0 and LayerNorm to 1/0: randomise so they matter
        for name, p in model.named_parameters():
            if name.endswith(".bias") or "ln_" in name:
                p.add_(0.05 * torch.randn_like(p))
    return model


def _model_shape():
    E = pkg("engine")
    return E.ModelShape(n_layer=SHAPE.n_layer, n_head=SHAPE.n_head, d_model=SHAPE.d_model,
                        vocab_size=SHAPE.vocab_size, n_positions=SHAPE.n_positions, eos_token_id=SHAPE.eos_token_id)


def test_state_dict_packs_into_the_blob_layout_and_ignores_the_extra_keys(tmp_path):
    W = pkg("weights")
    model = _hf_model()
    sd = model.state_dict()
    assert "lm_head.weight" in sd and all(k.startswith(("transformer.", "lm_head.")) for k in sd)
    sd = dict(sd)
    # older checkpoints also carry the causal-mask buffers; the packer must not trip over them
    sd["transformer.h.0.attn.bias"] = torch.ones(1, 1, 8, 8)
    sd["transformer.h.0.attn.masked_bias"] = torch.tensor(-1e4)
    blob = W.blob_from_state_dict(_model_shape(), sd)
    lib = pkg("_lib")
    import ctypes
    cfg = lib.Config(n_layer=SHAPE.n_layer, n_head=SHAPE.n_head, d_model=SHAPE.d_model,
                     vocab_size=SHAPE.vocab_size, n_positions=SHAPE.n_positions)
    assert blob.dtype == np.float32 and blob.size == lib.lib.b200llm_weight_blob_floats(ctypes.byref(cfg))
    # the oracle's packer (tests only) lays the same tensors out identically
    ref = O.pack_blob(SHAPE, {k: v for k, v in model.state_dict().items()}).numpy()
    assert np.array_equal(blob, ref)
    # keys without the `transformer.` prefix (GPT2Model checkpoints) pack to the same blob
    bare = {k.removeprefix("transformer."): v for k, v in model.state_dict().items() if k != "lm_head.weight"}
    assert np.array_equal(W.blob_from_state_dict(_model_shape(), bare), blob)
    # .safetensors round trip (tied tensors must be saved once: drop lm_head like HF's own save does)
    from safetensors.torch import save_file
    path = tmp_path / "model.safetensors"
    save_file({k: v.contiguous() for k, v in model.state_dict().items() if k != "lm_head.weight"}, str(path))
    assert np.array_equal(W.blob_from_safetensors(_model_shape(), str(path)), blob)
    # a wrong shape or a missing tensor is an error, not a silent mis-pack
    bad = dict(model.state_dict())
    bad["transformer.h.1.mlp.c_fc.weight"] = bad["transformer.h.1.mlp.c_fc.weight"].t().contiguous()
    with pytest.raises(ValueError):
        W.blob_from_state_dict(_model_shape(), bad)
    del bad["transformer.ln_f.bias"]
    with pytest.raises((KeyError, ValueError)):
        W.blob_from_state_dict(_model_shape(), bad)


def test_oracle_on_the_hf_state_dict_equals_the_hf_model():
    """The oracle consumes the same state dict the packer does: fp32 logits equal HF's."""
    model = _hf_model(seed=1)
    ids = O.synthetic_prompts(2, 20, vocab=SHAPE.vocab_size, seed=4)
    with torch.no_grad():
        hf = model(input_ids=ids).logits.float()
    mine = O.forward(SHAPE, {k: v.float() for k, v in model.state_dict().items()}, ids)
    assert (hf - mine).abs().max().item() <= 2e-4


@pytest.mark.gpu
@pytest.mark.timeout(300)
def test_engine_on_an_hf_state_dict_matches_the_hf_model(cuda, tmp_path):
    """HF model -> .safetensors -> blob_from_safetensors -> engine (the `server.py --weights`
    path) against the same HF model's logits and greedy tokens."""
    E, W = pkg("engine"), pkg("weights")
    from safetensors.torch import save_file
    model = _hf_model(seed=2)
    path = tmp_path / "m.safetensors"
    save_file({k: v.contiguous() for k, v in model.state_dict().items() if k != "lm_head.weight"}, str(path))
    blob = W.blob_from_safetensors(_model_shape(), str(path))
    ids = O.synthetic_prompts(3, 24, vocab=SHAPE.vocab_size - 1, seed=6)
    N = 8
    with torch.no_grad():
        ref_logits = model(input_ids=ids).logits.float()
        ref_tok = model.generate(input_ids=ids, attention_mask=torch.ones_like(ids), do_sample=False,
                                 max_new_tokens=N, min_new_tokens=N, pad_token_id=SHAPE.eos_token_id)[:, ids.shape[1]:]
    with E.Engine(_model_shape(), blob, max_seqs=4, max_seq_len=64, max_batch_tokens=128, max_new_tokens=N) as eng:
        logits = torch.zeros(ids.numel(), SHAPE.vocab_size, device=cuda)
        slots = [eng.seq_alloc(64) for _ in range(3)]
        eng.prefill(ids.tolist(), slots, logits_all_ptr=logits.data_ptr())
        eng.sync()
        err = (logits.cpu().view(3, 24, -1) - ref_logits).abs().max().item()
        print(f"engine on the HF state dict: max |dlogit| vs GPT2LMHeadModel = {err:.4f}")
        assert err <= 0.05
        for s in slots:
            eng.seq_free(s)
        got = np.array(eng.generate(ids.tolist(), E.SamplingParams(max_new_tokens=N, min_new_tokens=N)))
    ref = ref_tok.numpy()
    # the oracle on the same state dict reproduces HF's greedy tokens and supplies the margins
    sd = {k: v.float() for k, v in model.state_dict().items()}
    o_tok, _, margins = O.greedy_generate(SHAPE, sd, ids, N, min_new_tokens=N)
    assert (o_tok.numpy() == ref).all()
    same_prefix = np.cumprod(np.concatenate([np.ones((3, 1), bool), got[:, :-1] == ref[:, :-1]], 1), 1).astype(bool)
    decisive = margins.numpy() > 0.1
    assert (got[same_prefix & decisive] == ref[same_prefix & decisive]).all(), (got, ref)


PARAGRAPH = ("Summary: The team's 3 meetings (at 10:30, 14:00 & 16:45) were re-scheduled -- again!\n"
             "Key Points:\n- budget is $1,250.75\t(approx.)\n-   naïve café déjà-vu 你好 🙂\n"
             "I'll say we're done; they've said it's  fine.   ")


def _toy_vocab(tmp_path):
    """A GPT-2 style vocab.json / merges.txt: the 256 byte tokens plus merges learnt greedily from
    the paragraph itself (so multi-byte UTF-8, digits, punctuation and leading spaces all merge)."""
    tok = pkg("tokenizer")
    import regex
    b2u = tok.bytes_to_unicode()
    vocab = {b2u[b]: i for i, b in enumerate(b2u)}
    pat = regex.compile(r"'s|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+")
    words = [["".join(b2u[b] for b in ch.encode("utf-8")) for ch in piece] for piece in pat.findall(PARAGRAPH * 2)]
    words = [[c for part in w for c in part] for w in words]
    merges = []
    for _ in range(120):
        counts = {}
        for w in words:
            for a, b in zip(w, w[1:]):
                counts[(a, b)] = counts.get((a, b), 0) + 1
        if not counts:
            break
        best = max(sorted(counts), key=lambda p: counts[p])
        merges.append(best)
        vocab.setdefault(best[0] + best[1], len(vocab))
        new_words = []
        for w in words:
            out, i = [], 0
            while i < len(w):
                if i < len(w) - 1 and (w[i], w[i + 1]) == best:
                    out.append(w[i] + w[i + 1])
                    i += 2
                else:
                    out.append(w[i])
                    i += 1
            new_words.append(out)
        words = new_words
    vocab["<|endoftext|>"] = len(vocab)
    vj, mt = tmp_path / "vocab.json", tmp_path / "merges.txt"
    vj.write_text(json.dumps(vocab, ensure_ascii=False), encoding="utf-8")
    mt.write_text("#version: 0.2\n" + "\n".join(f"{a} {b}" for a, b in merges) + "\n", encoding="utf-8")
    return str(vj), str(mt), len(merges)


def test_bpe_tokenizer_equals_transformers_gpt2_tokenizer_on_the_same_files(tmp_path):
    tok = pkg("tokenizer")
    vj, mt, n_merges = _toy_vocab(tmp_path)
    assert n_merges >= 100
    mine = tok.load_tokenizer(vj, mt)
    from transformers.models.gpt2.tokenization_gpt2 import GPT2Tokenizer
    # transformers 5.x: GPT2Tokenizer is the `tokenizers` BPE + ByteLevel pre-tokenizer and takes the
    # parsed contents of the two files
    vocab = json.load(open(vj, encoding="utf-8"))
    merges = [tuple(l.split()) for l in open(mt, encoding="utf-8").read().splitlines()[1:] if l.strip()]
    hf = GPT2Tokenizer(vocab=vocab, merges=merges)
    texts = [PARAGRAPH, "hello world", " leading space", "trailing space ", "a\n\nb", "1234567890", "don't can't I'm",
             "tabs\tand  double  spaces", "ünïcödé — dash", ""]
    for t in texts:
        want = hf.encode(t, add_special_tokens=False) if t else []
        assert mine.encode(t) == want, t
        assert mine.decode(want) == t
    ids = mine.encode(PARAGRAPH)
    assert len(ids) < 0.7 * len(PARAGRAPH.encode("utf-8")), "merges were not applied"
