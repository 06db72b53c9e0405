import json

from conftest import pkg

tok = pkg("tokenizer")


def test_byte_level_ids_match_gpt2_vocabulary_positions():
    t = tok.ByteLevelTokenizer()
    # published GPT-2 vocab ids of single-byte tokens (SURVEY.md 7.3 probed the same three)
    assert t.encode("!") == [0] and t.encode("\n") == [198] and t.encode(" ") == [220]
    assert t.encode("A") == [32] and t.encode("a") == [64] and t.encode("~") == [93]


def test_byte_level_round_trip_unicode_and_eos_stop():
    t = tok.ByteLevelTokenizer()
    text = "Summary: héllo — 你好\n- point 1\t✓"
    ids = t.encode(text)
    assert all(0 <= i < 256 for i in ids) and len(ids) == len(text.encode("utf-8"))
    assert t.decode(ids) == text
    assert t.decode(ids[:5] + [t.eos_token_id] + ids[5:]) == text.encode("utf-8")[:5].decode()
    assert t.decode([300, 40000] + ids[:3]) == text[:3]  # merges cannot be rendered without vocab.json


def test_bpe_tokenizer_applies_merges_in_rank_order(tmp_path):
    b2u = tok.bytes_to_unicode()
    base = {b2u[b]: i for i, b in enumerate(b2u)}
    sp = b2u[ord(" ")]
    merges = [("l", "o"), ("h", "e"), ("he", "l"), (sp, "w"), ("hel", "lo")]
    vocab = dict(base)
    for a, b in merges:
        vocab[a + b] = len(vocab)
    vj, mt = tmp_path / "vocab.json", tmp_path / "merges.txt"
    vj.write_text(json.dumps(vocab), encoding="utf-8")
    mt.write_text("#version: 0.2\n" + "\n".join(f"{a} {b}" for a, b in merges) + "\n", encoding="utf-8")
    t = tok.load_tokenizer(str(vj), str(mt))
    ids = t.encode("hello world")
    assert ids[0] == vocab["hello"] and vocab[sp + "w"] in ids
    assert t.decode(ids) == "hello world"
    assert len(ids) < len("hello world")
