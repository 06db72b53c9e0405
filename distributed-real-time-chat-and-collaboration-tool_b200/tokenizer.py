"""GPT-2 tokenisation for the service.

No vocab.json / merges.txt exist offline, so the default is the zero-merge
byte-level tokenizer: every UTF-8 byte maps to the id GPT-2's vocabulary gives that single byte
(ids 0..255 follow GPT-2's bytes_to_unicode order: '!' -> 0, '\\n' -> 198, ' ' -> 220). That is a
valid GPT-2 tokenisation (just not the shortest one) and it round-trips. When the two vocabulary
files are supplied, ``GPT2BPETokenizer`` applies the real byte-pair merges.
"""
from __future__ import annotations

import json
from functools import lru_cache
from typing import Iterable, Sequence

EOS_ID = 50256


@lru_cache(maxsize=1)
def bytes_to_unicode() -> dict[int, str]:
    """GPT-2's reversible byte -> printable unicode table."""
    bs = list(range(ord("!"), ord("~") + 1)) + list(range(ord("¡"), ord("¬") + 1)) + list(range(ord("®"), ord("ÿ") + 1))
    cs = bs[:]
    n = 0
    for b in range(256):
        if b not in bs:
            bs.append(b)
            cs.append(256 + n)
            n += 1
    return dict(zip(bs, (chr(c) for c in cs)))


@lru_cache(maxsize=1)
def _byte_ids() -> tuple[list[int], dict[int, int]]:
    """byte -> id and id -> byte for the 256 single-byte tokens (vocab order = table order)."""
    order = list(bytes_to_unicode().keys())  # insertion order is the vocabulary order
    byte_to_id = [0] * 256
    for idx, b in enumerate(order):
        byte_to_id[b] = idx
    return byte_to_id, {idx: b for idx, b in enumerate(order)}


class ByteLevelTokenizer:
    """Zero-merge GPT-2 tokenizer (one token per UTF-8 byte)."""

    eos_token_id = EOS_ID
    vocab_size = 50257

    def encode(self, text: str) -> list[int]:
        table, _ = _byte_ids()
        return [table[b] for b in text.encode("utf-8")]

    def decode(self, ids: Iterable[int]) -> str:
        _, inverse = _byte_ids()
        out = bytearray()
        for i in ids:
            if i == self.eos_token_id:
                break
            b = inverse.get(int(i))
            if b is not None:
                out.append(b)
            # ids >= 256 are multi-byte merges whose spelling lives in vocab.json; without that
            # file they cannot be rendered and are skipped.
        return out.decode("utf-8", errors="replace")


class GPT2BPETokenizer:
    """Byte-pair tokenizer driven by GPT-2's vocab.json + merges.txt (when supplied)."""

    eos_token_id = EOS_ID

    def __init__(self, vocab: dict[str, int], merges: Sequence[tuple[str, str]]):
        import regex

        self.encoder = dict(vocab)
        self.decoder = {v: k for k, v in self.encoder.items()}
        self.ranks = {tuple(m): i for i, m in enumerate(merges)}
        self.byte_enc = bytes_to_unicode()
        self.byte_dec = {v: k for k, v in self.byte_enc.items()}
        self.vocab_size = max(self.decoder) + 1
        self.pattern = regex.compile(
            r"'s|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+")
        self._cache: dict[str, list[str]] = {}

    @classmethod
    def from_files(cls, vocab_json: str, merges_txt: str) -> "GPT2BPETokenizer":
        with open(vocab_json, encoding="utf-8") as f:
            vocab = json.load(f)
        with open(merges_txt, encoding="utf-8") as f:
            lines = [l.rstrip("\n") for l in f if l.strip() and not l.startswith("#version")]
        return cls(vocab, [tuple(l.split()) for l in lines])

    def _bpe(self, token: str) -> list[str]:
        if token in self._cache:
            return self._cache[token]
        word = list(token)
        while len(word) > 1:
            pairs = {(word[i], word[i + 1]) for i in range(len(word) - 1)}
            best = min(pairs, key=lambda p: self.ranks.get(p, float("inf")))
            if best not in self.ranks:
                break
            merged, i = [], 0
            while i < len(word):
                if i < len(word) - 1 and (word[i], word[i + 1]) == best:
                    merged.append(word[i] + word[i + 1])
                    i += 2
                else:
                    merged.append(word[i])
                    i += 1
            word = merged
        self._cache[token] = word
        return word

    def encode(self, text: str) -> list[int]:
        ids: list[int] = []
        for piece in self.pattern.findall(text):
            mapped = "".join(self.byte_enc[b] for b in piece.encode("utf-8"))
            ids.extend(self.encoder[t] for t in self._bpe(mapped))
        return ids

    def decode(self, ids: Iterable[int]) -> str:
        chars = []
        for i in ids:
            if i == self.eos_token_id:
                break
            chars.append(self.decoder.get(int(i), ""))
        data = bytearray(self.byte_dec[c] for c in "".join(chars) if c in self.byte_dec)
        return data.decode("utf-8", errors="replace")


def load_tokenizer(vocab_json: str | None = None, merges_txt: str | None = None):
    if vocab_json and merges_txt:
        return GPT2BPETokenizer.from_files(vocab_json, merges_txt)
    return ByteLevelTokenizer()
