"""CPU restatement of the algorithm on the hot path: GPT-2 causal-LM forward + greedy decode.

TEST INFRASTRUCTURE -- checker only, never the thing measured or shipped.

Parity status: the reference repository contains NO implementation of this path (its generator
is a remote Gemini call, llm_server/llm_server.py:32-33,167,231,287,403) and NO tests, so the
reference itself pins nothing here ("parity unpinned" by the reference). The arithmetic lives in
the third-party dependency the reference pins and never imports: transformers==4.35.0 /
torch==2.2.0 (reference requirements.txt:6-7). This file restates that library's published
GPT-2 algorithm function by function and is pinned against the installed transformers 5.5.0
GPT2LMHeadModel (same math; tests/test_oracle.py runs both, tests/golden/ holds its outputs).

$TF below = site-packages/transformers (5.5.0).

Written with plain fp32 torch CPU tensor ops (matmul, exp, tanh); no transformers import.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch


@dataclass(frozen=True)
class GPT2Shape:
    """Subset of $TF/models/gpt2/configuration_gpt2.py GPT2Config the path depends on."""
    n_layer: int = 12
    n_head: int = 12
    d_model: int = 768
    vocab_size: int = 50257
    n_positions: int = 1024
    ln_eps: float = 1e-5
    eos_token_id: int = 50256

    @property
    def head_dim(self) -> int:
        return self.d_model // self.n_head


SMALL = GPT2Shape()
MEDIUM = GPT2Shape(n_layer=24, n_head=16, d_model=1024)
LARGE = GPT2Shape(n_layer=36, n_head=20, d_model=1280)
SHAPES = {"small": SMALL, "medium": MEDIUM, "large": LARGE}

# Order of tensors in the flat fp32 weight blob (include/b200llm.h documents the same order).
LAYER_TENSORS = (
    ("ln_1.weight", lambda d: (d,)),
    ("ln_1.bias", lambda d: (d,)),
    ("attn.c_attn.weight", lambda d: (d, 3 * d)),
    ("attn.c_attn.bias", lambda d: (3 * d,)),
    ("attn.c_proj.weight", lambda d: (d, d)),
    ("attn.c_proj.bias", lambda d: (d,)),
    ("ln_2.weight", lambda d: (d,)),
    ("ln_2.bias", lambda d: (d,)),
    ("mlp.c_fc.weight", lambda d: (d, 4 * d)),
    ("mlp.c_fc.bias", lambda d: (4 * d,)),
    ("mlp.c_proj.weight", lambda d: (4 * d, d)),
    ("mlp.c_proj.bias", lambda d: (d,)),
)


def tensor_specs(shape: GPT2Shape) -> list[tuple[str, tuple[int, ...]]]:
    d = shape.d_model
    specs = [("transformer.wte.weight", (shape.vocab_size, d)),
             ("transformer.wpe.weight", (shape.n_positions, d))]
    for layer in range(shape.n_layer):
        for name, fn in LAYER_TENSORS:
            specs.append((f"transformer.h.{layer}.{name}", fn(d)))
    specs += [("transformer.ln_f.weight", (d,)), ("transformer.ln_f.bias", (d,))]
    return specs


def synthetic_weights(shape: GPT2Shape, seed: int = 0, hf_like: bool = False) -> dict[str, torch.Tensor]:
    """Deterministic random-init weights (no checkpoint exists offline).

    Follows GPT2PreTrainedModel._init_weights ($TF/models/gpt2/modeling_gpt2.py:433-458):
    N(0, 0.02) for embeddings and Conv1D weights, c_proj weights scaled by 1/sqrt(2*n_layer).
    With hf_like=False (parity tests) biases and LayerNorm parameters are also randomised so
    that every bias / gamma / beta code path is exercised; hf_like=True keeps HF's zeros/ones.
    """
    g = torch.Generator().manual_seed(seed)
    out: dict[str, torch.Tensor] = {}
    for name, shp in tensor_specs(shape):
        if name.endswith("ln_1.weight") or name.endswith("ln_2.weight") or name.endswith("ln_f.weight"):
            t = torch.ones(shp) if hf_like else 1.0 + 0.1 * torch.randn(shp, generator=g)
        elif name.endswith(".bias"):
            t = torch.zeros(shp) if hf_like else 0.02 * torch.randn(shp, generator=g)
        else:
            std = 0.02
            if name.endswith("c_proj.weight"):
                std = 0.02 / math.sqrt(2 * shape.n_layer)
            t = std * torch.randn(shp, generator=g)
        out[name] = t.contiguous()
    return out


def pack_blob(shape: GPT2Shape, weights: dict[str, torch.Tensor]) -> torch.Tensor:
    """Flat fp32 blob in the C ABI's order."""
    return torch.cat([weights[name].reshape(-1).float() for name, _ in tensor_specs(shape)])


def unpack_blob(shape: GPT2Shape, blob) -> dict[str, torch.Tensor]:
    """Inverse of pack_blob: the state dict a flat fp32 blob (C-ABI order) stands for."""
    flat = torch.as_tensor(blob, dtype=torch.float32).reshape(-1)
    out, at = {}, 0
    for name, shp in tensor_specs(shape):
        n = int(math.prod(shp))
        out[name] = flat[at:at + n].reshape(shp)
        at += n
    assert at == flat.numel(), "blob size does not match the shape"
    return out


def synthetic_prompts(batch: int, length: int, vocab: int = 50257, seed: int = 1234) -> torch.Tensor:
    """Token ids uniform in [0, vocab-1) (EOS excluded), SURVEY.md section 8d."""
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, vocab - 1, (batch, length), generator=g)


# ----------------------------------------------------------------------------------------
# Operators
# ----------------------------------------------------------------------------------------
def conv1d(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor) -> torch.Tensor:
    """Conv1D.forward: addmm(bias, x, W) with W stored [in, out]  ($TF/pytorch_utils.py:119-122)."""
    return x @ weight + bias


def layer_norm(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float) -> torch.Tensor:
    """nn.LayerNorm over the last dim: biased variance, eps inside the sqrt
    ($TF/models/gpt2/modeling_gpt2.py:252-254, 505)."""
    mean = x.mean(-1, keepdim=True)
    var = ((x - mean) ** 2).mean(-1, keepdim=True)
    return (x - mean) / torch.sqrt(var + eps) * gamma + beta


def gelu_new(x: torch.Tensor) -> torch.Tensor:
    """NewGELUActivation ($TF/activations.py:59-66)."""
    return 0.5 * x * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (x + 0.044715 * torch.pow(x, 3.0))))


def attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
    """eager_attention_forward ($TF/models/gpt2/modeling_gpt2.py:54-72): scores scaled by
    head_dim**-0.5, causal mask, fp32 softmax, weighted sum.
    q [B,H,T,dh]; k, v [B,H,S,dh] with the T queries occupying the last T of S positions."""
    dh = q.shape[-1]
    scores = (q @ k.transpose(-1, -2)) / math.sqrt(dh)
    T, S = q.shape[-2], k.shape[-2]
    causal = torch.ones(T, S, dtype=torch.bool).tril(S - T)
    scores = scores.masked_fill(~causal, torch.finfo(scores.dtype).min)
    return torch.softmax(scores, dim=-1) @ v


def split_heads(x: torch.Tensor, n_head: int) -> torch.Tensor:
    """[B,T,d] -> [B,H,T,dh]  ($TF/models/gpt2/modeling_gpt2.py:185-191)."""
    B, T, d = x.shape
    return x.view(B, T, n_head, d // n_head).transpose(1, 2)


class KVCache:
    """DynamicCache growth by concatenation ($TF/cache_utils.py:119-120)."""

    def __init__(self, n_layer: int):
        self.k: list[torch.Tensor | None] = [None] * n_layer
        self.v: list[torch.Tensor | None] = [None] * n_layer

    def append(self, layer: int, k: torch.Tensor, v: torch.Tensor):
        self.k[layer] = k if self.k[layer] is None else torch.cat([self.k[layer], k], dim=-2)
        self.v[layer] = v if self.v[layer] is None else torch.cat([self.v[layer], v], dim=-2)
        return self.k[layer], self.v[layer]

    def length(self) -> int:
        return 0 if self.k[0] is None else self.k[0].shape[-2]


def block(shape: GPT2Shape, w: dict[str, torch.Tensor], layer: int, x: torch.Tensor, cache: KVCache) -> torch.Tensor:
    """GPT2Block.forward ($TF/models/gpt2/modeling_gpt2.py:262-309): pre-LN attention and MLP
    sub-blocks, each followed by a residual add."""
    p = f"transformer.h.{layer}."
    h = layer_norm(x, w[p + "ln_1.weight"], w[p + "ln_1.bias"], shape.ln_eps)          # :273
    qkv = conv1d(h, w[p + "attn.c_attn.weight"], w[p + "attn.c_attn.bias"])           # :185
    q, k, v = qkv.split(shape.d_model, dim=-1)
    q, k, v = (split_heads(t, shape.n_head) for t in (q, k, v))
    k, v = cache.append(layer, k, v)                                                    # :194-196
    a = attention(q, k, v)                                                              # :54-72
    a = a.transpose(1, 2).reshape(x.shape)                                              # :222
    x = x + conv1d(a, w[p + "attn.c_proj.weight"], w[p + "attn.c_proj.bias"])         # :223, :282
    h = layer_norm(x, w[p + "ln_2.weight"], w[p + "ln_2.bias"], shape.ln_eps)          # :304
    f = gelu_new(conv1d(h, w[p + "mlp.c_fc.weight"], w[p + "mlp.c_fc.bias"]))         # :239-240
    return x + conv1d(f, w[p + "mlp.c_proj.weight"], w[p + "mlp.c_proj.bias"])        # :241, :307


def forward(shape: GPT2Shape, w: dict[str, torch.Tensor], ids: torch.Tensor, cache: KVCache | None = None,
            last_only: bool = False) -> torch.Tensor:
    """GPT2LMHeadModel.forward ($TF/models/gpt2/modeling_gpt2.py:658-725) on an unpadded
    batch ids [B,T]: positions continue from the cache length (:560-585), L blocks (:616-626),
    ln_f (:628), tied LM head without bias (:646-651, :705-706). Returns fp32 logits."""
    cache = cache if cache is not None else KVCache(shape.n_layer)
    B, T = ids.shape
    past = cache.length()
    pos = torch.arange(past, past + T)
    x = w["transformer.wte.weight"][ids] + w["transformer.wpe.weight"][pos]            # :577-585
    for layer in range(shape.n_layer):
        x = block(shape, w, layer, x, cache)
    x = layer_norm(x, w["transformer.ln_f.weight"], w["transformer.ln_f.bias"], shape.ln_eps)
    if last_only:
        x = x[:, -1:, :]
    return x @ w["transformer.wte.weight"].t()


def greedy_generate(shape: GPT2Shape, w: dict[str, torch.Tensor], ids: torch.Tensor, max_new_tokens: int,
                    min_new_tokens: int = 0, pad_token_id: int | None = None,
                    teacher: torch.Tensor | None = None):
    """GenerationMixin._sample with do_sample=False ($TF/generation/utils.py:2658-2841) for one
    unpadded batch of equal-length prompts: prefill (:2736), then per step the fp32
    last-position logits (:2762), argmax (:2793), pad for finished rows (:2796-2797), EOS
    bookkeeping ($TF/generation/stopping_criteria.py:450-471), stop when all rows finished.
    MinNewTokensLengthLogitsProcessor masks EOS until min_new_tokens exist.

    If `teacher` [B, max_new_tokens] is given, the fed-back token is the teacher's (teacher
    forcing) while the returned tokens are still this model's argmax.
    Returns (tokens [B, n], logits [B, n, V], margins [B, n]) -- margin = top1 - top2.
    """
    pad = shape.eos_token_id if pad_token_id is None else pad_token_id
    cache = KVCache(shape.n_layer)
    B = ids.shape[0]
    unfinished = torch.ones(B, dtype=torch.long)
    toks, all_logits, margins = [], [], []
    cur = ids
    for step in range(max_new_tokens):
        logits = forward(shape, w, cur, cache, last_only=True)[:, -1, :].float()
        scores = logits.clone()
        if step < min_new_tokens:
            scores[:, shape.eos_token_id] = -float("inf")
        top2 = scores.topk(2, dim=-1).values
        nxt = scores.argmax(dim=-1)
        nxt = nxt * unfinished + pad * (1 - unfinished)
        toks.append(nxt)
        all_logits.append(logits)
        margins.append(top2[:, 0] - top2[:, 1])
        unfinished = unfinished & (nxt != shape.eos_token_id).long()
        if unfinished.max() == 0:
            break
        fed = nxt if teacher is None else teacher[:, step]
        cur = fed[:, None]
    return torch.stack(toks, 1), torch.stack(all_logits, 1), torch.stack(margins, 1)
