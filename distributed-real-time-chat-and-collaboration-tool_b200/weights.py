"""Weight blobs for the engine (layout: include/b200llm.h).

No checkpoint can be downloaded in this environment, so serving and benchmarks run on
random-init weights of the right architecture; a Hugging Face GPT-2 state dict (or a
.safetensors file of one) is packed into the same blob when supplied.
"""
from __future__ import annotations

import math
from typing import Mapping

import numpy as np

from .engine import ModelShape

_LAYER = (
    ("ln_1.weight", lambda d: (d,)), ("ln_1.bias", lambda d: (d,)),
    ("attn.c_attn.weight", lambda d: (d, 3 * d)), ("attn.c_attn.bias", lambda d: (3 * d,)),
    ("attn.c_proj.weight", lambda d: (d, d)), ("attn.c_proj.bias", lambda d: (d,)),
    ("ln_2.weight", lambda d: (d,)), ("ln_2.bias", lambda d: (d,)),
    ("mlp.c_fc.weight", lambda d: (d, 4 * d)), ("mlp.c_fc.bias", lambda d: (4 * d,)),
    ("mlp.c_proj.weight", lambda d: (4 * d, d)), ("mlp.c_proj.bias", lambda d: (d,)),
)


def tensor_specs(shape: ModelShape) -> list[tuple[str, tuple[int, ...]]]:
    """(state-dict name, shape) in blob order. Conv1D weights keep HF's [in, out] layout."""
    d = shape.d_model
    specs = [("transformer.wte.weight", (shape.vocab_size, d)),
             ("transformer.wpe.weight", (shape.n_positions, d))]
    for layer in range(shape.n_layer):
        specs += [(f"transformer.h.{layer}.{n}", f(d)) for n, f in _LAYER]
    specs += [("transformer.ln_f.weight", (d,)), ("transformer.ln_f.bias", (d,))]
    return specs


def synthetic_blob(shape: ModelShape, seed: int = 0, hf_like: bool = True) -> np.ndarray:
    """Random-init blob following GPT-2's init scheme (N(0, 0.02); residual projections scaled
    by 1/sqrt(2 n_layer); LayerNorm 1/0; zero biases). hf_like=False also randomises biases and
    LayerNorm parameters."""
    rng = np.random.default_rng(seed)
    parts = []
    for name, shp in tensor_specs(shape):
        n = int(np.prod(shp))
        if name.endswith(("ln_1.weight", "ln_2.weight", "ln_f.weight")):
            t = np.ones(n, np.float32) if hf_like else 1 + 0.1 * rng.standard_normal(n, dtype=np.float32)
        elif name.endswith(".bias"):
            t = np.zeros(n, np.float32) if hf_like else 0.02 * rng.standard_normal(n, dtype=np.float32)
        else:
            std = 0.02 / math.sqrt(2 * shape.n_layer) if name.endswith("c_proj.weight") else 0.02
            t = rng.standard_normal(n, dtype=np.float32) * np.float32(std)
        parts.append(t.astype(np.float32, copy=False))
    return np.concatenate(parts)


def blob_from_state_dict(shape: ModelShape, state: Mapping[str, "np.ndarray"]) -> np.ndarray:
    """Pack a GPT2LMHeadModel state dict (numpy arrays or torch tensors). Keys may carry or
    omit the `transformer.` prefix; `lm_head.weight` is ignored (tied to wte)."""
    parts = []
    for name, shp in tensor_specs(shape):
        t = state.get(name, state.get(name.removeprefix("transformer.")))
        if t is None:
            raise KeyError(f"state dict lacks {name}")
        a = t.detach().cpu().float().numpy() if hasattr(t, "detach") else np.asarray(t, dtype=np.float32)
        if tuple(a.shape) != tuple(shp):
            raise ValueError(f"{name}: expected {shp}, got {tuple(a.shape)}")
        parts.append(np.ascontiguousarray(a, dtype=np.float32).reshape(-1))
    return np.concatenate(parts)


def blob_from_safetensors(shape: ModelShape, path: str) -> np.ndarray:
    from safetensors.numpy import load_file  # optional dependency, only when a file is given

    return blob_from_state_dict(shape, load_file(path))
