"""Operator-level calls into the C ABI with torch tensors as device buffers.

torch is plumbing here (allocation + streams); every computation happens in libb200llm.so.
Used by the kernel parity tests and the profiling scripts.
"""
from __future__ import annotations

import torch

from ._lib import check, lib

EPI_BIAS_BF16 = 0
EPI_BIAS_GELU_BF16 = 1
EPI_BIAS_RESID_F32 = 2
EPI_F32 = 3
EPI_PARTIAL_F32 = 4


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: torch.Tensor | None) -> int | None:
    if t is None:
        return None
    assert t.is_cuda and t.is_contiguous(), "device-resident contiguous tensor required"
    return t.data_ptr()


def gemm(act: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None, out: torch.Tensor,
         epilogue: int, split_k: int = 1, orientation: int = -1) -> torch.Tensor:
    """out[token, feature] (+)= act[token, :] . w[feature, :]  (+ epilogue)."""
    tokens, k = act.shape
    features, k2 = w.shape
    assert k == k2 and act.dtype == torch.bfloat16 and w.dtype == torch.bfloat16
    ld_out = out.shape[-1]
    check(lib.b200llm_op_gemm(_ptr(act), _ptr(w), _ptr(bias), _ptr(out), tokens, features, k,
                              ld_out, epilogue, split_k, orientation, _stream()))
    return out


def layernorm(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float = 1e-5,
              partials: torch.Tensor | None = None, add_bias: torch.Tensor | None = None) -> torch.Tensor:
    rows, d = x.shape
    assert x.dtype == torch.float32
    out = torch.empty(rows, d, dtype=torch.bfloat16, device=x.device)
    n_partials = 0 if partials is None else partials.shape[0]
    check(lib.b200llm_op_layernorm(_ptr(x), _ptr(partials), n_partials, _ptr(add_bias), _ptr(gamma),
                                   _ptr(beta), _ptr(out), rows, d, eps, _stream()))
    return out


def embed(wte: torch.Tensor, wpe: torch.Tensor, ids: torch.Tensor, pos: torch.Tensor) -> torch.Tensor:
    n_tok = ids.numel()
    vocab, d = wte.shape
    x = torch.empty(n_tok, d, dtype=torch.float32, device=wte.device)
    check(lib.b200llm_op_embed(_ptr(wte), _ptr(wpe), _ptr(ids), _ptr(pos), _ptr(x), n_tok, d, vocab,
                               wpe.shape[0], _stream()))
    return x


def prefill_attention(qkv: torch.Tensor, cu_seqlens: torch.Tensor, max_seq_len: int,
                      n_head: int) -> torch.Tensor:
    n_tok = qkv.shape[0]
    d = n_head * 64
    assert qkv.shape[1] == 3 * d and qkv.dtype == torch.bfloat16
    out = torch.zeros(n_tok, d, dtype=torch.bfloat16, device=qkv.device)
    check(lib.b200llm_op_prefill_attention(_ptr(qkv), _ptr(cu_seqlens), cu_seqlens.numel() - 1,
                                           max_seq_len, n_head, _ptr(out), _stream()))
    return out


def decode_attention_scratch_floats(batch: int, n_head: int, n_splits: int) -> int:
    return int(lib.b200llm_decode_attention_scratch_floats(batch, n_head, n_splits))


def decode_attention(qkv: torch.Tensor, seq_len: torch.Tensor, page_table: torch.Tensor,
                     k_pool: torch.Tensor, v_pool: torch.Tensor, n_head: int, n_splits: int = 1,
                     scratch: torch.Tensor | None = None, cluster: bool = False,
                     streaming: bool = False, counters: torch.Tensor | None = None) -> torch.Tensor:
    """`cluster`: merge the 2..8 KV splits inside the kernel over a thread-block cluster (what the
    engine does for a few live rows) instead of through scratch + a combine launch. `streaming`: the
    persistent-warp / TMA form of the kernel (large batches, long contexts); with `counters` (zeroed
    int32 [batch * n_head]) its KV splits are merged inside the kernel, without the combine launch."""
    batch = qkv.shape[0]
    d = n_head * 64
    out = torch.zeros(batch, d, dtype=torch.bfloat16, device=qkv.device)
    if n_splits > 1 and scratch is None:
        scratch = torch.empty(decode_attention_scratch_floats(batch, n_head, n_splits),
                              dtype=torch.float32, device=qkv.device)
    if streaming:
        check(lib.b200llm_op_decode_attention_stream(_ptr(qkv), _ptr(seq_len), _ptr(page_table),
                                                     page_table.shape[1], _ptr(k_pool), _ptr(v_pool),
                                                     k_pool.shape[0], _ptr(out), batch, n_head, n_splits,
                                                     _ptr(scratch), _ptr(counters), _stream()))
        return out
    check(lib.b200llm_op_decode_attention(_ptr(qkv), _ptr(seq_len), _ptr(page_table),
                                          page_table.shape[1], _ptr(k_pool), _ptr(v_pool), _ptr(out),
                                          batch, n_head, -n_splits if cluster else n_splits, _ptr(scratch),
                                          _stream()))
    return out


def kv_scatter(qkv: torch.Tensor, row_pos: torch.Tensor, row_slot: torch.Tensor, page_table: torch.Tensor,
               k_pool: torch.Tensor, v_pool: torch.Tensor, n_head: int) -> None:
    check(lib.b200llm_op_kv_scatter(_ptr(qkv), _ptr(row_pos), _ptr(row_slot), qkv.shape[0], n_head,
                                    _ptr(page_table), page_table.shape[1], _ptr(k_pool), _ptr(v_pool), _stream()))


def prefill_attention_paged(qkv: torch.Tensor, cu_seqlens: torch.Tensor, prefix_len: torch.Tensor,
                            slots: torch.Tensor, max_suffix_len: int, n_head: int, page_table: torch.Tensor,
                            k_pool: torch.Tensor, v_pool: torch.Tensor, impl: int = 0) -> torch.Tensor:
    """Causal attention of prompt suffixes over the paged cache (prefix + suffix keys in the pages).
    impl 0: the engine's tcgen05 kernel; 1: the mma.sync version of the same computation."""
    out = torch.zeros(qkv.shape[0], n_head * 64, dtype=torch.bfloat16, device=qkv.device)
    check(lib.b200llm_op_prefill_attention_paged(_ptr(qkv), qkv.shape[0], _ptr(cu_seqlens), _ptr(prefix_len),
                                                 _ptr(slots), cu_seqlens.numel() - 1, max_suffix_len, n_head,
                                                 _ptr(page_table), page_table.shape[1], _ptr(k_pool),
                                                 _ptr(v_pool), k_pool.shape[0], _ptr(out), impl, _stream()))
    return out


def lm_head_argmax_groups(vocab: int, tokens: int) -> int:
    return int(lib.b200llm_lm_head_argmax_groups(vocab, tokens))


def lm_head_argmax(act: torch.Tensor, wte: torch.Tensor, eos_id: int, logits: torch.Tensor | None = None,
                   work: tuple | None = None):
    """The decode step's LM head with the fused greedy argmax. Returns (amax_val, amax_idx,
    eos_logit): per-group candidates [groups, tokens] and the EOS logit per token."""
    tokens, k = act.shape
    vocab = wte.shape[0]
    if work is None:
        g = lm_head_argmax_groups(vocab, tokens)
        work = (torch.full((g, tokens), float("-inf"), device=act.device),
                torch.zeros(g, tokens, dtype=torch.int32, device=act.device),
                torch.zeros(tokens, device=act.device))
    val, idx, eos = work
    check(lib.b200llm_op_lm_head_argmax(_ptr(act), _ptr(wte), tokens, vocab, k, eos_id, _ptr(logits), _ptr(val),
                                        _ptr(idx), _ptr(eos), val.shape[1], _stream()))
    return val, idx, eos


def argmax(logits: torch.Tensor, vocab: int | None = None) -> torch.Tensor:
    rows, ld = logits.shape
    out = torch.empty(rows, dtype=torch.int32, device=logits.device)
    check(lib.b200llm_op_argmax(_ptr(logits), rows, ld, vocab or ld, _ptr(out), _stream()))
    return out


def convert_transpose(src: torch.Tensor) -> torch.Tensor:
    rows, cols = src.shape
    dst = torch.empty(cols, rows, dtype=torch.bfloat16, device=src.device)
    check(lib.b200llm_op_convert_transpose(_ptr(src), _ptr(dst), rows, cols, _stream()))
    return dst


def ln_gemm(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None,
            gelu: bool = False, eps: float = 1e-5, split: int = 1) -> torch.Tensor:
    """out = (gelu_new)(LayerNorm(x) @ w.T + bias), LayerNorm folded into the GEMM (decode, <= 64 rows).
    split > 1: returns the fp32 partial planes [split, rows, features] (no bias) instead."""
    rows, d = x.shape
    features = w.shape[0]
    if split > 1:
        planes = torch.zeros(split, rows, features, dtype=torch.float32, device=x.device)
        check(lib.b200llm_op_ln_gemm(_ptr(x), _ptr(gamma), _ptr(beta), eps, _ptr(w), None, None, rows, features, d,
                                     features, 0, split, _ptr(planes), _stream()))
        return planes
    out = torch.zeros(rows, features, dtype=torch.bfloat16, device=x.device)
    check(lib.b200llm_op_ln_gemm(_ptr(x), _ptr(gamma), _ptr(beta), eps, _ptr(w), _ptr(bias), _ptr(out), rows,
                                 features, d, features, int(gelu), 1, None, _stream()))
    return out


def gemm_redadd(act: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None, x: torch.Tensor, rows: int,
                split: int = 0) -> torch.Tensor:
    """x[:rows] += act[:rows] @ w.T + bias, K splits added into x in L2 (decode, <= 64 rows). `act` must
    hold the whole 16 / 32 / 64-row token tile."""
    check(lib.b200llm_op_gemm_redadd(_ptr(act), _ptr(w), _ptr(bias), _ptr(x), rows, w.shape[0], w.shape[1],
                                     split, _stream()))
    return x
