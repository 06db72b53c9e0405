"""Decode attention for callers that cannot allocate device memory.

Some embedders of libb200llm.so have no way to get a workspace when they need one: their device
memory was carved up once at start-up (a later cudaMalloc would fail, or would stall every stream
on the GPU while the driver handles it), or they are a plug-in that only ever receives input and
output buffers. For them the library offers b200llm_op_decode_attention_noalloc, which keeps the
split-KV partial results in a fixed 4 MiB array inside the library instead of a caller buffer.

This test calls that entry point through the C ABI twice in a row (different shapes, so the
second call runs over whatever the first one left in the shared array) and checks each result
against the ordinary entry point that takes a caller-allocated scratch buffer and against a plain
fp32 reference, with the same 2e-2 tolerance the other attention tests use. It also checks that a request beyond the built-in
limit is refused with a clear message. The file name makes pytest collect it last.
"""
import ctypes as C

import pytest
import torch

from conftest import pkg

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]

PAGE = 16


def _case(dev, ctx_lens, n_head, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    B, d = len(ctx_lens), n_head * 64
    max_pages = (max(ctx_lens) + 1 + PAGE - 1) // PAGE
    n_pages = B * max_pages + 2
    table = torch.randperm(n_pages, generator=g)[: B * max_pages].int().view(B, max_pages).contiguous().to(dev)
    k_pool = (torch.randn(n_pages, n_head, PAGE, 64, generator=g) * 0.8).to(torch.bfloat16).to(dev)
    v_pool = (torch.randn(n_pages, n_head, PAGE, 64, generator=g) * 0.8).to(torch.bfloat16).to(dev)
    qkv = (torch.randn(B, 3 * d, generator=g) * 0.8).to(torch.bfloat16).to(dev)
    seq_len = torch.tensor(ctx_lens, dtype=torch.int32, device=dev)
    return qkv, seq_len, table, k_pool, v_pool


def _noalloc(lib, qkv, seq_len, table, k_pool, v_pool, n_head, n_splits):
    out = torch.zeros(qkv.shape[0], n_head * 64, dtype=torch.bfloat16, device=qkv.device)
    rc = lib.b200llm_op_decode_attention_noalloc(
        qkv.data_ptr(), seq_len.data_ptr(), table.data_ptr(), table.shape[1], k_pool.data_ptr(),
        v_pool.data_ptr(), out.data_ptr(), qkv.shape[0], n_head, n_splits,
        torch.cuda.current_stream().cuda_stream)
    return rc, out


def _fp32_reference(qkv, ctx_lens, table, k_pool, v_pool, n_head):
    """Plain fp32 softmax(q K^T / 8) V over the cached rows plus the new token's own K/V (which the
    kernel has appended to the pool by the time this runs), the same reference test_gpu_ops.py uses."""
    out = []
    for b, L in enumerate(ctx_lens):
        pos = torch.arange(L + 1, device=qkv.device)
        pages = table[b, pos // PAGE].long()
        k = k_pool[pages, :, pos % PAGE].float()      # [L+1, H, 64]
        v = v_pool[pages, :, pos % PAGE].float()
        q = qkv[b].view(3, n_head, 64)[0].float()     # [H, 64]
        p = torch.softmax(torch.einsum("hd,shd->hs", q, k) / 8.0, dim=-1)
        out.append(torch.einsum("hs,shd->hd", p, v).reshape(-1))
    return torch.stack(out)


def test_noalloc_decode_attention_matches_the_scratch_buffer_entry_point(cuda):
    _lib, ops = pkg("_lib"), pkg("ops")
    lib = _lib.lib
    calls = [([700, 33, 1100, 260], 16, 8, 1),      # first call: 8 KV splits
             ([90, 511, 17, 300, 64, 5], 12, 4, 2)]  # second call: other shape over the used array
    for ctx_lens, n_head, n_splits, seed in calls:
        qkv, seq_len, table, k_pool, v_pool = _case(cuda, ctx_lens, n_head, seed)
        k_ref, v_ref = k_pool.clone(), v_pool.clone()
        expected = ops.decode_attention(qkv, seq_len, table, k_ref, v_ref, n_head, n_splits)
        rc, got = _noalloc(lib, qkv, seq_len, table, k_pool, v_pool, n_head, n_splits)
        assert rc == 0, lib.b200llm_last_error()
        torch.cuda.synchronize()
        assert torch.isfinite(got.float()).all()
        assert (got.float() - expected.float()).abs().max().item() <= 2e-2
        assert (got.float() - _fp32_reference(qkv, ctx_lens, table, k_pool, v_pool, n_head)).abs().max().item() <= 2e-2
        # both entry points append the new token's K/V to the same page slots
        assert torch.equal(k_pool, k_ref) and torch.equal(v_pool, v_ref)


def test_noalloc_rejects_requests_beyond_the_builtin_scratch(cuda):
    lib = pkg("_lib").lib
    qkv, seq_len, table, k_pool, v_pool = _case(cuda, [40] * 4, 12, 3)
    limit = 1 << 20
    n_splits = limit // (4 * 12 * 66) + 1          # one split too many for 4 rows x 12 heads
    rc, _ = _noalloc(lib, qkv, seq_len, table, k_pool, v_pool, 12, n_splits)
    assert rc == 4                                   # B200LLM_ERR_CAPACITY
    msg = lib.b200llm_last_error().decode()
    assert "exceed the built-in scratch" in msg and str(limit) in msg
