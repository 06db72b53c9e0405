"""Kernel-level parity: each CUDA op, called through the C ABI, against a plain PyTorch fp32
reference of the same op on the same inputs (oracle citations live in include/b200llm.h).

Tolerances: GEMM inputs are bf16 and accumulate in fp32 -> the only error vs the fp32 reference
of the SAME bf16 inputs is summation order (1e-3 relative on O(1) outputs) plus the bf16
rounding of bf16 outputs (2^-8 relative)."""
import math

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


def gelu_new(x):
    return 0.5 * x * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (x + 0.044715 * x.pow(3))))


def _gemm_inputs(tokens, features, k, dev, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    act = (torch.randn(tokens, k, generator=g) * 0.5).to(torch.bfloat16).to(dev)
    w = (torch.randn(features, k, generator=g) * 0.05).to(torch.bfloat16).to(dev)
    bias = (torch.randn(features, generator=g) * 0.1).to(dev)
    return act, w, bias


@pytest.mark.parametrize("tokens,features,k,orientation", [
    (64, 2304, 768, 1),     # decode QKV, swap-AB
    (1, 768, 768, 1),       # single token
    (13, 3072, 768, 1),     # ragged batch -> 16-wide N tile
    (40, 1024, 1024, 1),    # 64-wide tile with padding rows
    (100, 1280, 5120, 1),   # large-model MLP down-proj
    (256, 768, 3072, 1),
    (300, 768, 768, 1),     # > 256 tokens in swap-AB: two N tiles
    (128, 2304, 768, 0),    # prefill orientation, single M tile
    (1000, 3072, 768, 0),   # ragged M, 128-wide N
    (8192, 2304, 768, 0),   # BASELINE config 2 prefill QKV -> 256-wide N tiles, persistent
    (333, 768, 3072, -1),   # auto
])
def test_gemm_bias_bf16(ops, cuda, tokens, features, k, orientation):
    act, w, bias = _gemm_inputs(tokens, features, k, cuda)
    out = torch.zeros(tokens, features, dtype=torch.bfloat16, device=cuda)
    ops.gemm(act, w, bias, out, ops.EPI_BIAS_BF16, orientation=orientation)
    torch.cuda.synchronize()
    ref = act.float() @ w.float().t() + bias
    err = (out.float() - ref).abs().max().item()
    scale = ref.abs().max().item()
    assert err <= 2e-3 * scale + 2 ** -8 * scale, (err, scale)


@pytest.mark.parametrize("tokens,features,k,epi,split", [
    (8192, 2304, 768, "bias", 1),    # BASELINE config 2 prefill QKV: 288 pair tiles, persistent clusters
    (512, 3072, 768, "gelu", 1),
    (640, 768, 3072, "resid", 1),    # TMA reduce-add into the fp32 residual stream
    (300, 768, 768, "resid", 1),     # ragged: the second CTA of the last pair has only out-of-range rows
    (129, 1024, 1024, "bias", 1),
    (8192, 768, 3072, "resid", 3),   # config 2 MLP down-projection: 96 tiles x 3 K splits = 288 work items
    (640, 768, 3072, "resid", 4),
    (300, 1024, 4096, "resid", 2),   # ragged rows with K splits (the bias is added by the first split only)
    (1000, 768, 3072, "resid", 5),   # a split count that does not divide the 48 k-blocks (10, 10, 10, 10, 8)
    (4096, 768, 3072, "resid", -1),  # the split count by the wave model (what the engine passes)
])
def test_gemm_cta_pair_umma(ops, cuda, tokens, features, k, epi, split):
    """tcgen05.mma.cta_group::2: two CTAs share one 256 x 256 tile (orientation code 2). `split` > 1: the K
    range of a tile is cut into work items whose partial products meet in the TMA reduce-add."""
    act, w, bias = _gemm_inputs(tokens, features, k, cuda, seed=8)
    ref = act.float() @ w.float().t() + bias
    if epi == "resid":
        resid = torch.randn(tokens, features, device=cuda)
        out = resid.clone()
        ops.gemm(act, w, bias, out, ops.EPI_BIAS_RESID_F32, split_k=split, orientation=2)
        ref, got = ref + resid, out
    else:
        out = torch.zeros(tokens, features, dtype=torch.bfloat16, device=cuda)
        ops.gemm(act, w, bias, out, ops.EPI_BIAS_GELU_BF16 if epi == "gelu" else ops.EPI_BIAS_BF16, orientation=2)
        if epi == "gelu":
            ref = gelu_new(ref)
        got = out.float()
    torch.cuda.synchronize()
    scale = max(1.0, ref.abs().max().item())
    assert (got - ref).abs().max().item() <= 1e-2 * scale


@pytest.mark.parametrize("orientation,tokens", [(1, 64), (0, 512)])
def test_gemm_gelu(ops, cuda, orientation, tokens):
    act, w, bias = _gemm_inputs(tokens, 3072, 768, cuda, seed=1)
    out = torch.zeros(tokens, 3072, dtype=torch.bfloat16, device=cuda)
    ops.gemm(act, w, bias, out, ops.EPI_BIAS_GELU_BF16, orientation=orientation)
    torch.cuda.synchronize()
    ref = gelu_new(act.float() @ w.float().t() + bias)
    assert (out.float() - ref).abs().max().item() <= 1e-2 * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("orientation,tokens", [(1, 64), (1, 7), (0, 640)])
def test_gemm_residual_f32(ops, cuda, orientation, tokens):
    act, w, bias = _gemm_inputs(tokens, 768, 3072, cuda, seed=2)
    resid = torch.randn(tokens, 768, device=cuda)
    out = resid.clone()
    ops.gemm(act, w, bias, out, ops.EPI_BIAS_RESID_F32, orientation=orientation)
    torch.cuda.synchronize()
    ref = resid + act.float() @ w.float().t() + bias
    assert (out - ref).abs().max().item() <= 2e-3 * ref.abs().max().item()


@pytest.mark.parametrize("orientation,tokens", [(1, 64), (1, 3), (0, 260)])
def test_gemm_lm_head_f32_odd_vocab(ops, cuda, orientation, tokens):
    vocab = 50257  # not a multiple of anything: masked tail tile + unaligned rows
    act, w, _ = _gemm_inputs(tokens, vocab, 768, cuda, seed=3)
    out = torch.full((tokens, vocab), float("nan"), device=cuda)
    ops.gemm(act, w, None, out, ops.EPI_F32, orientation=orientation)
    torch.cuda.synchronize()
    ref = act.float() @ w.float().t()
    assert torch.isfinite(out).all()
    assert (out - ref).abs().max().item() <= 2e-3 * ref.abs().max().item()


@pytest.mark.parametrize("split", [2, 6, 12, 24])
def test_gemm_split_k_partials_then_layernorm(ops, cuda, split):
    tokens, d, k = 64, 768, 3072
    if (k // 64) % split:
        pytest.skip("split must divide k blocks")
    act, w, bias = _gemm_inputs(tokens, d, k, cuda, seed=4)
    partials = torch.zeros(split, tokens, d, device=cuda)
    ops.gemm(act, w, None, partials, ops.EPI_PARTIAL_F32, split_k=split, orientation=1)
    x = torch.randn(tokens, d, device=cuda)
    x0 = x.clone()
    gamma = torch.rand(d, device=cuda) + 0.5
    beta = torch.randn(d, device=cuda) * 0.1
    h = ops.layernorm(x, gamma, beta, 1e-5, partials=partials, add_bias=bias)
    torch.cuda.synchronize()
    ref_x = x0 + act.float() @ w.float().t() + bias
    assert (x - ref_x).abs().max().item() <= 2e-3 * ref_x.abs().max().item()
    ref_h = torch.nn.functional.layer_norm(ref_x, (d,), gamma, beta, 1e-5)
    assert (h.float() - ref_h).abs().max().item() <= 3e-2


@pytest.mark.parametrize("tokens,features,k,cluster,epi", [
    (64, 2304, 768, 6, "bias"),      # decode QKV: 18 tiles x cluster of 6
    (64, 3072, 768, 6, "gelu"),      # MLP up
    (64, 768, 768, 6, "resid"),      # attention out-proj, uneven token split (64 / 6)
    (64, 768, 3072, 8, "resid"),     # MLP down: 6 k-blocks per CTA
    (13, 1024, 1024, 4, "bias"),     # ragged batch on a 16-wide tile
    (200, 1280, 5120, 8, "resid"),   # 256-wide tile, 10 k-blocks per CTA (> ring depth)
    (1, 768, 768, 2, "gelu"),
])
def test_gemm_cluster_split_k_dsmem_reduce(ops, cuda, tokens, features, k, cluster, epi):
    """K split over a thread-block cluster, reduced through distributed shared memory."""
    act, w, bias = _gemm_inputs(tokens, features, k, cuda, seed=6)
    ref = act.float() @ w.float().t() + bias
    if epi == "resid":
        resid = torch.randn(tokens, features, device=cuda)
        out = resid.clone()
        ops.gemm(act, w, bias, out, ops.EPI_BIAS_RESID_F32, split_k=-cluster, orientation=1)
        ref = ref + resid
        got = out
    else:
        out = torch.zeros(tokens, features, dtype=torch.bfloat16, device=cuda)
        code = ops.EPI_BIAS_GELU_BF16 if epi == "gelu" else ops.EPI_BIAS_BF16
        ops.gemm(act, w, bias, out, code, split_k=-cluster, orientation=1)
        if epi == "gelu":
            ref = gelu_new(ref)
        got = out.float()
    torch.cuda.synchronize()
    scale = max(1.0, ref.abs().max().item())
    assert (got - ref).abs().max().item() <= 1e-2 * scale
    # deterministic: a second run is bit-identical
    if epi != "resid":
        out2 = torch.zeros_like(out)
        ops.gemm(act, w, bias, out2, code, split_k=-cluster, orientation=1)
        torch.cuda.synchronize()
        assert torch.equal(out, out2)


@pytest.mark.parametrize("rows,d", [(1, 768), (64, 768), (257, 1024), (33, 1280), (5, 128)])
def test_layernorm(ops, cuda, rows, d):
    x = torch.randn(rows, d, device=cuda) * 3 + 0.7
    gamma = torch.rand(d, device=cuda) + 0.5
    beta = torch.randn(d, device=cuda)
    out = ops.layernorm(x, gamma, beta, 1e-5)
    torch.cuda.synchronize()
    ref = torch.nn.functional.layer_norm(x, (d,), gamma, beta, 1e-5)
    # bf16 output rounding: 2^-8 relative
    assert (out.float() - ref).abs().max().item() <= 2 ** -7 * ref.abs().max().item()


def test_embed(ops, cuda):
    vocab, npos, d, n = 1000, 64, 768, 77
    wte = torch.randn(vocab, d, device=cuda)
    wpe = torch.randn(npos, d, device=cuda)
    ids = torch.randint(0, vocab, (n,), dtype=torch.int32, device=cuda)
    pos = torch.randint(0, npos, (n,), dtype=torch.int32, device=cuda)
    x = ops.embed(wte, wpe, ids, pos)
    torch.cuda.synchronize()
    assert torch.equal(x, wte[ids.long()] + wpe[pos.long()])


def test_convert_transpose(ops, cuda):
    src = torch.randn(768, 2304, device=cuda)
    dst = ops.convert_transpose(src)
    torch.cuda.synchronize()
    assert torch.equal(dst, src.t().contiguous().to(torch.bfloat16))


def test_argmax_ties_and_tail(ops, cuda):
    logits = torch.randn(9, 50257, device=cuda)
    logits[0, 50256] = 100.0          # last element
    logits[1, 0] = 100.0              # first element
    logits[2, 77] = 50.0
    logits[2, 99] = 50.0              # tie -> lowest index
    out = ops.argmax(logits)
    torch.cuda.synchronize()
    ref = logits.argmax(-1).int()
    assert out[0] == 50256 and out[1] == 0 and out[2] == 77
    assert torch.equal(out[3:], ref[3:])


def _ref_attention(q, k, v):
    # q [T, H, 64], k/v [S, H, 64], causal with q occupying the last T positions of S
    T, S = q.shape[0], k.shape[0]
    s = torch.einsum("thd,shd->hts", q.float(), k.float()) / 8.0
    mask = torch.ones(T, S, dtype=torch.bool, device=q.device).tril(S - T)
    s = s.masked_fill(~mask, float("-inf"))
    p = torch.softmax(s, dim=-1)
    return torch.einsum("hts,shd->thd", p, v.float())


@pytest.mark.parametrize("lens,n_head", [
    ([128] * 4, 12),            # BASELINE config 2 shape (subset of batch)
    ([1, 5, 64, 65, 200], 12),  # ragged, tile boundaries
    ([1024], 16),               # long prompt (config 3)
    ([37, 300], 20),
])
def test_prefill_attention(ops, cuda, lens, n_head):
    d = n_head * 64
    n_tok = sum(lens)
    qkv = (torch.randn(n_tok, 3 * d, device=cuda) * 0.8).to(torch.bfloat16)
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device=cuda)
    out = ops.prefill_attention(qkv, cu, max(lens), n_head)
    torch.cuda.synchronize()
    start = 0
    for L in lens:
        blk = qkv[start:start + L].view(L, 3, n_head, 64)
        ref = _ref_attention(blk[:, 0], blk[:, 1], blk[:, 2]).reshape(L, d)
        got = out[start:start + L].float()
        assert (got - ref).abs().max().item() <= 2e-2, (L, (got - ref).abs().max().item())
        start += L


@pytest.mark.parametrize("ctx_lens,n_head,n_splits", [
    ([0, 1, 15, 16, 17, 159, 160], 12, 1),
    ([127] * 8, 12, 1),
    # several 128-row staging passes in one CTA, incl. the register-held head of the 2nd pass
    ([126, 127, 128, 129, 143, 144, 160, 161, 255, 256, 257, 511, 640], 12, 1),
    ([62, 63, 64, 65, 191, 192, 193], 12, 1),
    ([1100, 3, 64, 700], 16, 4),
    ([1100, 3, 64, 700], 16, 8),
    ([40, 500], 12, 3),
    ([1100, 3, 64, 700], 16, 16),
    ([300, 299], 20, 2),
    ([3000, 2047, 1], 12, 8),      # several page-list refills per split (> 512 tokens each)
    ([4100, 4095], 12, 16),        # beyond the 256 page ids one CTA stages in shared memory
])
@pytest.mark.parametrize("mode", ["scratch-combine", "cluster-combine", "streaming", "streaming-merge"])
def test_decode_attention_paged(ops, cuda, ctx_lens, n_head, n_splits, mode):
    """seq_len[b] tokens are cached; the new token (position seq_len[b]) is appended by the
    kernel, then attends over everything including itself. `cluster` = 1 makes 2..8 KV splits merge
    inside the kernel (the engine's setting) instead of through scratch + a combine launch; `streaming`
    is the persistent-warp / TMA form of the kernel the engine runs for large batches and long contexts
    (`streaming-merge`: its splits merged inside the kernel through zeroed counters, as the engine does)."""
    cluster = mode == "cluster-combine"
    if cluster and not 2 <= n_splits <= 8:
        pytest.skip("the in-cluster combine only exists for 2..8 splits")
    d = n_head * 64
    B = len(ctx_lens)
    page = 16
    max_pages = (max(ctx_lens) + 1 + page - 1) // page
    n_pages = B * max_pages + 3
    g = torch.Generator(device="cpu").manual_seed(5)
    perm = torch.randperm(n_pages, generator=g)[: B * max_pages].int()
    page_table = perm.view(B, max_pages).contiguous().to(cuda)  # scattered physical pages
    k_pool = torch.zeros(n_pages, n_head, page, 64, dtype=torch.bfloat16, device=cuda)
    v_pool = torch.zeros_like(k_pool)
    ks, vs = [], []
    for b, L in enumerate(ctx_lens):
        k = (torch.randn(L, n_head, 64, generator=g) * 0.8).to(torch.bfloat16).to(cuda)
        v = (torch.randn(L, n_head, 64, generator=g) * 0.8).to(torch.bfloat16).to(cuda)
        ks.append(k)
        vs.append(v)
        for t in range(L):
            p = page_table[b, t // page].item()
            k_pool[p, :, t % page] = k[t]
            v_pool[p, :, t % page] = v[t]
    qkv = (torch.randn(B, 3 * d, generator=g) * 0.8).to(torch.bfloat16).to(cuda)
    seq_len = torch.tensor(ctx_lens, dtype=torch.int32, device=cuda)
    counters = torch.zeros(B * n_head, dtype=torch.int32, device=cuda) if mode == "streaming-merge" else None
    out = ops.decode_attention(qkv, seq_len, page_table, k_pool, v_pool, n_head, n_splits, cluster=cluster,
                               streaming=mode.startswith("streaming"), counters=counters)
    torch.cuda.synchronize()
    if counters is not None:
        assert int(counters.abs().sum()) == 0  # left at zero for the next launch ...
        again = ops.decode_attention(qkv, seq_len, page_table, k_pool, v_pool, n_head, n_splits, streaming=True,
                                     counters=counters)
        torch.cuda.synchronize()
        assert torch.equal(out, again)         # ... whose merge runs in split order whoever finishes last
    for b, L in enumerate(ctx_lens):
        cur = qkv[b].view(3, n_head, 64)
        k_all = torch.cat([ks[b], cur[1][None]], 0)
        v_all = torch.cat([vs[b], cur[2][None]], 0)
        ref = _ref_attention(cur[0][None], k_all, v_all).reshape(d)
        got = out[b].float()
        assert (got - ref).abs().max().item() <= 2e-2, (b, L)
        # the new token's K/V landed in its page
        p = page_table[b, L // page].item()
        assert torch.equal(k_pool[p, :, L % page], cur[1])
        assert torch.equal(v_pool[p, :, L % page], cur[2])


@pytest.mark.parametrize("seed", range(12))
def test_decode_attention_kernels_agree_on_random_shapes(ops, cuda, seed):
    """The two decode attention kernels (one CTA per unit / persistent warps with TMA tiles) on random row
    counts, head counts, ragged contexts and split counts: each against the fp32 reference, and the K/V
    append of both. Contexts are drawn around the tile, page and pass boundaries as well as uniformly."""
    g = torch.Generator(device="cpu").manual_seed(1000 + seed)
    n_head = [12, 16, 20][seed % 3]
    B = int(torch.randint(1, 41, (1,), generator=g))
    edges = torch.tensor([0, 1, 15, 16, 17, 31, 32, 33, 127, 128, 129, 159, 160, 161, 255, 256, 257, 511, 512, 513])
    ctx = [int(edges[int(torch.randint(0, len(edges), (1,), generator=g))]) if int(torch.randint(0, 2, (1,), generator=g))
           else int(torch.randint(0, 700, (1,), generator=g)) for _ in range(B)]
    n_splits = [1, 1, 2, 3, 5, 8][int(torch.randint(0, 6, (1,), generator=g))]
    d, page = n_head * 64, 16
    max_pages = (max(ctx) + 1 + page - 1) // page
    n_pages = B * max_pages + 7
    page_table = torch.randperm(n_pages, generator=g)[: B * max_pages].int().view(B, max_pages).contiguous().to(cuda)
    k0 = torch.zeros(n_pages, n_head, page, 64, dtype=torch.bfloat16, device=cuda)
    v0 = torch.zeros_like(k0)
    ks, vs = [], []
    for b, L in enumerate(ctx):
        k = (torch.randn(L, n_head, 64, generator=g) * 0.8).to(torch.bfloat16).to(cuda)
        v = (torch.randn(L, n_head, 64, generator=g) * 0.8).to(torch.bfloat16).to(cuda)
        ks.append(k)
        vs.append(v)
        if L:
            pos = torch.arange(L, device=cuda)
            pg = page_table[b, (pos // page).long()].long()
            k0[pg, :, (pos % page).long()] = k
            v0[pg, :, (pos % page).long()] = v
    qkv = (torch.randn(B, 3 * d, generator=g) * 0.8).to(torch.bfloat16).to(cuda)
    seq_len = torch.tensor(ctx, dtype=torch.int32, device=cuda)
    outs = []
    for streaming in (False, True):
        kp, vp = k0.clone(), v0.clone()
        out = ops.decode_attention(qkv, seq_len, page_table, kp, vp, n_head, n_splits, streaming=streaming)
        torch.cuda.synchronize()
        outs.append(out)
        for b, L in enumerate(ctx):
            cur = qkv[b].view(3, n_head, 64)
            ref = _ref_attention(cur[0][None], torch.cat([ks[b], cur[1][None]], 0),
                                 torch.cat([vs[b], cur[2][None]], 0)).reshape(d)
            assert (out[b].float() - ref).abs().max().item() <= 2e-2, (streaming, b, L, n_splits)
            p = page_table[b, L // page].item()
            assert torch.equal(kp[p, :, L % page], cur[1]) and torch.equal(vp[p, :, L % page], cur[2])
    assert (outs[0].float() - outs[1].float()).abs().max().item() <= 2e-2


@pytest.mark.parametrize("prefix_suffix,n_head", [
    ([(0, 40), (64, 1), (16, 100), (496, 130)], 12),    # no prefix / one token / ragged tiles / long prefix
    ([(1008, 16), (32, 33)], 16),                       # Summarize-like: almost everything cached
    ([(48, 64), (64, 64), (80, 65), (0, 200)], 20),     # prefix not a multiple of the key tile
    ([(0, 128)] * 4, 12),                               # BASELINE config 2 prompts (cold)
    ([(0, 1024), (0, 127), (0, 129), (0, 257)], 16),    # config 3 prompt: 8 query tiles x up to 8 key tiles
    ([(1, 1), (15, 2), (127, 1), (128, 1), (129, 130)], 12),
])
@pytest.mark.parametrize("impl", [0, 1], ids=["tcgen05", "mma.sync"])
def test_prefill_attention_over_a_cached_prefix(ops, cuda, prefix_suffix, n_head, impl):
    """The engine's prefill attention. Sequence b has `prefix` tokens in its pages already (0 for a
    cold prompt); its `suffix` rows are scattered to the pages by kv_scatter and then attend over
    prefix + suffix keys read from the pages. Reference: fp32 causal attention of the suffix queries
    over all keys. impl 0 is the tcgen05 / TMEM / TMA kernel the engine runs, impl 1 the same
    computation with mma.sync."""
    d, page = n_head * 64, 16
    B = len(prefix_suffix)
    totals = [p + s for p, s in prefix_suffix]
    max_pages = (max(totals) + page - 1) // page
    n_pages = B * max_pages + 5
    g = torch.Generator(device="cpu").manual_seed(11)
    page_table = torch.randperm(n_pages, generator=g)[: B * max_pages].int().view(B, max_pages).contiguous().to(cuda)
    k_pool = torch.zeros(n_pages, n_head, page, 64, dtype=torch.bfloat16, device=cuda)
    v_pool = torch.zeros_like(k_pool)
    n_rows = sum(s for _, s in prefix_suffix)
    qkv = (torch.randn(n_rows, 3 * d, generator=g) * 0.8).to(torch.bfloat16).to(cuda)
    ks, vs, row_pos, row_slot, cu = [], [], [], [], [0]
    for b, (P, S) in enumerate(prefix_suffix):
        k = (torch.randn(P, n_head, 64, generator=g) * 0.8).to(torch.bfloat16).to(cuda)
        v = (torch.randn(P, n_head, 64, generator=g) * 0.8).to(torch.bfloat16).to(cuda)
        for t in range(P):
            pg = page_table[b, t // page].item()
            k_pool[pg, :, t % page] = k[t]
            v_pool[pg, :, t % page] = v[t]
        ks.append(k)
        vs.append(v)
        row_pos += list(range(P, P + S))
        row_slot += [b] * S
        cu.append(cu[-1] + S)
    row_pos = torch.tensor(row_pos, dtype=torch.int32, device=cuda)
    row_slot = torch.tensor(row_slot, dtype=torch.int32, device=cuda)
    cu_t = torch.tensor(cu, dtype=torch.int32, device=cuda)
    prefix = torch.tensor([p for p, _ in prefix_suffix], dtype=torch.int32, device=cuda)
    slots = torch.arange(B, dtype=torch.int32, device=cuda)
    k_before, v_before = k_pool.clone(), v_pool.clone()
    if impl == 1:   # the mma.sync partner reads every key from the pages; the tcgen05 kernel fills them itself
        ops.kv_scatter(qkv, row_pos, row_slot, page_table, k_pool, v_pool, n_head)
    out = ops.prefill_attention_paged(qkv, cu_t, prefix, slots, max(s for _, s in prefix_suffix), n_head,
                                      page_table, k_pool, v_pool, impl=impl)
    torch.cuda.synchronize()
    for b, (P, S) in enumerate(prefix_suffix):
        blk = qkv[cu[b]:cu[b + 1]].view(S, 3, n_head, 64)
        k_all = torch.cat([ks[b], blk[:, 1]], 0)
        v_all = torch.cat([vs[b], blk[:, 2]], 0)
        ref = _ref_attention(blk[:, 0], k_all, v_all).reshape(S, d)
        got = out[cu[b]:cu[b + 1]].float()
        assert (got - ref).abs().max().item() <= 2e-2, (b, P, S, (got - ref).abs().max().item())
        # the scatter wrote exactly the suffix rows, the cached prefix rows are untouched
        for t in (0, S // 2, S - 1):
            pg = page_table[b, (P + t) // page].item()
            assert torch.equal(k_pool[pg, :, (P + t) % page], blk[t, 1])
            assert torch.equal(v_pool[pg, :, (P + t) % page], blk[t, 2])
        for t in range(0, P, 7):
            pg = page_table[b, t // page].item()
            assert torch.equal(k_pool[pg, :, t % page], k_before[pg, :, t % page])
            assert torch.equal(v_pool[pg, :, t % page], v_before[pg, :, t % page])


@pytest.mark.parametrize("tokens,vocab,k", [(64, 50257, 768), (4, 50257, 768), (32, 50257, 1280), (1, 1000, 128),
                                            (200, 12345, 256)])
def test_lm_head_with_fused_argmax(ops, cuda, tokens, vocab, k):
    """The kernel instantiation the decode step launches (kEpiLogitsArgmax, balanced row ranges for
    one token tile): candidates reduce to torch's argmax of the fp32 logits of the same bf16 inputs,
    ties to the lowest index, EOS kept apart; optional logits equal the plain LM-head GEMM."""
    g = torch.Generator(device="cpu").manual_seed(tokens + vocab)
    act = (torch.randn(tokens, k, generator=g) * 0.5).to(torch.bfloat16).to(cuda)
    wte = (torch.randn(vocab, k, generator=g) * 0.05).to(torch.bfloat16).to(cuda)
    eos = vocab - 1
    wte[eos] = act[0] * 0.2                      # EOS would win row 0 if it were not kept apart
    wte[7] = wte[vocab // 2]                     # an exact tie across two groups -> lowest index wins
    logits = torch.zeros(tokens, vocab, device=cuda)
    for store in (False, True):
        val, idx, eos_logit = ops.lm_head_argmax(act, wte, eos, logits if store else None)
        torch.cuda.synchronize()
        ref = act.float() @ wte.float().t()
        assert (eos_logit - ref[:, eos]).abs().max().item() <= 2e-3 * ref.abs().max().item()
        masked = ref.clone()
        masked[:, eos] = float("-inf")
        want = masked.argmax(-1)
        best = val.max(0)
        # lowest index among the groups that reach the maximum
        cand = torch.where(val == best.values[None, :], idx, torch.full_like(idx, vocab))
        got = cand.min(0).values.long()
        margin = masked.topk(2, -1).values
        decisive = (margin[:, 0] - margin[:, 1]) > 1e-3 * ref.abs().max().item()
        assert (got[decisive] == want[decisive]).all()
        assert (got != eos).all()
        if store:
            assert (logits - ref).abs().max().item() <= 2e-3 * ref.abs().max().item()


@pytest.mark.parametrize("rows,d,features,gelu", [
    (64, 768, 2304, False),    # GPT-2 small ln_1 + c_attn at the headline batch
    (64, 768, 3072, True),     # ln_2 + c_fc + gelu_new
    (16, 1024, 3072, False),   # medium: the weight ring recycles (16 k-blocks, 12 stages)
    (32, 1280, 5120, True),    # large
    (1, 768, 2304, False), (13, 128, 384, True), (50, 256, 1024, True),
])
def test_layernorm_folded_into_the_gemm(ops, cuda, rows, d, features, gelu):
    """ln_gemm_kernel: out = (gelu_new)(LayerNorm(x) W^T + b) with the LayerNorm computed inside the
    GEMM kernel (fp32 statistics) and handed to the tensor core through shared memory. Reference:
    fp32 layer_norm -> bf16 (the operand precision) -> fp32 matmul."""
    g = torch.Generator(device="cpu").manual_seed(rows + d)
    x = (torch.randn(rows, d, generator=g) * 3.0 + 1.5).to(cuda)     # non-zero mean, std 3
    gamma = (1.0 + 0.1 * torch.randn(d, generator=g)).to(cuda)
    beta = (0.1 * torch.randn(d, generator=g)).to(cuda)
    w = (torch.randn(features, d, generator=g) * 0.05).to(torch.bfloat16).to(cuda)
    bias = (torch.randn(features, generator=g) * 0.1).to(cuda)
    out = ops.ln_gemm(x, gamma, beta, w, bias, gelu=gelu)
    torch.cuda.synchronize()
    h = torch.nn.functional.layer_norm(x, (d,), gamma, beta, 1e-5).to(torch.bfloat16).float()
    ref = h @ w.float().t() + bias
    if gelu:
        ref = gelu_new(ref)
    err = (out.float() - ref).abs().max().item()
    scale = ref.abs().max().item()
    # bf16 output rounding + an occasional 1-ulp difference of a normalised operand element
    assert err <= 4e-3 * scale + 2 ** -8 * scale, (err, scale)
    # the K-split form (light CTAs, fp32 partial planes for the consumer to reduce): same sum, no bias
    if not gelu:
        n_kb = d // 64
        for split in [s for s in (2, 4) if n_kb % s == 0]:
            planes = ops.ln_gemm(x, gamma, beta, w, None, split=split)
            torch.cuda.synchronize()
            got = planes.sum(0) + bias
            err = (got - (h @ w.float().t() + bias)).abs().max().item()
            assert err <= 4e-3 * scale, (split, err, scale)


@pytest.mark.parametrize("rows,features,k,split", [
    (64, 768, 768, 0), (64, 768, 3072, 0),      # GPT-2 small out-proj / MLP-down, automatic split
    (64, 768, 3072, 6), (64, 768, 768, 2),
    (16, 1024, 4096, 0), (32, 1280, 5120, 0),   # medium / large
    (3, 768, 768, 12), (40, 256, 1024, 4), (64, 128, 512, 8),
])
def test_split_k_gemm_reduce_added_into_the_residual_stream(ops, cuda, rows, features, k, split):
    """gemm_redadd_kernel: x += act W^T + b with the K splits added into x in L2 by TMA reduce-add.
    Rows of x beyond the live batch must stay untouched."""
    g = torch.Generator(device="cpu").manual_seed(features + k)
    tile = 16 if rows <= 16 else 32 if rows <= 32 else 64
    act = (torch.randn(tile, k, generator=g) * 0.5).to(torch.bfloat16).to(cuda)
    w = (torch.randn(features, k, generator=g) * 0.05).to(torch.bfloat16).to(cuda)
    bias = (torch.randn(features, generator=g) * 0.1).to(cuda)
    x0 = torch.randn(tile, features, generator=g).to(cuda)
    x = x0.clone()
    ops.gemm_redadd(act, w, bias, x, rows, split)
    torch.cuda.synchronize()
    ref = x0[:rows] + act[:rows].float() @ w.float().t() + bias
    err = (x[:rows] - ref).abs().max().item()
    assert err <= 2e-3 * ref.abs().max().item(), err
    assert torch.equal(x[rows:], x0[rows:]), "rows beyond the batch were modified"
    # twice more: the additions accumulate
    ops.gemm_redadd(act, w, None, x, rows, split)
    torch.cuda.synchronize()
    ref2 = ref + act[:rows].float() @ w.float().t()
    assert (x[:rows] - ref2).abs().max().item() <= 3e-3 * ref2.abs().max().item()
