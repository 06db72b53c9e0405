"""Continuous batching + routing logic on a fake engine (host logic only)."""
import threading

import pytest

from conftest import pkg
from fakes import FakeEngine

sched = pkg("scheduler")
SamplingParams = pkg("engine").SamplingParams


def _router(*engines, **kw):
    return sched.Router([sched.EngineWorker(e, **kw) for e in engines])


def test_concurrent_requests_are_batched_and_all_complete():
    eng = FakeEngine(max_seqs=8, max_new_tokens=12)
    r = _router(eng, decode_burst=4)
    futs = [r.submit([i, i + 1, i + 2], SamplingParams(max_new_tokens=10)) for i in range(20)]
    res = [f.result(timeout=10) for f in futs]
    for i, x in enumerate(res):
        base = 3 * i + 3
        assert x.tokens == [(base + s) % 50000 for s in range(10)]
        assert x.finish_reason == 2 and x.prompt_tokens == 3 and x.total_ms >= x.queue_ms >= 0
    assert eng.max_active_seen > 1, "requests were never batched together"
    assert not eng.slots, "KV slots leaked"
    r.shutdown()


def test_a_lone_request_runs_no_more_decode_steps_than_it_needs():
    """32 new tokens = 1 from prefill + 31 decode steps: bursts of 8 + 8 + 8 + 7, not 4 x 8."""
    eng = FakeEngine(max_seqs=4, max_seq_len=128, max_new_tokens=64)
    r = _router(eng, decode_burst=8)
    res = r.generate([1, 2, 3], SamplingParams(max_new_tokens=32), timeout=10)
    assert len(res.tokens) == 32
    steps = [c[2] for c in eng.calls if c[0] == "decode"]
    assert sum(steps) == 31 and max(steps) <= 8, steps
    r.shutdown()


def test_eos_ends_a_request_early_with_finish_reason_stop():
    eng = FakeEngine(max_seqs=4, max_new_tokens=16, eos_every=7)
    r = _router(eng)
    out = r.generate([3], SamplingParams(max_new_tokens=16), timeout=5)   # base 3: step 4 hits 7
    assert out.finish_reason == 1 and out.tokens == [3, 4, 5, 6]
    r.shutdown()


def test_admission_respects_slot_and_token_budgets():
    eng = FakeEngine(max_seqs=2, max_batch_tokens=10, max_new_tokens=4)
    r = _router(eng, decode_burst=2)
    futs = [r.submit([1] * 6, SamplingParams(max_new_tokens=4)) for _ in range(5)]
    assert all(len(f.result(timeout=10).tokens) == 4 for f in futs)
    prefill_sizes = [c[1] for c in eng.calls if c[0] == "prefill"]
    assert all(sum(s) <= 10 for s in prefill_sizes) and max(len(s) for s in prefill_sizes) == 1
    assert eng.max_active_seen <= 2
    r.shutdown()


def test_oversized_prompt_and_engine_failure_become_exceptions_not_hangs():
    eng = FakeEngine(max_seqs=2, max_seq_len=16)
    r = _router(eng)
    with pytest.raises(ValueError):
        r.generate([1] * 16, SamplingParams(max_new_tokens=4), timeout=5)
    r.shutdown()
    bad = FakeEngine(fail_prefill=True)
    r2 = _router(bad)
    with pytest.raises(RuntimeError):
        r2.generate([1, 2], SamplingParams(max_new_tokens=4), timeout=5)
    assert not bad.slots
    r2.shutdown()


def test_router_shards_requests_over_replicas_with_identical_results():
    """Replicas only: the same request set gives the same tokens on 1 or 2 workers."""
    prompts = [[i, 2 * i + 1] for i in range(24)]
    sp = SamplingParams(max_new_tokens=6)
    one = _router(FakeEngine(max_seqs=4))
    ref = [one.generate(p, sp, timeout=10).tokens for p in prompts]
    one.shutdown()
    engines = [FakeEngine(max_seqs=4), FakeEngine(max_seqs=4)]
    two = _router(*engines)
    futs = [two.submit(p, sp) for p in prompts]
    got = [f.result(timeout=10).tokens for f in futs]
    assert got == ref
    used = [sum(1 for c in e.calls if c[0] == "prefill") for e in engines]
    assert all(u > 0 for u in used), f"one replica stayed idle: {used}"
    two.shutdown()


def test_many_threads_submitting_concurrently():
    eng = FakeEngine(max_seqs=8, max_new_tokens=8)
    r = _router(eng)
    results = {}

    def client(i):
        results[i] = r.generate([i], SamplingParams(max_new_tokens=5), timeout=10).tokens

    threads = [threading.Thread(target=client, args=(i,)) for i in range(32)]
    [t.start() for t in threads]
    [t.join() for t in threads]
    assert all(results[i] == [(i + s) % 50000 for s in range(5)] for i in range(32))
    r.shutdown()
