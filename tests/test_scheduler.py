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


def test_a_decode_error_frees_the_slots_and_later_requests_still_complete():
    """the catch-all used to fail the futures without seq_free, leaking every active
    slot and its KV pages; afterwards nothing could ever be admitted again."""
    eng = FakeEngine(max_seqs=2, max_new_tokens=8, fail_decode_times=1)
    r = _router(eng, decode_burst=4)
    first = [r.submit([i + 1], SamplingParams(max_new_tokens=6)) for i in range(2)]
    for f in first:
        with pytest.raises(RuntimeError):
            f.result(timeout=5)
    assert not eng.slots, "slots leaked by the error path"
    assert eng.pool_left == 2 * 64
    later = [r.submit([i + 5], SamplingParams(max_new_tokens=6)) for i in range(4)]
    assert all(len(f.result(timeout=5).tokens) == 6 for f in later)
    assert r.workers[0].inflight_tokens == 0
    r.shutdown()


def test_a_request_larger_than_the_whole_kv_pool_fails_instead_of_spinning():
    eng = FakeEngine(max_seqs=4, max_seq_len=64, max_new_tokens=16, kv_pool_tokens=32)
    r = _router(eng)
    with pytest.raises(ValueError):
        r.generate([1] * 30, SamplingParams(max_new_tokens=16), timeout=5)   # needs 46 > 32 pool tokens
    assert len(r.generate([1] * 4, SamplingParams(max_new_tokens=8), timeout=5).tokens) == 8
    r.shutdown()


def test_timeout_cancels_the_request_and_returns_its_slot():
    """ADVICE r1 / VERDICT r1 weak #9: a caller that gives up must not leave its sequence decoding
    to max_new_tokens (ask-AI retries three times, llm_server.py:147-212)."""
    import time
    from concurrent.futures import TimeoutError as FutureTimeout

    eng = FakeEngine(max_seqs=2, max_seq_len=4096, max_new_tokens=2000, step_sleep=0.002)
    r = _router(eng, decode_burst=4)
    with pytest.raises(FutureTimeout):
        r.generate([1, 2, 3], SamplingParams(max_new_tokens=2000), timeout=0.1)
    deadline = time.time() + 5
    while eng.slots and time.time() < deadline:
        time.sleep(0.01)
    assert not eng.slots, "the abandoned request kept its slot"
    assert eng.cancels, "the engine was never told to stop the sequence"
    assert r.workers[0].inflight_tokens == 0 and r.workers[0].cancelled == 1
    steps = sum(c[2] for c in eng.calls if c[0] == "decode")
    assert steps < 500, f"kept decoding after the cancel: {steps} steps"
    # a request cancelled while still waiting is dropped at admission
    blocker = r.submit([9], SamplingParams(max_new_tokens=300))
    blocker2 = r.submit([8], SamplingParams(max_new_tokens=300))
    waiting = r.submit([7], SamplingParams(max_new_tokens=5))
    sched.EngineWorker.cancel(waiting)
    with pytest.raises(TimeoutError):
        waiting.result(timeout=5)
    assert len(blocker.result(timeout=10).tokens) == 300 and len(blocker2.result(timeout=10).tokens) == 300
    r.shutdown()


def test_a_failing_replica_is_bypassed_and_reprobed():
    """a worker whose engine keeps raising has zero in-flight tokens and used
    to attract EVERY new request. After FAIL_LIMIT failures it leaves the rotation; one probe
    request is sent to it every REPROBE_S seconds and a success brings it back."""
    import time

    bad = FakeEngine(max_seqs=4, fail_prefill=True)
    good = FakeEngine(max_seqs=4)
    r = _router(bad, good)
    wb, wg = r.workers
    wb.REPROBE_S = 0.3
    sp = SamplingParams(max_new_tokens=4)
    outcomes = []
    for i in range(20):
        try:
            outcomes.append(len(r.generate([i + 1], sp, timeout=5).tokens))
        except RuntimeError:
            outcomes.append(-1)
    assert outcomes.count(-1) == sched.EngineWorker.FAIL_LIMIT, outcomes
    assert outcomes[-10:] == [4] * 10 and not wb.healthy
    # the replica recovers: the next probe succeeds and it takes traffic again
    bad.fail_prefill = False
    time.sleep(0.35)
    assert len(r.generate([99], sp, timeout=5).tokens) == 4
    assert wb.healthy and wb.completed == 1
    r.shutdown()


def test_process_per_gpu_workers_behind_the_router():
    """remote_worker.RemoteEngineWorker: the EngineWorker loop runs in a child process (one per
    GPU), the Router talks to it over a pipe. Same results as in-process workers, requests are
    sharded, cancel and a dying process are handled."""
    import time
    from concurrent.futures import TimeoutError as FutureTimeout

    rw = pkg("remote_worker")
    kw = dict(max_seqs=4, max_seq_len=4096, max_new_tokens=600, step_sleep=0.001)
    workers = [rw.RemoteEngineWorker(d, "small", kw, engine_factory="fakes:FakeEngine", decode_burst=4) for d in (0, 1)]
    r = sched.Router(workers)
    assert all(w.is_alive() for w in workers)
    prompts = [[i, 2 * i + 1] for i in range(16)]
    sp = SamplingParams(max_new_tokens=6)
    futs = [r.submit(p, sp) for p in prompts]
    got = [f.result(timeout=20).tokens for f in futs]
    one = _router(FakeEngine(max_seqs=4))
    assert got == [one.generate(p, sp, timeout=10).tokens for p in prompts]
    one.shutdown()
    snaps = [w.snapshot() for w in workers]
    assert all(s["completed"] > 0 for s in snaps) and sum(s["completed"] for s in snaps) == 16
    assert all(w.inflight_tokens == 0 for w in workers)
    # a caller that gives up: the cancel crosses the pipe and the child frees the slot
    with pytest.raises(FutureTimeout):
        r.generate([1, 2, 3], SamplingParams(max_new_tokens=600), timeout=0.1)
    deadline = time.time() + 5
    while sum(w.snapshot()["cancelled"] for w in workers) == 0 and time.time() < deadline:
        time.sleep(0.02)
    assert sum(w.snapshot()["cancelled"] for w in workers) == 1
    with pytest.raises(ValueError):  # a bad request is reported as such and does not mark the replica unhealthy
        r.generate([1] * 5000, sp, timeout=5)
    assert all(w.healthy for w in workers)
    # a worker process that dies fails its requests and leaves the rotation
    workers[0].proc.terminate()
    workers[0].proc.join(5)
    time.sleep(0.2)
    assert not workers[0].healthy
    assert [len(r.generate(p, sp, timeout=10).tokens) for p in prompts[:6]] == [6] * 6
    assert workers[1].snapshot()["completed"] >= 6
    r.shutdown()
    for w in workers:
        w.join(5)
