"""The drop-in, proven on the reference's OWN class.

The unmodified /root/reference/llm_server/llm_server.py is imported as shipped (only
`google.generativeai`, which is not installed, is stubbed -- the same stub oracle/make_host_golden.py
uses), its LLMServicer is constructed, and the one attribute the maintainer-side change touches,
`servicer.model` (llm_server.py:29-34), is replaced by this repository's LocalGenerativeModel.
All four handlers (llm_server.py:45-145) are then called with the reference's generated protobuf
messages. Nothing else of the reference is altered.

Two flavours:
  * CPU (`-m "not gpu"`): the model sits on the scheduler with the fake engine of tests/fakes.py --
    this proves the seam (call shape, response shape, threading) wherever the reference checkout
    exists, i.e. in the build container on every round;
  * GPU (`-m gpu`): the same with the CUDA engine underneath, asserting kernel launches.
The reference checkout does not exist on the GPU box. The GPU flavour therefore also looks for a
staged copy (directories llm_server/ and generated/) under $B200LLM_REFERENCE_DIR or
baseline/_ref/ and skips when there is none; profiles/r2_reference_dropin_gpu.log is the record of
such a staged run.
"""
import os
import threading
from pathlib import Path

import pytest

from conftest import ROOT, pkg

CANDIDATES = [os.environ.get("B200LLM_REFERENCE_DIR", ""), "/root/reference", os.path.join(ROOT, "baseline", "_ref")]


def _reference_root():
    for c in CANDIDATES:
        if c and (Path(c) / "llm_server" / "llm_server.py").exists() and (Path(c) / "generated").is_dir():
            return Path(c)
    return None


@pytest.fixture(scope="module")
def reference():
    root = _reference_root()
    if root is None:
        pytest.skip("no reference checkout (/root/reference) or staged copy on this machine")
    from oracle.make_host_golden import load_reference

    mod, pb2 = load_reference(root)
    return mod, pb2


def _exercise_all_four_handlers(mod, pb2, model):
    servicer = mod.LLMServicer("x")                 # reference constructor, stubbed genai inside
    servicer.model = model                          # <- the whole maintainer-side change
    msgs = [pb2.Message(sender="alice", content="Are we still on for the design review at 3pm?"),
            pb2.Message(sender="bob", content="Yes, I booked room 4B. Bring the latency numbers.")]
    r = servicer.GetSmartReply(pb2.SmartReplyRequest(request_id="s1", recent_messages=msgs, user_id="bob"), None)
    # three parsed lines, or the reference's own padding ("I agree", "Interesting point", llm_server.py:251-260)
    # when the random-weight model's text has fewer usable lines
    assert r.request_id == "s1" and 2 <= len(r.suggestions) <= 3 and all(isinstance(x, str) and x for x in r.suggestions)
    r = servicer.SummarizeConversation(pb2.SummarizeRequest(request_id="m1", messages=msgs, max_length=120), None)
    # (a random-weight model's text rarely contains "Summary:": the reference then returns what its
    # parser found, possibly an empty summary, plus its default key points, llm_server.py:316-348)
    assert r.request_id == "m1" and len(r.summary) <= 120 and 1 <= len(r.key_points) <= 3
    r = servicer.GetLLMAnswer(pb2.LLMRequest(request_id="a1", query="When is the review?", context=["alice: 3pm"]), None)
    assert r.request_id == "a1" and isinstance(r.answer, str) and r.answer
    assert r.confidence in (0.0, pytest.approx(0.95))
    r = servicer.GetContextSuggestions(pb2.ContextRequest(request_id="c1", context=msgs, current_input="I will"), None)
    assert r.request_id == "c1" and 1 <= len(r.suggestions) <= 5 and 1 <= len(r.topics) <= 3
    # the reference calls the model from a thread pool without a lock (llm_server.py:501)
    out = []

    def call(i):
        out.append(servicer.GetSmartReply(pb2.SmartReplyRequest(request_id=str(i), recent_messages=msgs), None))
    ts = [threading.Thread(target=call, args=(i,)) for i in range(8)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert sorted(int(o.request_id) for o in out) == list(range(8)) and all(2 <= len(o.suggestions) <= 3 for o in out)
    return servicer


def test_reference_servicer_with_the_local_model_on_a_fake_engine(reference):
    from fakes import FakeEngine

    mod, pb2 = reference
    sched, lm, tok = pkg("scheduler"), pkg("local_model"), pkg("tokenizer")

    class ByteFake(FakeEngine):
        """tokens in the printable byte range so that the reference's parsers get real text"""
        def _emit(self, st):
            super()._emit(st)
            if st["tokens"] and st["tokens"][-1] != self.shape.eos_token_id:
                st["tokens"][-1] = 33 + st["tokens"][-1] % 60

    eng = ByteFake(max_seqs=8, max_seq_len=2048, max_batch_tokens=8192, max_new_tokens=160)
    router = sched.Router([sched.EngineWorker(eng, decode_burst=8)])
    model = lm.LocalGenerativeModel(router, tok.ByteLevelTokenizer(), default_max_new_tokens=24,
                                    max_prompt_tokens=1800)
    try:
        servicer = _exercise_all_four_handlers(mod, pb2, model)
        assert type(servicer).__module__ == "llm_server", "this must be the reference's class, not our mirror"
        assert any(c[0] == "prefill" for c in eng.calls) and not eng.slots
    finally:
        router.shutdown()


@pytest.mark.gpu
@pytest.mark.timeout(600)
def test_reference_servicer_with_the_local_model_on_the_cuda_engine(reference, cuda):
    from oracle import gpt2_oracle as O

    mod, pb2 = reference
    E, sched, lm, tok = pkg("engine"), pkg("scheduler"), pkg("local_model"), pkg("tokenizer")
    w = O.synthetic_weights(O.SMALL, seed=0)
    eng = E.Engine(E.GPT2_SMALL, O.pack_blob(O.SMALL, w).numpy(), max_seqs=16, max_seq_len=1024,
                   max_batch_tokens=8192, max_new_tokens=160, prefix_cache=True)
    router = sched.Router([sched.EngineWorker(eng, decode_burst=8)])
    model = lm.LocalGenerativeModel(router, tok.ByteLevelTokenizer(), default_max_new_tokens=24,
                                    max_prompt_tokens=800)
    try:
        launches0 = eng.kernel_launches
        servicer = _exercise_all_four_handlers(mod, pb2, model)
        assert type(servicer).__module__ == "llm_server"
        assert eng.kernel_launches > launches0, "the CUDA engine did not run"
        print(f"reference LLMServicer on the CUDA engine: {eng.kernel_launches - launches0} kernel launches, "
              f"prefix cache {eng.prefix_cache_stats()}")
    finally:
        router.shutdown()
        eng.close()
