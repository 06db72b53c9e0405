"""gRPC integration on 127.0.0.1: all four llm.LLMService RPCs through the build-owned bootstrap,
LocalGenerativeModel -> Router -> (fake) engine, and wire compatibility with the reference's own
generated stub when the reference tree is present."""
import os
import sys
from concurrent import futures

import grpc
import pytest

from conftest import pkg
from fakes import FakeEngine

server_mod = pkg("server")
svc = pkg("llm_service")
pb = pkg("llm_proto")
sched = pkg("scheduler")
local_model = pkg("local_model")
tokenizer = pkg("tokenizer")


class CannedModel:
    def __init__(self, text, finish_reason=1):
        self.text, self.finish_reason, self.prompts = text, finish_reason, []

    def generate_content(self, prompt, generation_config=None):
        self.prompts.append((prompt, generation_config))
        return local_model.GenerateContentResponse(
            candidates=[local_model.Candidate(self.finish_reason,
                                              local_model.Content([local_model.Part(self.text)]))])


@pytest.fixture()
def served():
    servers = []

    def start(model):
        server = grpc.server(futures.ThreadPoolExecutor(max_workers=8))
        server_mod.add_llm_service(svc.LLMServicer(model), server)
        port = server.add_insecure_port("127.0.0.1:0")
        server.start()
        servers.append(server)
        return f"127.0.0.1:{port}"

    yield start
    for s in servers:
        s.stop(0)


def test_all_four_rpcs_are_registered_and_answer(served):
    model = CannedModel("Sounds good\nSee you there\nThanks!")
    with grpc.insecure_channel(served(model)) as ch:
        stub = server_mod.LLMServiceStub(ch)
        msgs = [pb.Message(sender="alice", content="lunch at noon?"), pb.Message(sender="bob", content="sure")]
        r = stub.GetSmartReply(pb.SmartReplyRequest(request_id="1", recent_messages=msgs, user_id="bob"), timeout=5)
        assert r.request_id == "1" and list(r.suggestions) == ["Sounds good", "See you there", "Thanks!"]
        model.text = "Summary: Lunch is at noon.\nKey Points:\n- noon\n- bob agrees"
        r = stub.SummarizeConversation(pb.SummarizeRequest(request_id="2", messages=msgs), timeout=5)
        assert r.summary == "Lunch is at noon." and list(r.key_points) == ["noon", "bob agrees"]
        assert "in under 200 characters" in model.prompts[-1][0]          # default max_length (llm_server.py:96)
        model.text = "  At noon.  "
        r = stub.GetLLMAnswer(pb.LLMRequest(request_id="3", query="when?", context=["alice: lunch at noon?"]), timeout=5)
        assert r.answer == "At noon." and abs(r.confidence - 0.95) < 1e-6
        assert model.prompts[-1][1] == {"temperature": 0.7, "max_output_tokens": 150}
        model.text = "COMPLETIONS:\n- be there\n- bring food\nTOPICS:\n- lunch"
        r = stub.GetContextSuggestions(pb.ContextRequest(request_id="4", context=msgs, current_input="I will"), timeout=5)
        assert list(r.suggestions) == ["be there", "bring food"] and list(r.topics) == ["lunch"]


def test_backend_failure_never_becomes_a_grpc_error(served):
    class Broken:
        def generate_content(self, *a, **k):
            raise RuntimeError("GPU on fire")

    svc._sleep = lambda s: None
    with grpc.insecure_channel(served(Broken())) as ch:
        stub = server_mod.LLMServiceStub(ch)
        m = [pb.Message(sender="a", content="b")]
        assert list(stub.GetSmartReply(pb.SmartReplyRequest(request_id="x", recent_messages=m), timeout=5).suggestions) == \
            ["I agree", "That's interesting", "Tell me more"]
        assert "trouble connecting" in stub.GetLLMAnswer(pb.LLMRequest(request_id="y", query="q"), timeout=5).answer


def test_local_model_through_router_returns_gemini_shaped_response():
    eng = FakeEngine(max_seqs=4, max_seq_len=512, max_batch_tokens=1024, max_new_tokens=32)
    router = sched.Router([sched.EngineWorker(eng)])
    model = local_model.LocalGenerativeModel(router, tokenizer.ByteLevelTokenizer(), default_max_new_tokens=8)
    resp = model.generate_content("hello")
    assert resp and resp.candidates and resp.candidates[0].finish_reason == 2
    assert resp.text == resp.candidates[0].content.parts[0].text
    assert resp.usage["prompt_tokens"] == 5 and resp.usage["generated_tokens"] == 8
    resp = model.generate_content("hello", generation_config={"temperature": 0.7, "max_output_tokens": 3})
    assert resp.usage["generated_tokens"] == 3
    # and the whole stack end to end: servicer -> local model -> router -> engine
    out = svc.generate_smart_replies(model, [pb.Message(sender="a", content="hi")])
    assert len(out) >= 2
    router.shutdown()


@pytest.mark.skipif(not os.path.isdir("/root/reference/generated"), reason="reference tree not present")
def test_reference_generated_stub_can_call_this_server(served):
    """Drop-in at the wire: the Raft node's own LLMServiceStub (unmodified generated code) calls
    a server built from llm_proto.py, including the two RPCs the reference fails to register."""
    sys.path.insert(0, "/root/reference")
    import generated.llm_service_pb2 as ref_pb2
    import generated.llm_service_pb2_grpc as ref_grpc

    model = CannedModel("one\ntwo\nthree")
    with grpc.insecure_channel(served(model)) as ch:
        stub = ref_grpc.LLMServiceStub(ch)
        r = stub.GetSmartReply(ref_pb2.SmartReplyRequest(
            request_id="r", recent_messages=[ref_pb2.Message(sender="a", content="b")], user_id="u"), timeout=5)
        assert list(r.suggestions) == ["one", "two", "three"]
        model.text = "fine"
        r = stub.GetLLMAnswer(ref_pb2.LLMRequest(request_id="q", query="how?", parameters={"k": "v"}), timeout=5)
        assert r.answer == "fine"   # UNIMPLEMENTED with the reference's own server
        model.text = "COMPLETIONS:\n- a\nTOPICS:\n- t"
        r = stub.GetContextSuggestions(ref_pb2.ContextRequest(request_id="c", current_input="x"), timeout=5)
        assert list(r.suggestions) == ["a"] and list(r.topics) == ["t"]
