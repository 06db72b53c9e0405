"""Generate tests/golden/host_golden.json by RUNNING the unmodified reference host code.

TEST INFRASTRUCTURE. Runs only in the build container (needs /root/reference, which does not
exist on the GPU box); the JSON it writes is committed and is what the tests read.

The reference's llm_server/llm_server.py is imported as shipped, with a stub module standing in
for `google.generativeai` (not installed). `servicer.model` is replaced by a
scripted fake that records the prompt and generation_config it receives and returns a
Gemini-shaped response. Every case records: the request, the scripted backend behaviour, the
exact prompt string(s) and generation_config the reference produced, and the parsed result.
These pin the drop-in boundary: prompt construction, response validation,
parsing and every canned fallback.

    python -m oracle.make_host_golden
"""
from __future__ import annotations

import json
import sys
import types
from pathlib import Path

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent.parent / "tests" / "golden" / "host_golden.json"


def load_reference(root: Path = REF):
    """Import the reference's llm_server.py exactly as shipped from `root` (the read-only checkout,
    or a staged copy of its llm_server/ and generated/ directories)."""
    stub = types.ModuleType("google.generativeai")
    stub.configure = lambda **kw: None
    stub.GenerativeModel = lambda name: object()
    import google.protobuf  # noqa: F401  (the real `google` namespace package must stay importable)
    import google

    google.generativeai = stub
    sys.modules["google.generativeai"] = stub
    sys.path.insert(0, str(root))
    sys.path.insert(0, str(root / "llm_server"))
    import importlib

    mod = importlib.import_module("llm_server")
    # back-off sleeps are not part of the result: give the MODULE a stand-in for `time` whose
    # sleep() returns at once (patching time.sleep itself would leak into the whole process)
    real_time = mod.time

    class _NoSleepTime:
        def __getattr__(self, name):
            return getattr(real_time, name)

        @staticmethod
        def sleep(_seconds):
            return None

    if not isinstance(real_time, _NoSleepTime) and getattr(real_time, "__name__", "") == "time":
        mod.time = _NoSleepTime()
    import generated.llm_service_pb2 as pb2

    return mod, pb2


class _Part:
    def __init__(self, text):
        self.text = text


class _Content:
    def __init__(self, parts):
        self.parts = parts


class _Cand:
    def __init__(self, finish_reason, parts):
        self.finish_reason = finish_reason
        self.content = _Content(parts)


class FakeResponse:
    """Gemini-shaped response. `text_attr=False` makes `.text` raise AttributeError, the
    branch the reference handles with candidates[0].content.parts[0].text."""

    def __init__(self, spec):
        self._spec = spec
        self.candidates = [] if spec.get("no_candidates") else [
            _Cand(spec.get("finish_reason", 1), [_Part(t) for t in spec.get("parts", [spec.get("text", "")])])
        ]
        if spec.get("empty_parts"):
            self.candidates[0].content.parts = []

    def __getattr__(self, name):
        if name == "text":
            if self._spec.get("text_attr", True):
                return self._spec.get("text", "")
            raise AttributeError("text")
        raise AttributeError(name)


class FakeModel:
    def __init__(self, script):
        self.script = list(script)
        self.calls = []

    def generate_content(self, prompt, generation_config=None):
        self.calls.append({"prompt": prompt, "generation_config": generation_config})
        spec = self.script.pop(0) if len(self.script) > 1 else self.script[0]
        if spec.get("raise"):
            raise RuntimeError(spec["raise"])
        if spec.get("none"):
            return None
        return FakeResponse(spec)


CONV = [
    {"sender": "alice", "content": "Are we still on for the design review at 3pm?"},
    {"sender": "bob", "content": "Yes, I booked room 4B. Bring the latency numbers."},
    {"sender": "charlie", "content": "I can't make it before 3:30, start without me."},
    {"sender": "alice", "content": "OK. Bob, can you share the doc beforehand?"},
    {"sender": "bob", "content": "Sure, sending it now."},
    {"sender": "dave", "content": "Can I join too?"},
    {"sender": "alice", "content": "Of course!"},
]

BACKENDS = {
    "plain_lines": {"text": "Sounds good to me\nI'll be there\nThanks for the update"},
    "numbered": {"text": "1. Sounds good\n2) See you then\n- Will do\n* extra\n• bullet"},
    "two_lines": {"text": "  Only one reply  \n\nSecond one"},
    "one_line": {"text": "Just this"},
    "empty_text": {"text": "   "},
    "summary_full": {"text": "Summary: The team confirmed the 3pm design review in room 4B.\nCharlie joins late.\n\nKey Points:\n- Review at 3pm in 4B\n- Charlie arrives 3:30\n- Bob shares the doc\n- Fourth point ignored"},
    "summary_lower_bullets": {"text": "Summary: Short one.\nKey points:\n• first\n* not a bullet start\n- second"},
    "summary_unparsable": {"text": "The conversation was about scheduling a meeting and sharing a document before it starts."},
    "summary_long": {"text": "Summary: " + "very long summary text " * 20 + "\nKey Points:\n- a"},
    "ctx_full": {"text": "COMPLETIONS:\n- be there at three\n- bring my notes\n- join remotely\n- four\n- five\n- six\n\nTOPICS:\n- meeting agenda\n- latency numbers\n- third\n- fourth"},
    "ctx_suggestions_word": {"text": "Suggestions:\n• say hello\n- ask about the doc\nTopics\n- docs"},
    "ctx_no_sections": {"text": "- orphan bullet\nnothing else"},
    "answer": {"text": "  The review is at 3pm in room 4B. Bob will share the document first.  "},
    "max_tokens": {"text": "Cut off answer", "finish_reason": 2},
    "blocked": {"text": "x", "finish_reason": 3},
    "no_candidates": {"no_candidates": True, "text": "ignored"},
    "none_response": {"none": True},
    "parts_fallback": {"text_attr": False, "parts": ["From the parts\nSecond line\nThird line"]},
    "parts_empty": {"text_attr": False, "empty_parts": True},
    "raises": {"raise": "backend exploded"},
}


def msgs(pb2, items):
    return [pb2.Message(sender=m["sender"], content=m["content"]) for m in items]


def main():
    mod, pb2 = load_reference()
    servicer = mod.LLMServicer("unused-key")
    cases = []

    def run(kind, request, backend_names, fn):
        model = FakeModel([BACKENDS[b] for b in backend_names])
        servicer.model = model
        result = fn()
        cases.append({"kind": kind, "request": request, "backend": backend_names,
                      "calls": model.calls, "result": result})

    # ---- smart replies (llm_server.py:214-264) ----
    for n_msgs in (0, 1, 2, 5, 7):
        for b in ("plain_lines", "numbered", "two_lines", "one_line", "empty_text", "max_tokens",
                  "blocked", "no_candidates", "none_response", "parts_fallback", "parts_empty", "raises"):
            if n_msgs in (1, 7) and b not in ("plain_lines", "numbered"):
                continue
            req = {"messages": CONV[:n_msgs]}
            run("smart_reply", req, [b],
                lambda: list(servicer._generate_smart_replies(msgs(pb2, CONV[:n_msgs]))))

    # ---- summarize (llm_server.py:266-356) ----
    for n_msgs, max_length in ((0, 200), (3, 200), (7, 200), (7, 60)):
        for b in ("summary_full", "summary_lower_bullets", "summary_unparsable", "summary_long",
                  "blocked", "no_candidates", "none_response", "parts_fallback", "parts_empty", "raises"):
            req = {"messages": CONV[:n_msgs], "max_length": max_length}

            def fn():
                s, k = servicer._summarize_conversation(msgs(pb2, CONV[:n_msgs]), max_length)
                # set() ordering of participants is hash-seed dependent in the reference:
                # record participants separately so tests compare order-insensitively.
                return {"summary": s, "key_points": list(k)}
            run("summarize", req, [b], fn)

    # ---- ask-AI (llm_server.py:147-212) ----
    for ctx in ([], ["alice: hi", "bob: hello"], [f"user{i}: line {i}" for i in range(8)]):
        for b in (["answer"], ["max_tokens"], ["blocked"], ["no_candidates"], ["none_response"],
                  ["parts_fallback"], ["parts_empty"], ["raises"], ["raises", "answer"],
                  ["no_candidates", "answer"]):
            req = {"query": "When is the design review?", "context": ctx}
            run("ask", req, b, lambda: servicer._generate_response(req["query"], list(ctx)))

    # ---- context suggestions (llm_server.py:358-473) ----
    for n_msgs, cur in ((0, ""), (0, "I will"), (3, ""), (7, "I think we should")):
        for b in ("ctx_full", "ctx_suggestions_word", "ctx_no_sections", "blocked", "no_candidates",
                  "none_response", "parts_fallback", "parts_empty", "raises"):
            req = {"messages": CONV[:n_msgs], "current_input": cur}

            def fn():
                s, t = servicer._get_context_suggestions(msgs(pb2, CONV[:n_msgs]), cur)
                return {"suggestions": list(s), "topics": list(t)}
            run("suggest", req, [b], fn)

    # ---- the four RPC handlers end to end (llm_server.py:45-145) ----
    def rpc(kind, req_dict, backend, call):
        model = FakeModel([BACKENDS[b] for b in backend])
        servicer.model = model
        cases.append({"kind": kind, "request": req_dict, "backend": backend, "calls": model.calls,
                      "result": call()})

    for b in (["plain_lines"], ["raises"]):
        rpc("rpc_smart_reply", {"request_id": "r1", "messages": CONV[:3], "user_id": "alice"}, b,
            lambda: (lambda r: {"request_id": r.request_id, "suggestions": list(r.suggestions)})(
                servicer.GetSmartReply(pb2.SmartReplyRequest(request_id="r1", recent_messages=msgs(pb2, CONV[:3]),
                                                             user_id="alice"), None)))
    for b, ml in ((["summary_full"], 0), (["summary_full"], 50), (["raises"], 0)):
        rpc("rpc_summarize", {"request_id": "r2", "messages": CONV[:4], "max_length": ml}, b,
            lambda: (lambda r: {"request_id": r.request_id, "summary": r.summary, "key_points": list(r.key_points)})(
                servicer.SummarizeConversation(pb2.SummarizeRequest(request_id="r2", messages=msgs(pb2, CONV[:4]),
                                                                    max_length=ml), None)))
    for b in (["answer"], ["raises"]):
        rpc("rpc_ask", {"request_id": "r3", "query": "What time?", "context": ["alice: 3pm works"]}, b,
            lambda: (lambda r: {"request_id": r.request_id, "answer": r.answer, "confidence": round(r.confidence, 4)})(
                servicer.GetLLMAnswer(pb2.LLMRequest(request_id="r3", query="What time?",
                                                     context=["alice: 3pm works"]), None)))
    for b in (["ctx_full"], ["raises"]):
        rpc("rpc_suggest", {"request_id": "r4", "messages": CONV[:2], "current_input": "Let's"}, b,
            lambda: (lambda r: {"request_id": r.request_id, "suggestions": list(r.suggestions), "topics": list(r.topics)})(
                servicer.GetContextSuggestions(pb2.ContextRequest(request_id="r4", context=msgs(pb2, CONV[:2]),
                                                                  current_input="Let's"), None)))

    OUT.parent.mkdir(parents=True, exist_ok=True)
    OUT.write_text(json.dumps({"generated_by": "oracle/make_host_golden.py",
                               "reference": "llm_server/llm_server.py (unmodified, genai stubbed)",
                               "backends": BACKENDS, "cases": cases}, indent=1, ensure_ascii=False))
    print(f"wrote {len(cases)} cases to {OUT}")


if __name__ == "__main__":
    main()
