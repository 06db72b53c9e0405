"""Host side of llm.LLMService: prompt construction, response validation and output parsing.

Behavioural mirror of the reference's LLMServicer (llm_server/llm_server.py:23-473) with one
difference: the generator behind ``self.model`` is whatever is passed in -- by default the local
B200 engine (local_model.LocalGenerativeModel) instead of a remote Gemini client
(llm_server.py:29-34). Everything the Raft node can observe is kept: the four RPC names and
message types, the exact prompt text sent to the generator, the ``generation_config`` ask-AI
passes, the finish_reason/candidates validation idiom, the line parsers, every canned fallback,
and "never surface an exception or a gRPC error status" (llm_server.py:62-68, 85-90, 109-115,
137-145). tests/test_host_golden.py replays 155 recorded runs of the unmodified reference
through this module.

Structure (table-driven rather than four hand-unrolled copies):
  PROMPTS      the templates, keyed by use case
  _ask()       one generator call + the validation idiom -> (status, text)
  _FALLBACKS   canned outputs per (use case, status)
  parsers      smart-reply lines, summary / key points, completions / topics
"""
from __future__ import annotations

import logging
import time
from typing import Callable, Sequence

from . import llm_proto as pb

logger = logging.getLogger(__name__)

FINISH_STOP = 1        # llm_server.py:184 "1=STOP, 2=MAX_TOKENS"
FINISH_MAX_TOKENS = 2
_ACCEPTED_FINISH = (FINISH_STOP, FINISH_MAX_TOKENS)

ASK_GENERATION_CONFIG = {"temperature": 0.7, "max_output_tokens": 150}   # llm_server.py:169-172
ASK_MAX_ATTEMPTS = 3                                                      # llm_server.py:164
CONTEXT_WINDOW = 5                                                        # last-5 slices :152, :220, :363

# --------------------------------------------------------------------------------------------
# Prompt templates (text identical to llm_server.py:153-161, :221-229, :273-285, :369-401)
# --------------------------------------------------------------------------------------------
PROMPTS = {
    "ask_with_context": (
        "Based on this recent conversation context:\n\n{context}\n\n"
        "User's question: {query}\n"
        "Do not give more than 2 sentences in your response.\n"
        "Provide a helpful, short, informative response that considers the conversation context:"
    ),
    "ask_plain": "{query}\n\nProvide a short, helpful answer in 2 sentences or less.",
    "smart_reply": (
        "Based on this conversation:\n{conversation}\n\n"
        "Generate exactly 3 short, natural reply suggestions. Each suggestion should be:\n"
        "- Under 10 words\n- Contextually relevant\n- Natural and conversational\n\n"
        "Format: Just list the 3 suggestions, one per line, no numbering or bullets."
    ),
    "summarize": (
        "Summarize this conversation concisely in under {max_length} characters:\n\n{conversation}\n\n"
        "Then provide 3 key bullet points about the discussion.\n\n"
        "Format:\nSummary: [your summary here]\n\nKey Points:\n- point 1\n- point 2\n- point 3"
    ),
    "suggest_typing": (
        "Based on this conversation context:\n{context}\n\n"
        "User started typing: \"{current_input}\"\n\n"
        "Provide 3 natural completions for what they might want to say next, completing their thought.\n"
        "Also suggest 2 related topics they could discuss.\n\n"
        "Format as simple lists:\nCOMPLETIONS:\n- completion 1\n- completion 2  \n- completion 3\n\n"
        "TOPICS:\n- topic 1\n- topic 2"
    ),
    "suggest_idle": (
        "Based on this conversation context:\n{context}\n\n"
        "Suggest 3 natural things the user might want to say next.\n"
        "Also suggest 2 related topics they could discuss.\n\n"
        "Format as simple lists:\nCOMPLETIONS:\n- suggestion 1\n- suggestion 2\n- suggestion 3\n\n"
        "TOPICS:\n- topic 1\n- topic 2"
    ),
}

BULLET_CHARS = "-•* "
NUMBERING_CHARS = "0123456789.-•*) "  # llm_server.py:256


def _transcript(messages: Sequence) -> str:
    return "\n".join(f"{m.sender}: {m.content}" for m in messages)


# --------------------------------------------------------------------------------------------
# One generator call + the reference's validation idiom (llm_server.py:176-194, 234-249, ...)
# --------------------------------------------------------------------------------------------
OK, EMPTY, BLOCKED, NO_TEXT = "ok", "empty", "blocked", "no_text"


def _ask(model, prompt: str, generation_config: dict | None = None) -> tuple[str, str]:
    """Returns (status, text). Exceptions from the generator propagate to the caller, which
    owns the catch-all (as in the reference)."""
    response = model.generate_content(prompt, generation_config=generation_config) \
        if generation_config is not None else model.generate_content(prompt)
    if not response or not response.candidates:
        return EMPTY, ""
    first = response.candidates[0]
    if first.finish_reason not in _ACCEPTED_FINISH:
        logger.warning("generation blocked: finish_reason=%s", first.finish_reason)
        return BLOCKED, ""
    try:
        return OK, response.text.strip()
    except AttributeError:
        parts = first.content.parts
        if parts:
            return OK, parts[0].text.strip()
        return NO_TEXT, ""


def _participants(messages: Sequence) -> list[str]:
    return list(set(m.sender for m in messages))  # unordered in the reference too (:293)


# --------------------------------------------------------------------------------------------
# Use cases
# --------------------------------------------------------------------------------------------
_sleep = time.sleep  # the retry back-off's clock (llm_server.py:164, :204); tests replace THIS, not time.sleep


def generate_response(model, query: str, context_list: Sequence[str],
                      sleep: Callable[[float], None] | None = None) -> str:
    """Ask-AI (llm_server.py:147-212): up to three attempts with 1 s / 2 s back-off."""
    sleep = sleep or _sleep
    try:
        if context_list:
            prompt = PROMPTS["ask_with_context"].format(context="\n".join(context_list[-CONTEXT_WINDOW:]), query=query)
        else:
            prompt = PROMPTS["ask_plain"].format(query=query)
        for attempt in range(ASK_MAX_ATTEMPTS):
            last = attempt == ASK_MAX_ATTEMPTS - 1
            try:
                status, text = _ask(model, prompt, dict(ASK_GENERATION_CONFIG))
            except Exception as exc:  # noqa: BLE001 - the reference retries on anything
                if last:
                    raise
                logger.warning("generator retry %d/%d: %s", attempt + 1, ASK_MAX_ATTEMPTS, str(exc)[:80])
                sleep(2 ** attempt)
                continue
            if status == OK:
                return text
            if status == BLOCKED:
                return "I cannot provide a response to that query. Please try asking something else."
            if not last:
                sleep(2 ** attempt)
                continue
            if status == EMPTY:
                return "I'm having trouble generating a response. Please try rephrasing your question."
            return "I received an empty response. Please try again."
    except Exception as exc:  # noqa: BLE001
        logger.error("generator failed after retries: %s", exc)
        return ("I'm having trouble connecting to the AI service right now. "
                "The API might be overloaded. Please try again in a moment.")


_SMART_REPLY_FALLBACK = {
    EMPTY: ["I agree", "That's interesting", "Tell me more"],
    BLOCKED: ["Sounds good", "I understand", "Interesting"],
    NO_TEXT: ["I agree", "That makes sense", "Good point"],
}
_SMART_REPLY_ERROR = ["I agree", "That's interesting", "Tell me more"]
_SMART_REPLY_PADDING = ["I agree", "Interesting point"]


def generate_smart_replies(model, messages: Sequence) -> list[str]:
    """Smart replies (llm_server.py:214-264)."""
    if not messages:
        return ["Hello!", "How can I help?", "What's on your mind?"]
    try:
        prompt = PROMPTS["smart_reply"].format(conversation=_transcript(messages[-CONTEXT_WINDOW:]))
        status, text = _ask(model, prompt)
        if status != OK:
            return list(_SMART_REPLY_FALLBACK[status])
        cleaned = []
        for line in text.split("\n"):
            line = line.strip()
            if not line:
                continue
            line = line.lstrip(NUMBERING_CHARS)
            if line:
                cleaned.append(line)
        if len(cleaned) >= 3:
            return cleaned[:3]
        return cleaned + _SMART_REPLY_PADDING[: 3 - len(cleaned)]
    except Exception as exc:  # noqa: BLE001
        logger.error("smart reply error: %s", exc)
        return list(_SMART_REPLY_ERROR)


def summarize_conversation(model, messages: Sequence, max_length: int) -> tuple[str, list[str]]:
    """Summary + key points (llm_server.py:266-356)."""
    if not messages:
        return "No messages to summarize", []
    try:
        prompt = PROMPTS["summarize"].format(max_length=max_length, conversation=_transcript(messages))
        status, text = _ask(model, prompt)
        n = len(messages)
        if status == EMPTY:
            return f"Conversation with {n} messages", [f"{n} messages exchanged"]
        if status == BLOCKED:
            return f"Discussion between {', '.join(_participants(messages))}", [f"{n} messages", "Active conversation"]
        if status == NO_TEXT:
            return f"Conversation between {', '.join(_participants(messages))}", [f"{n} messages"]

        summary, key_points, in_points = "", [], False
        for raw in text.split("\n"):
            line = raw.strip()
            if line.startswith("Summary:"):
                summary = line.replace("Summary:", "").strip()
            elif "Key Points:" in line or "Key points:" in line:
                in_points = True
            elif in_points and (line.startswith("-") or line.startswith("•")):
                point = line.lstrip(BULLET_CHARS).strip()
                if point:
                    key_points.append(point)
            elif not in_points and summary and line:
                summary += " " + line
        if len(summary) > max_length:
            summary = summary[: max_length - 3] + "..."
        if not summary:
            summary = text[: max_length - 3] + "..." if len(text) > max_length else text
        if not key_points:
            key_points = [f"{n} messages exchanged",
                          f"Participants: {', '.join(_participants(messages)[:3])}",
                          "Active discussion"]
        return summary, key_points[:3]
    except Exception as exc:  # noqa: BLE001
        logger.error("summarize error: %s", exc)
        people = _participants(messages)
        return (f"Conversation between {', '.join(people)}",
                [f"{len(messages)} messages", f"Participants: {len(people)}"])


_SUGGEST_FALLBACK = {
    EMPTY: (["continue the conversation", "ask for details", "share thoughts"], ["discussion topic"]),
    BLOCKED: (["continue the thought", "ask a question", "share more"], ["current topic"]),
    NO_TEXT: (["continue the conversation", "ask for clarification"], ["related subjects"]),
}
_SUGGEST_ERROR = (["continue the conversation", "ask for clarification", "share thoughts"],
                  ["discussion topic", "related subjects"])


def get_context_suggestions(model, messages: Sequence, current_input: str) -> tuple[list[str], list[str]]:
    """Completions + topics (llm_server.py:358-473)."""
    try:
        context = _transcript(messages[-CONTEXT_WINDOW:]) if messages else "No previous context"
        if current_input:
            prompt = PROMPTS["suggest_typing"].format(context=context, current_input=current_input)
        else:
            prompt = PROMPTS["suggest_idle"].format(context=context)
        status, text = _ask(model, prompt)
        if status != OK:
            s, t = _SUGGEST_FALLBACK[status]
            return list(s), list(t)

        buckets = {"suggestions": [], "topics": []}
        section = None
        for raw in text.split("\n"):
            line = raw.strip()
            upper = line.upper()
            if "COMPLETION" in upper or "SUGGESTION" in upper:
                section = "suggestions"
            elif "TOPIC" in upper:
                section = "topics"
            elif line.startswith("-") or line.startswith("•"):
                item = line.lstrip(BULLET_CHARS).strip()
                if item and section is not None:
                    buckets[section].append(item)
        suggestions, topics = buckets["suggestions"], buckets["topics"]
        if not suggestions:
            if current_input:
                suggestions = [f"{current_input} be the best option", f"{current_input} work well",
                               f"{current_input} make sense"]
            else:
                suggestions = ["continue the thought", "ask a question", "share more details"]
        if not topics:
            topics = ["current discussion", "related ideas"]
        return suggestions[:5], topics[:3]
    except Exception as exc:  # noqa: BLE001
        logger.error("context suggestions error: %s", exc)
        return list(_SUGGEST_ERROR[0]), list(_SUGGEST_ERROR[1])


# --------------------------------------------------------------------------------------------
# gRPC servicer (method names and signatures of the reference's LLMServicer)
# --------------------------------------------------------------------------------------------
class LLMServicer:
    """llm.LLMService handlers. ``model`` is the drop-in seam (reference: self.model,
    llm_server.py:33): any object with ``generate_content(prompt, generation_config=None)``."""

    def __init__(self, model, backend: str = "b200-local"):
        self.model = model
        self.backend = backend

    # llm_server.py:45-68
    def GetLLMAnswer(self, request, context):
        request_id = request.request_id
        try:
            answer = generate_response(self.model, request.query, list(request.context) if request.context else [])
            return pb.LLMResponse(request_id=request_id, answer=answer, confidence=0.95)
        except Exception as exc:  # noqa: BLE001
            logger.error("GetLLMAnswer failed: %s", exc)
            return pb.LLMResponse(
                request_id=request_id,
                answer="I apologize, but I'm having trouble processing your request. Please try again.",
                confidence=0.0)

    # llm_server.py:70-90
    def GetSmartReply(self, request, context):
        request_id = request.request_id
        try:
            suggestions = generate_smart_replies(self.model, list(request.recent_messages))
            return pb.SmartReplyResponse(request_id=request_id, suggestions=suggestions)
        except Exception as exc:  # noqa: BLE001
            logger.error("GetSmartReply failed: %s", exc)
            return pb.SmartReplyResponse(request_id=request_id,
                                         suggestions=["I agree", "That's interesting", "Tell me more"])

    # llm_server.py:92-115
    def SummarizeConversation(self, request, context):
        request_id = request.request_id
        max_length = request.max_length if request.max_length > 0 else 200
        try:
            summary, key_points = summarize_conversation(self.model, list(request.messages), max_length)
            return pb.SummarizeResponse(request_id=request_id, summary=summary, key_points=key_points)
        except Exception as exc:  # noqa: BLE001
            logger.error("SummarizeConversation failed: %s", exc)
            return pb.SummarizeResponse(request_id=request_id, summary="Unable to generate summary at this time.",
                                        key_points=[])

    # llm_server.py:117-145
    def GetContextSuggestions(self, request, context):
        request_id = request.request_id
        try:
            suggestions, topics = get_context_suggestions(self.model, list(request.context), request.current_input)
            return pb.SuggestionsResponse(request_id=request_id, suggestions=suggestions, topics=topics)
        except Exception as exc:  # noqa: BLE001
            logger.error("GetContextSuggestions failed: %s", exc)
            return pb.SuggestionsResponse(request_id=request_id,
                                          suggestions=["sounds interesting", "tell me more", "I see"], topics=[])
