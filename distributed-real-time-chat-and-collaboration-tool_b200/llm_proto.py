"""Wire schema of the reference's llm.LLMService, built at import time.

The reference ships protoc-generated modules (generated/llm_service_pb2.py) whose service
registration is incomplete (only 2 of 4 RPCs) and protoc is not available here,
so the schema is restated as data and turned into message classes with protobuf's descriptor
API. Field names, numbers and types are those of protos/llm_service.proto:12-62 plus the
GetLLMAnswer method the serialized descriptor in generated/llm_service_pb2.py carries; the wire
format is therefore identical and the reference's own LLMServiceStub can call a server built
from these classes (tests/test_service.py checks both directions when the reference is present).

A private DescriptorPool keeps these definitions from colliding with the reference's generated
module if both are imported into one process.
"""
from __future__ import annotations

from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

PACKAGE = "llm"
SERVICE = "LLMService"

_F = descriptor_pb2.FieldDescriptorProto
_STRING, _INT32, _FLOAT, _MSG = _F.TYPE_STRING, _F.TYPE_INT32, _F.TYPE_FLOAT, _F.TYPE_MESSAGE

# message -> [(field, number, type, repeated, message type name)]
_SCHEMA = {
    "Message": [("sender", 1, _STRING, False, None), ("content", 2, _STRING, False, None)],
    "LLMRequest": [("request_id", 1, _STRING, False, None), ("query", 2, _STRING, False, None),
                   ("context", 3, _STRING, True, None), ("parameters", 4, _MSG, True, "LLMRequest.ParametersEntry")],
    "LLMResponse": [("request_id", 1, _STRING, False, None), ("answer", 2, _STRING, False, None),
                    ("confidence", 3, _FLOAT, False, None)],
    "SmartReplyRequest": [("request_id", 1, _STRING, False, None), ("recent_messages", 2, _MSG, True, "Message"),
                          ("user_id", 3, _STRING, False, None)],
    "SmartReplyResponse": [("request_id", 1, _STRING, False, None), ("suggestions", 2, _STRING, True, None)],
    "SummarizeRequest": [("request_id", 1, _STRING, False, None), ("messages", 2, _MSG, True, "Message"),
                         ("max_length", 3, _INT32, False, None)],
    "SummarizeResponse": [("request_id", 1, _STRING, False, None), ("summary", 2, _STRING, False, None),
                          ("key_points", 3, _STRING, True, None)],
    "ContextRequest": [("request_id", 1, _STRING, False, None), ("context", 2, _MSG, True, "Message"),
                       ("current_input", 3, _STRING, False, None)],
    "SuggestionsResponse": [("request_id", 1, _STRING, False, None), ("suggestions", 2, _STRING, True, None),
                            ("topics", 3, _STRING, True, None)],
}

# rpc -> (request message, response message); all unary-unary
METHODS = {
    "GetLLMAnswer": ("LLMRequest", "LLMResponse"),
    "GetSmartReply": ("SmartReplyRequest", "SmartReplyResponse"),
    "SummarizeConversation": ("SummarizeRequest", "SummarizeResponse"),
    "GetContextSuggestions": ("ContextRequest", "SuggestionsResponse"),
}


def _build_file() -> descriptor_pb2.FileDescriptorProto:
    fd = descriptor_pb2.FileDescriptorProto(name="b200llm/llm_service.proto", package=PACKAGE, syntax="proto3")
    for msg_name, fields in _SCHEMA.items():
        m = fd.message_type.add(name=msg_name)
        for fname, number, ftype, repeated, type_name in fields:
            f = m.field.add(name=fname, number=number, type=ftype,
                            label=_F.LABEL_REPEATED if repeated else _F.LABEL_OPTIONAL)
            if type_name:
                f.type_name = f".{PACKAGE}.{type_name}"
        if msg_name == "LLMRequest":  # map<string, string> parameters = 4
            entry = m.nested_type.add(name="ParametersEntry")
            entry.field.add(name="key", number=1, type=_STRING, label=_F.LABEL_OPTIONAL)
            entry.field.add(name="value", number=2, type=_STRING, label=_F.LABEL_OPTIONAL)
            entry.options.map_entry = True
    svc = fd.service.add(name=SERVICE)
    for rpc, (req, resp) in METHODS.items():
        svc.method.add(name=rpc, input_type=f".{PACKAGE}.{req}", output_type=f".{PACKAGE}.{resp}")
    return fd


_POOL = descriptor_pool.DescriptorPool()
_FILE = _POOL.Add(_build_file()) if hasattr(_POOL, "Add") else None
if _FILE is None:  # newer protobuf: AddSerializedFile
    _POOL.AddSerializedFile(_build_file().SerializeToString())


def _cls(name: str):
    return message_factory.GetMessageClass(_POOL.FindMessageTypeByName(f"{PACKAGE}.{name}"))


Message = _cls("Message")
LLMRequest = _cls("LLMRequest")
LLMResponse = _cls("LLMResponse")
SmartReplyRequest = _cls("SmartReplyRequest")
SmartReplyResponse = _cls("SmartReplyResponse")
SummarizeRequest = _cls("SummarizeRequest")
SummarizeResponse = _cls("SummarizeResponse")
ContextRequest = _cls("ContextRequest")
SuggestionsResponse = _cls("SuggestionsResponse")

MESSAGE_CLASSES = {n: _cls(n) for n in _SCHEMA}


def method_path(rpc: str) -> str:
    return f"/{PACKAGE}.{SERVICE}/{rpc}"
