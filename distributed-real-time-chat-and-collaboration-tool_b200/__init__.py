"""B200-native local text generation for the chat tool's llm.LLMService.

Drop-in replacement for the remote generator behind the reference's
``LLMServicer.model.generate_content`` (reference llm_server/llm_server.py:32-33, :167, :231,
:287, :403): GPT-2-class causal-LM generation executed by hand-written sm_100a CUDA kernels
behind a C ABI (include/b200llm.h). See DESIGN.md and INTEGRATION.md.

    _lib          ctypes binding of libb200llm.so (fails loudly if it is missing)
    engine        Engine: model executor + paged KV-cache manager + sampler (device side)
    ops           operator-level entry points (tests / profiling)
    weights       weight blobs (random init, HF state dict, safetensors)
    tokenizer     GPT-2 byte-level / BPE tokenisation
    scheduler     per-GPU continuous batching + request router
    local_model   LocalGenerativeModel.generate_content (the drop-in seam)
    llm_service   LLMServicer: prompts, validation, parsers, fallbacks (reference behaviour)
    llm_proto     llm.LLMService wire schema
    server        gRPC bootstrap registering all four RPCs
"""
import sys as _sys

from . import _lib  # noqa: F401  (fails loudly when libb200llm.so is missing)
from ._lib import B200Error  # noqa: F401
from . import engine, llm_proto, llm_service, local_model, scheduler, tokenizer, weights  # noqa: F401,E402

# `import b200llm.<sub>` resolves to these same module objects (see /b200llm.py).
for _name in ("_lib", "engine", "llm_proto", "llm_service", "local_model", "scheduler", "tokenizer", "weights"):
    _sys.modules.setdefault("b200llm." + _name, _sys.modules[__name__ + "." + _name])

__all__ = ["B200Error", "engine", "llm_proto", "llm_service", "local_model", "scheduler", "tokenizer", "weights"]
