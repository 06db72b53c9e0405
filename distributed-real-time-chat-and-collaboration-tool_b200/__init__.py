"""B200-native local text generation for the chat tool's llm.LLMService.

Drop-in replacement for the remote generator behind the reference's
``LLMServicer.model.generate_content`` (reference llm_server/llm_server.py:32-33, :167, :231,
:287, :403): GPT-2-class causal-LM generation executed by hand-written sm_100a CUDA kernels
behind a C ABI (include/b200llm.h). See DESIGN.md and INTEGRATION.md.
"""
from . import _lib  # noqa: F401  (fails loudly when libb200llm.so is missing)
from ._lib import B200Error  # noqa: F401

__all__ = ["B200Error"]
