"""ctypes binding of libb200llm.so (the C ABI declared in include/b200llm.h).

The product path has no CPU or PyTorch fallback: if the shared library is missing or a symbol
is absent, importing this module raises. Build it with `python __graft_entry__.py build` or
`python distributed-real-time-chat-and-collaboration-tool_b200/build.py`.
"""
from __future__ import annotations

import ctypes as C
import re
from pathlib import Path

HERE = Path(__file__).resolve().parent
import os as _os

# B200LLM_LIB selects another build of the same library (profiling variants); default in-tree.
LIB_PATH = Path(_os.environ["B200LLM_LIB"]) if _os.environ.get("B200LLM_LIB") else HERE / "libb200llm.so"
HEADER_PATH = HERE.parent / "include" / "b200llm.h"


class B200Error(RuntimeError):
    """A C-ABI call returned a non-zero status."""

    def __init__(self, code: int, message: str):
        super().__init__(f"b200llm error {code}: {message}")
        self.code = code
        self.message = message


class Config(C.Structure):
    """b200llm_config, field for field (tests/test_cabi_exports.py compares this with the header).
    struct_size is filled in on construction."""
    _fields_ = [
        ("struct_size", C.c_uint32),
        ("n_layer", C.c_int32),
        ("n_head", C.c_int32),
        ("d_model", C.c_int32),
        ("vocab_size", C.c_int32),
        ("n_positions", C.c_int32),
        ("ln_eps", C.c_float),
        ("eos_token_id", C.c_int32),
        ("pad_token_id", C.c_int32),
        ("max_seqs", C.c_int32),
        ("max_seq_len", C.c_int32),
        ("max_batch_tokens", C.c_int32),
        ("max_new_tokens", C.c_int32),
        ("kv_pool_tokens", C.c_int64),
        ("device", C.c_int32),
        ("use_cuda_graphs", C.c_int32),
        ("use_pdl", C.c_int32),
        ("decode_split_k", C.c_int32),
        ("keep_logits", C.c_int32),
        ("decode_reduce", C.c_int32),
        ("prefix_cache", C.c_int32),
    ]

    def __init__(self, *args, **kw):
        super().__init__(*args, **kw)
        if not self.struct_size:
            self.struct_size = C.sizeof(Config)


class Sampling(C.Structure):
    _fields_ = [
        ("max_new_tokens", C.c_int32),
        ("min_new_tokens", C.c_int32),
        ("temperature", C.c_float),
        ("top_k", C.c_int32),
        ("seed", C.c_uint64),
    ]


def declared_symbols(header: Path = HEADER_PATH) -> list[str]:
    """Every function name include/b200llm.h declares (used by the CPU-side export test)."""
    text = header.read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200llm_[a-z0-9_]+)\s*\(", text)))


def _load() -> C.CDLL:
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} is missing: the CUDA extension must be built (python __graft_entry__.py "
            "build). There is no CPU fallback for the generation path."
        )
    lib = C.CDLL(str(LIB_PATH))
    vp, i32, i64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float
    P = C.POINTER
    sigs = {
        "b200llm_last_error": (C.c_char_p, []),
        "b200llm_abi_version": (C.c_int, []),
        "b200llm_weight_blob_floats": (C.c_size_t, [P(Config)]),
        "b200llm_engine_create": (C.c_int, [P(Config), vp, C.c_size_t, P(vp)]),
        "b200llm_engine_destroy": (C.c_int, [vp]),
        "b200llm_engine_set_stream": (C.c_int, [vp, vp]),
        "b200llm_seq_alloc": (C.c_int, [vp, i32, P(i32)]),
        "b200llm_seq_free": (C.c_int, [vp, i32]),
        "b200llm_seq_alloc_with_prefix": (C.c_int, [vp, vp, i32, i32, P(i32), P(i32)]),
        "b200llm_prefix_cache_stats": (C.c_int, [vp, P(i64), P(i64), P(i64)]),
        "b200llm_seq_cancel": (C.c_int, [vp, i32]),
        "b200llm_read_tokens_batch": (C.c_int, [vp, vp, i32, i32, vp, i32, vp, vp]),
        "b200llm_op_lm_head_argmax": (C.c_int, [vp, vp, i32, i32, i32, i32, vp, vp, vp, vp, i32, vp]),
        "b200llm_lm_head_argmax_groups": (i32, [i32, i32]),
        "b200llm_op_prefill_attention_paged": (C.c_int, [vp, i32, vp, vp, vp, i32, i32, i32, vp, i32, vp, vp, i32,
                                                         vp, i32, vp]),
        "b200llm_op_ln_gemm": (C.c_int, [vp, vp, vp, f32, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp, vp]),
        "b200llm_op_gemm_redadd": (C.c_int, [vp, vp, vp, vp, i32, i32, i32, i32, vp]),
        "b200llm_op_kv_scatter": (C.c_int, [vp, vp, vp, i32, i32, vp, i32, vp, vp, vp]),
        "b200llm_kv_pages_free": (C.c_int, [vp, P(i64), P(i64)]),
        "b200llm_prefill": (C.c_int, [vp, vp, i32, vp, vp, vp, i32, vp]),
        "b200llm_decode": (C.c_int, [vp, vp, i32, i32]),
        "b200llm_sync": (C.c_int, [vp]),
        "b200llm_read_tokens": (C.c_int, [vp, i32, vp, i32, P(i32), P(i32)]),
        "b200llm_read_status": (C.c_int, [vp, vp, i32, vp, vp]),
        "b200llm_read_last_logits": (C.c_int, [vp, i32, vp]),
        "b200llm_generate": (C.c_int, [vp, vp, vp, vp, i32, vp, i32, vp]),
        "b200llm_kernel_launches": (i64, [vp]),
        "b200llm_op_gemm": (C.c_int, [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp]),
        "b200llm_op_layernorm": (C.c_int, [vp, vp, i32, vp, vp, vp, vp, i32, i32, f32, vp]),
        "b200llm_op_embed": (C.c_int, [vp, vp, vp, vp, vp, i32, i32, i32, i32, vp]),
        "b200llm_op_prefill_attention": (C.c_int, [vp, vp, i32, i32, i32, vp, vp]),
        "b200llm_op_decode_attention": (C.c_int, [vp, vp, vp, i32, vp, vp, vp, i32, i32, i32, vp, vp]),
        "b200llm_op_decode_attention_stream": (C.c_int, [vp, vp, vp, i32, vp, vp, i32, vp, i32, i32, i32, vp, vp, vp]),
        "b200llm_decode_attention_scratch_floats": (C.c_size_t, [i32, i32, i32]),
        "b200llm_op_decode_attention_noalloc": (C.c_int, [vp, vp, vp, i32, vp, vp, vp, i32, i32, i32, vp]),
        "b200llm_op_argmax": (C.c_int, [vp, i32, i32, i32, vp, vp]),
        "b200llm_op_convert_transpose": (C.c_int, [vp, vp, i32, i32, vp]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


def check(code: int) -> None:
    if code != 0:
        raise B200Error(code, lib.b200llm_last_error().decode("utf-8", "replace"))
