"""CPU-side: the C-ABI library loads and exports every symbol include/b200llm.h declares, and
the three descriptions of b200llm_config (header, _lib.Config, INTEGRATION.md's stub) agree."""
import ctypes
import re
from pathlib import Path

from conftest import pkg


def test_library_exports_every_declared_symbol():
    _lib = pkg("_lib")
    names = _lib.declared_symbols()
    assert len(names) >= 20
    raw = ctypes.CDLL(str(_lib.LIB_PATH))
    missing = [n for n in names if not hasattr(raw, n)]
    assert not missing, missing
    assert raw.b200llm_abi_version() == 2


def test_blob_size_matches_gpt2_small_parameter_count():
    _lib = pkg("_lib")
    cfg = _lib.Config(n_layer=12, n_head=12, d_model=768, vocab_size=50257, n_positions=1024)
    # 124,439,808 parameters with the tied head counted once (124.44 M)
    assert _lib.lib.b200llm_weight_blob_floats(ctypes.byref(cfg)) == 124439808


def test_invalid_arguments_report_errors_without_a_gpu():
    _lib = pkg("_lib")
    rc = _lib.lib.b200llm_engine_create(None, None, 0, None)
    assert rc == 1
    assert b"null" in _lib.lib.b200llm_last_error()


C_TYPES = {"int32_t": ctypes.c_int32, "uint32_t": ctypes.c_uint32, "int64_t": ctypes.c_int64,
           "uint64_t": ctypes.c_uint64, "float": ctypes.c_float}


def _header_struct(name: str):
    _lib = pkg("_lib")
    text = re.sub(r"/\*.*?\*/", "", _lib.HEADER_PATH.read_text(), flags=re.S)
    body = re.search(r"typedef struct " + name + r"\s*\{(.*?)\}\s*" + name + r"\s*;", text, flags=re.S).group(1)
    fields = re.findall(r"\b(u?int32_t|u?int64_t|float)\s+(\w+)\s*;", body)
    assert fields, f"no fields parsed for {name}"
    return [(n, C_TYPES[t]) for t, n in fields]


def test_ctypes_structs_match_the_header_field_for_field():
    """b200llm_config grew a field and a documented binding went stale without anything
    failing. Names, order and types of both ABI structs are now compared with the header."""
    _lib = pkg("_lib")
    assert _lib.Config._fields_ == _header_struct("b200llm_config")
    assert _lib.Sampling._fields_ == _header_struct("b200llm_sampling")
    assert _header_struct("b200llm_config")[0] == ("struct_size", ctypes.c_uint32)
    assert _lib.Config(n_layer=1).struct_size == ctypes.sizeof(_lib.Config)


def test_integration_md_stub_matches_the_header():
    """The ctypes stub printed in INTEGRATION.md section C is executed and compared as well."""
    _lib = pkg("_lib")
    md = (Path(_lib.HEADER_PATH).parent.parent / "INTEGRATION.md").read_text()
    block = next(b for b in re.findall(r"```python\n(.*?)```", md, flags=re.S) if "class Config(C.Structure)" in b)
    start = block.index("class Config(C.Structure)")
    end = block.index("\n\n", start)
    ns = {"C": ctypes}
    exec(block[start:end], ns)  # noqa: S102  the documented class definition only
    assert ns["Config"]._fields_ == _header_struct("b200llm_config")
    assert ctypes.sizeof(ns["Config"]) == ctypes.sizeof(_lib.Config)
    # the documented constructor call fills every field and sets struct_size
    m = re.search(r"cfg = Config\((.*?)\)\n", block, flags=re.S)
    cfg = eval("Config(" + m.group(1) + ")", {"Config": ns["Config"], "C": ctypes})  # noqa: S307
    assert cfg.struct_size == ctypes.sizeof(_lib.Config) and cfg.n_layer == 12 and cfg.vocab_size == 50257
    assert cfg.max_seq_len <= cfg.n_positions


def test_a_config_built_against_another_header_is_refused():
    _lib = pkg("_lib")
    cfg = _lib.Config(n_layer=12, n_head=12, d_model=768, vocab_size=50257, n_positions=1024)
    cfg.struct_size = ctypes.sizeof(_lib.Config) - 4            # a binding that lacks the last field
    assert _lib.lib.b200llm_weight_blob_floats(ctypes.byref(cfg)) == 0
    eng = ctypes.c_void_p()
    dummy = (ctypes.c_float * 1)()
    rc = _lib.lib.b200llm_engine_create(ctypes.byref(cfg), dummy, 1, ctypes.byref(eng))
    assert rc == 1 and b"struct_size" in _lib.lib.b200llm_last_error()
